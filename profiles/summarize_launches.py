"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel name."""
import csv, collections, sys
def main(path, note=""):
    with open(path) as f:
        lines = [l for l in f if not l.startswith('==')]
    agg = collections.defaultdict(lambda: [0, 0.0, 0.0])
    for row in csv.DictReader(lines):
        name = row['Kernel Name'].split('(')[0][:60]
        try: v = float(row['Metric Value'].replace(',', ''))
        except Exception: continue
        unit = row['Metric Unit']
        v = v / 1e3 if unit == 'ns' else (v * 1e3 if unit == 'ms' else v)
        a = agg[name]; a[0] += 1; a[1] += v; a[2] = max(a[2], v)
    tot = sum(a[1] for a in agg.values())
    if note: print("# " + note)
    for k, a in sorted(agg.items(), key=lambda x: -x[1][1]):
        print(f"{k:60s} n={a[0]:5d} total={a[1]/1e3:9.3f} ms  mean={a[1]/a[0]:9.2f} us max={a[2]:9.2f} us share={a[1]/tot:.3f}")
if __name__ == "__main__":
    main(sys.argv[1], " ".join(sys.argv[2:]))
