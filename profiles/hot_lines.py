"""Aggregate warp-stall samples per CUDA source line from
`ncu -i X.ncu-rep --page source --print-source sass,cuda --csv --kernel-name regex:K`."""
import csv, sys, collections
def main(path, top=25):
    rows = list(csv.reader(open(path)))
    fileName = ""; cur = None
    agg = collections.Counter(); text = {}
    for r in rows:
        if len(r) >= 2 and r[0] == 'File Path': fileName = r[1].split('/')[-1]; continue
        if len(r) >= 2 and r[0] in ('Function Name', 'Line No'): continue
        if len(r) > 4 and r[0] != '':
            cur = (fileName, r[0]); text[cur] = ','.join(r[1:]).strip()[:100]; continue
        if len(r) > 6 and r[0] == '' and cur is not None:
            try: agg[cur] += int(r[6])
            except ValueError: pass
    tot = sum(agg.values())
    print("total samples", tot)
    for k, v in agg.most_common(top):
        print(f"{v:6d} {100.0*v/max(tot,1):5.1f}%  {k[0]}:{k[1]:>4s}  {text[k]}")
if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 25)
