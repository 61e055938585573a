"""The C++17 host mirror of small_thing::Optimizer (include/small_thing/*.h) against the Python
builder: a synthetic session is handed to tests/cpp/host_mirror_main.cpp, which rebuilds it through
Variable<T> / addFactor / registerVariable / registeredVariablesToEliminationRange the way a user of
the reference's lib/small_thing would, and

* (no GPU) the graph the mirror hands to the C ABI must equal the Python builder's graph array by
  array — registration indices, per-store variable tables, constants, losses, preintegration
  records, rolling-shutter tables — and give the same symbolic phase (order, dataSize, supernodes);
  a compute call must throw std::runtime_error without a device (no host numerics);
* (GPU) Optimizer::optimize / computeCost through the mirror must reproduce the Python path's run
  and leave the optimised values in the user's own Variables.

Like the reference (one FactorStore per functor type, Optimizer.h:115-175) the mirror keeps one store
per (factor type, arity, loss); the Python session is merged the same way before comparing."""
import os
import subprocess

import numpy as np
import pytest

from visual_inertial_bundle_adjustment_b200 import synth, capi, graph as G

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "visual_inertial_bundle_adjustment_b200")
SRC = os.path.join(ROOT, "tests", "cpp", "host_mirror_main.cpp")
RS_STRIDE = synth.RS_STRIDE


@pytest.fixture(scope="module")
def mirror_bin():
    capi.load()  # builds / locates libviba_b200.so
    out_dir = os.path.join(ROOT, "tests", "cpp", "_build")
    os.makedirs(out_dir, exist_ok=True)
    exe = os.path.join(out_dir, "host_mirror_main")
    hdrs = [os.path.join(ROOT, "include", "small_thing", h) for h in ("Optimizer.h", "Factor.h", "Variable.h", "SoftLoss.h")]
    deps = [SRC, os.path.join(ROOT, "include", "viba_b200.h")] + hdrs
    if not os.path.exists(exe) or os.path.getmtime(exe) < max(os.path.getmtime(d) for d in deps):
        cmd = ["g++", "-std=c++17", "-O1", "-Wall", "-Wextra", "-Werror", "-I" + os.path.join(ROOT, "include"), SRC, "-o", exe,
               "-L" + PKG, "-lviba_b200", "-Wl,-rpath," + PKG]
        r = subprocess.run(cmd, capture_output=True, text=True)
        assert r.returncode == 0, r.stderr[-4000:]
    return exe


def merged(g, point_ids):
    """same variables, one store per (type, arity, loss) in first-use order, re-registered"""
    m = G.Graph()
    m._kind, m._tdim, m._flags, m._aux, m._data, m._const = g._kind, g._tdim, g._flags, g._aux, g._data, g._const
    m.nv = g.nv
    groups = {}
    for st in g.stores:
        groups.setdefault((st.type, st.vars.shape[1], st.loss), []).append(st)
    for (t, _, loss), sts in groups.items():
        aux = None
        if sts[0].aux is not None:
            assert len(sts) == 1 or t in (G.F_INERTIAL, G.F_INERTIAL_SEC, G.F_INERTIAL_SEC_SPLIT)
            aux = np.concatenate([s.aux for s in sts])
        m.add_store(G.FactorStore(t, np.concatenate([s.vars for s in sts]), np.concatenate([s.consts for s in sts]), loss=loss, aux=aux))
    return m.finalize(point_ids)


def dump(g, point_ids, path):
    const = np.concatenate(g._const).astype(np.uint8)
    with open(path, "wb") as f:
        np.int64(g.nv).tofile(f)
        g.var_kind.astype(np.int32).tofile(f); g.var_tdim.astype(np.int32).tofile(f); g.var_flags.astype(np.uint32).tofile(f)
        g.var_aux.astype(np.float64).tofile(f); const.tofile(f)
        np.int64(g.var_data.size).tofile(f); g.var_data.tofile(f)
        np.int64(len(point_ids)).tofile(f); np.asarray(point_ids, np.int32).tofile(f)
        np.int32(len(g.stores)).tofile(f)
        for st in g.stores:
            np.array([st.type, st.vars.shape[1]], np.int32).tofile(f)
            np.int64(st.count).tofile(f)
            np.int32(st.consts.shape[1]).tofile(f)
            np.array(st.loss, np.float64).tofile(f)
            np.int64(0 if st.aux is None else st.aux.size).tofile(f)
            st.vars.tofile(f); st.consts.tofile(f)
            if st.aux is not None:
                st.aux.tofile(f)


class _R:
    def __init__(self, path):
        self.b = open(path, "rb").read()
        self.o = 0

    def take(self, dt, n=None):
        k = 1 if n is None else int(n)
        a = np.frombuffer(self.b, dtype=dt, count=k, offset=self.o)
        self.o += a.nbytes
        return a[0] if n is None else a


def load_cpp_graph(path):
    r = _R(path)
    d = {}
    n = d["nv"] = int(r.take(np.int64))
    d["kind"], d["tdim"], d["flags"] = r.take(np.int32, n), r.take(np.int32, n), r.take(np.uint32, n)
    d["index"], d["off"], d["aux"] = r.take(np.int64, n), r.take(np.int64, n), r.take(np.float64, n)
    d["data"] = r.take(np.float64, int(r.take(np.int64)))
    d["num_points"] = int(r.take(np.int64))
    d["py2cpp"] = r.take(np.int32, int(r.take(np.int64)))
    d["stores"] = []
    for _ in range(int(r.take(np.int32))):
        s = {}
        s["type"], s["nvars"] = (int(x) for x in r.take(np.int32, 2))
        s["count"] = int(r.take(np.int64)); s["cpf"] = int(r.take(np.int32))
        s["loss"] = tuple(float(x) for x in r.take(np.float64, 2))
        al = int(r.take(np.int64))
        s["vars"] = r.take(np.int32, s["count"] * s["nvars"]).reshape(s["count"], s["nvars"])
        s["consts"] = r.take(np.float64, s["count"] * s["cpf"]).reshape(s["count"], s["cpf"])
        s["aux"] = r.take(np.float64, al)
        d["stores"].append(s)
    d["order"], d["data_size"], d["num_lumps"], d["var_dim"], d["err_dim"] = (int(x) for x in r.take(np.int64, 5))
    assert r.o == len(r.b)
    return d


@pytest.mark.parametrize("name", ["tiny", "tiny_rs", "small_rs", "tiny_lc"])
def test_mirror_builds_the_same_graph(name, mirror_bin, tmp_path):
    s = synth.make_session(**synth.CONFIGS[name])
    g = merged(s.graph, s.ids["point"])
    dump(g, s.ids["point"], tmp_path / "in.bin")
    r = subprocess.run([mirror_bin, str(tmp_path / "in.bin"), str(tmp_path / "out.bin")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-2000:]
    c = load_cpp_graph(tmp_path / "out.bin")
    m = c["py2cpp"]
    used = m >= 0
    referenced = np.zeros(g.nv, bool)
    referenced[np.asarray(s.ids["point"])] = True
    for st in g.stores:
        referenced[st.vars.ravel()] = True
    assert np.array_equal(used, referenced) and c["nv"] == int(used.sum())
    assert len(np.unique(m[used])) == c["nv"]
    # variables: same kind / tangent dim / flags / registration index / values
    for key, ref in (("kind", g.var_kind), ("tdim", g.var_tdim), ("flags", g.var_flags), ("index", g.var_index), ("aux", g.var_aux)):
        assert np.array_equal(c[key][m[used]], ref[used]), key
    dd = np.array([G.DATA_DIM[k] for k in range(6)])[g.var_kind]
    for i in np.flatnonzero(used)[:: max(1, g.nv // 500)]:
        o = c["off"][m[i]]
        assert np.array_equal(c["data"][o:o + dd[i]], g.var_data[g.var_off[i]:g.var_off[i] + dd[i]])
    assert c["data"].size == int(dd[used].sum())
    assert c["num_points"] == g.num_points
    # stores: same order, tables, constants, losses, store-level data
    assert len(c["stores"]) == len(g.stores)
    for cs, st in zip(c["stores"], g.stores):
        assert (cs["type"], cs["nvars"], cs["count"], cs["cpf"]) == (st.type, st.vars.shape[1], st.count, st.consts.shape[1])
        assert cs["loss"] == st.loss
        assert np.array_equal(cs["vars"], m[st.vars])
        assert np.array_equal(cs["consts"], st.consts)
        if st.type == G.F_VISUAL_RS:
            rigs = np.unique(st.consts[:, 6].astype(np.int64))
            assert np.array_equal(cs["aux"][:3], st.aux[:3]) and int(cs["aux"][3]) == rigs.max() + 1
            a = cs["aux"][4:].reshape(-1, RS_STRIDE)
            b = st.aux[4:].reshape(-1, RS_STRIDE)
            assert np.array_equal(a[rigs], b[rigs])
        elif st.aux is not None:
            assert np.array_equal(cs["aux"], st.aux)
        else:
            assert cs["aux"].size == 0
    # and the symbolic phase sees the same problem
    order, data_size, lumps, _ = capi.symbolic_only(g)
    assert (c["order"], c["data_size"], c["num_lumps"]) == (order, data_size, lumps)
    assert c["var_dim"] == order
    assert c["err_dim"] == sum(st.count * G.ERR_SIZE[st.type] for st in g.stores)


def test_mirror_host_contract(mirror_bin):
    """addFactor argument checks, store introspection, registration order, constants, loss-by-pointer semantics,
    elimination-range checks: std::runtime_error wherever the reference's BASPACHO_CHECK would throw"""
    r = subprocess.run([mirror_bin, "--selftest"], capture_output=True, text=True)
    assert r.returncode == 0 and "selftest: ok" in r.stdout, r.stdout + r.stderr


def test_mirror_compute_call_throws_without_device(mirror_bin, tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a device is present")
    s = synth.make_session(**synth.CONFIGS["tiny"])
    dump(merged(s.graph, s.ids["point"]), s.ids["point"], tmp_path / "in.bin")
    r = subprocess.run([mirror_bin, str(tmp_path / "in.bin"), str(tmp_path / "out.bin"), "--optimize", "2"], capture_output=True, text=True)
    assert r.returncode == 3 and "runtime_error: viba_b200:" in r.stderr  # no host numerics behind the mirror


@pytest.mark.gpu
@pytest.mark.parametrize("name,iters", [("small_rs", 6)])
def test_mirror_optimize_matches_python_path(name, iters, mirror_bin, tmp_path):
    s = synth.make_session(**synth.CONFIGS[name])
    g = merged(s.graph, s.ids["point"])
    dump(g, s.ids["point"], tmp_path / "in.bin")
    r = subprocess.run([mirror_bin, str(tmp_path / "in.bin"), str(tmp_path / "out.bin"), "--optimize", str(iters)],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    out = np.fromfile(tmp_path / "out.bin", np.float64)
    init_cost, final_cost, n_it, cost0, cost1 = out[:5]
    vals = out[5:]
    ctx = capi.Context(0)
    p = capi.Problem(ctx, g)
    summ = p.optimize(G.default_settings(iters))
    ref = p.vars
    p.close(); ctx.close()
    assert int(n_it) == summ.num_iterations
    assert abs(init_cost - summ.initial_cost) <= 1e-12 * abs(summ.initial_cost)
    assert abs(final_cost - summ.final_cost) <= 1e-9 * abs(summ.final_cost)
    assert abs(cost0 - init_cost) <= 1e-10 * abs(init_cost)      # computeCost before == initial cost
    assert abs(cost1 - final_cost) <= 1e-9 * abs(final_cost)     # and after: values went back and forth intact
    assert vals.size == ref.size and np.abs(vals - ref).max() <= 1e-6
    assert final_cost < 0.9 * init_cost


@pytest.mark.gpu
@pytest.mark.parametrize("flag,solver_type", [("--pcg-jacobi", G.SOLVER_PCG_JACOBI), ("--pcg-gauss-seidel", G.SOLVER_PCG_GAUSS_SEIDEL)])
def test_mirror_optimize_with_pcg(flag, solver_type, mirror_bin, tmp_path):
    """Settings::solverType = Solver_PCG_Jacobi / Solver_PCG_GaussSeidel through the C++ mirror (Optimizer.cpp:232-331): the
    same run as the Python path with the same settings, and the direct solver's optimum"""
    iters = 6
    s = synth.make_session(**synth.CONFIGS["small_rs"])
    g = merged(s.graph, s.ids["point"])
    dump(g, s.ids["point"], tmp_path / "in.bin")
    r = subprocess.run([mirror_bin, str(tmp_path / "in.bin"), str(tmp_path / "out.bin"), "--optimize", str(iters), flag],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    out = np.fromfile(tmp_path / "out.bin", np.float64)
    init_cost, final_cost, n_it = out[:3]
    vals = out[5:]
    ctx = capi.Context(0)
    st = G.default_settings(iters)
    p = capi.Problem(ctx, g); direct = p.optimize(st); p.close()
    st.solver_type, st.pcg_max_iterations, st.pcg_desired_residual = solver_type, 8000, 1e-12
    p = capi.Problem(ctx, g); summ = p.optimize(st); ref = p.vars; last = p.pcg_last(); p.close()
    ctx.close()
    assert last[0] > 1 and last[1] < 1e-12
    assert int(n_it) == summ.num_iterations == direct.num_iterations
    assert abs(final_cost - summ.final_cost) <= 1e-8 * abs(summ.final_cost)
    assert abs(final_cost - direct.final_cost) <= 1e-6 * abs(direct.final_cost)
    assert vals.size == ref.size and np.abs(vals - ref).max() <= 1e-6


@pytest.mark.gpu
def test_mirror_covariances_and_marginal_problem(mirror_bin, tmp_path):
    """the recipe of the reference's TestOptimizer.cpp:52-83 through the C++ mirror: computeCovariances,
    computeJointCovariances and computeMarginalProblem agree on offsets, marginal information x covariance = I, the
    per-variable covariances are the diagonal blocks of the joint one, duplicated indices throw"""
    s = synth.make_session(**synth.CONFIGS["tiny_rs"])
    g = merged(s.graph, s.ids["point"])
    dump(g, s.ids["point"], tmp_path / "in.bin")
    r = subprocess.run([mirror_bin, str(tmp_path / "in.bin"), str(tmp_path / "out.bin"), "--covariance"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    D, delta, diag_diff, diag_scale, m_cost, cost, threw = np.fromfile(tmp_path / "out.bin", np.float64)[:7]
    assert D >= 3 and threw == 1.0
    assert delta <= 1e-6                      # 1e-9 in the reference, on a 6-variable spring chain
    assert diag_diff <= 1e-9 * diag_scale
    assert 0 <= m_cost <= cost * (1 + 1e-12)  # the eliminated variables can only remove cost


@pytest.mark.gpu
def test_mirror_pre_step_callback_matches_oracle(oracle, mirror_bin, tmp_path):
    """Settings::preStepCallback through the C++ mirror: user code rewrites a preintegration record at iteration 2 from
    the calibration it reads in its own Variable at that moment; the oracle driven the same way gives the same run"""
    from tests.oracle_lib import Oracle
    iters = 6
    s = synth.make_session(**synth.CONFIGS["small_rs"])
    g = merged(s.graph, s.ids["point"])
    dump(g, s.ids["point"], tmp_path / "in.bin")
    r = subprocess.run([mirror_bin, str(tmp_path / "in.bin"), str(tmp_path / "out.bin"), "--optimize", str(iters), "--callback"],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    out = np.fromfile(tmp_path / "out.bin", np.float64)
    init_cost, final_cost, n_it, cost0, cost1, log_lines, callbacks, calib_seen = out[:8]
    vals = out[8:]
    # the oracle with the same callback
    o = Oracle(oracle, g)
    store = next(i for i, st in enumerate(g.stores) if st.type == G.F_INERTIAL)
    calib_var = next(v for v in range(g.nv) if g.var_kind[v] == G.VAR_IMUCALIB and g.var_index[v] >= 0)
    seen = []

    def pre(it):
        if it != 2:
            return
        aux = o.store_aux(store)
        aux[4] += 1e-3
        cur = o.vars[g.var_off[calib_var]:g.var_off[calib_var] + 32]
        aux[218:250] = cur
        seen.append(cur[6])
        o.set_store_aux(store, aux)
    so, _ = o.optimize_ex(G.default_settings(iters), pre_step=pre)
    assert int(n_it) == so.num_iterations and int(log_lines) == so.num_iterations and int(callbacks) >= so.num_iterations
    assert abs(final_cost - so.final_cost) <= 1e-9 * abs(so.final_cost)
    assert abs(calib_seen - seen[0]) <= 1e-9 * max(1e-3, abs(seen[0])) and calib_seen != g.var_data[g.var_off[calib_var] + 6]
    assert vals.size == o.vars.size and np.abs(vals - o.vars).max() <= 1e-6


@pytest.mark.gpu
def test_baspacho_facade_over_the_mirror(mirror_bin, tmp_path):
    """include/baspacho/baspacho/{Solver,Accessor}.h: order / dataSize / accessor / factor / solve over the mirror's device
    problem give the plain ABI's result (to rounding: the Schur update sums with fp64 reductions in no fixed order); the
    partial-factor family throws"""
    s = synth.make_session(**synth.CONFIGS["small_rs"])
    g = merged(s.graph, s.ids["point"])
    dump(g, s.ids["point"], tmp_path / "in.bin")
    r = subprocess.run([mirror_bin, str(tmp_path / "in.bin"), str(tmp_path / "out.bin"), "--facade"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    n, ds, checked, worst, scale, threw, cost = np.fromfile(tmp_path / "out.bin", np.float64)
    ctx = capi.Context(0)
    p = capi.Problem(ctx, g)
    assert (int(n), int(ds)) == (p.order, p.data_size) and int(checked) == p.order
    assert abs(cost - p.compute_grad_hess()) <= 1e-12 * abs(cost)
    p.close(); ctx.close()
    assert worst <= 1e-9 * scale and scale > 0 and threw == 1.0
