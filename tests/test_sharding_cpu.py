"""CPU tests of the N>1 path's host logic (no GPU): the shard plan exported by the C ABI
(vb_shard_plan, host-only) and the exchange algebra of DESIGN.md §7, run as a real world_size-2
torch.distributed job on the gloo backend with the CPU oracle standing in for the kernels:
each rank assembles only its own factors, eliminates only its own landmarks, the reduced camera
system / camera right-hand side / landmark steps are all-reduced, and the result must equal the
undivided solve."""
import os
import socket

import numpy as np
import pytest

from visual_inertial_bundle_adjustment_b200 import synth, capi, sharding, graph as G

LAM = 1e-5


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("name,nranks", [("small", 2), ("small_rs", 3), ("tiny_rs", 4), ("small", 8), ("small_lc_rs", 2)])
def test_shard_plan_partitions_everything_once(name, nranks):
    g = synth.make_session(**synth.CONFIGS[name]).graph
    owner = capi.shard_plan(g, nranks)
    assert owner.shape == (g.num_points,) and owner.min() >= 0 and owner.max() < nranks
    assert len(np.unique(owner)) == nranks  # nobody idle
    seen = [np.zeros(st.count, np.int32) for st in g.stores]
    for r in range(nranks):
        for s, keep in zip(seen, sharding.owned_factors(g, owner, r, nranks)):
            s[keep] += 1
    assert all((s == 1).all() for s in seen)  # every factor evaluated by exactly one rank
    # all observations of a landmark live on its owner, and the observation load is balanced
    load = np.zeros(nranks)
    for st in g.stores:
        if st.type in (G.F_VISUAL, G.F_VISUAL_RS):
            load += np.bincount(owner[g.var_index[st.vars[:, 0]]], minlength=nranks)
    assert load.max() <= 1.6 * load.mean()
    assert np.array_equal(capi.shard_plan(g, nranks), owner)  # deterministic
    assert (capi.shard_plan(g, 1) == 0).all()


def _worker(rank, world, port, name, out_dir):
    import torch
    import torch.distributed as dist
    from tests.oracle_lib import load_oracle, Oracle
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = synth.make_session(**synth.CONFIGS[name]).graph
        owner = capi.shard_plan(g, world)
        lib = load_oracle()
        o = Oracle(lib, sharding.subgraph(g, owner, rank, world))
        cost = o.eval_grad_hess()
        H, grad = o.dense_hessian(), o.grad.copy()
        n3 = 3 * g.num_points
        mine = np.repeat(owner == rank, 3)
        # addDamping on a shard: d(1+lam) everywhere, "+lam" on owned landmarks and, for cameras, on rank 0 only
        d = np.diag(H).copy() * (1.0 + LAM)
        d[:n3] += LAM * mine
        if rank == 0:
            d[n3:] += LAM
        H[np.diag_indices_from(H)] = d
        # local landmark elimination (owned landmarks only)
        Hpp = H[:n3, :n3][np.ix_(mine, mine)]
        Hcp = H[n3:, :n3][:, mine]
        S = H[n3:, n3:] - Hcp @ np.linalg.solve(Hpp, Hcp.T)
        gc = grad[n3:] - Hcp @ np.linalg.solve(Hpp, grad[:n3][mine])
        # exchange: reduced camera system, camera right-hand side, scalars
        tS, tg, tc = torch.from_numpy(S), torch.from_numpy(gc), torch.tensor([cost], dtype=torch.float64)
        dist.all_reduce(tS); dist.all_reduce(tg); dist.all_reduce(tc)
        sc = np.linalg.solve(S, gc)
        tsc = torch.from_numpy(sc)
        dist.broadcast(tsc, src=0)
        sp = np.zeros(n3)
        sp[mine] = np.linalg.solve(Hpp, grad[:n3][mine] - Hcp.T @ sc)
        tsp = torch.from_numpy(sp)
        dist.all_reduce(tsp)
        step = np.concatenate([sp, sc])
        dots = torch.tensor([float(step @ grad)], dtype=torch.float64)  # full step . partial gradient
        dist.all_reduce(dots)
        if rank == 0:
            full = Oracle(lib, g)
            c_full = full.eval_grad_hess()
            g_full = full.grad.copy()
            full.damp(LAM)
            ref = np.linalg.solve(full.dense_hessian(), g_full)
            np.savez(os.path.join(out_dir, "r0.npz"), step=step, ref=ref, cost=tc.numpy(), c_full=c_full,
                     dot=dots.numpy(), dot_ref=float(ref @ g_full))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("name,world", [("tiny_rs", 2), ("tiny", 2), ("tiny_lc", 2)])
def test_exchange_algebra_gloo(tmp_path, oracle, name, world):
    import torch.multiprocessing as mp
    mp.spawn(_worker, args=(world, _free_port(), name, str(tmp_path)), nprocs=world, join=True)
    z = np.load(tmp_path / "r0.npz")
    assert abs(z["cost"][0] - z["c_full"]) <= 1e-12 * abs(z["c_full"])
    assert np.abs(z["step"] - z["ref"]).max() <= 1e-7 * np.abs(z["ref"]).max()
    assert abs(z["dot"][0] - z["dot_ref"]) <= 1e-8 * abs(z["dot_ref"])


def test_sharded_entry_points_fail_loudly_without_gpu():
    lib = capi.load()
    import ctypes as C
    h = C.c_void_p()
    assert lib.vb_ctx_create_on_stream(0, None, C.byref(h)) == 2  # VB_ERR_CUDA: no CPU fallback
    g = synth.make_session(**synth.CONFIGS["tiny"]).graph
    desc = g.desc()
    out = np.zeros(g.num_points + 1, np.int32)
    assert lib.vb_shard_plan(C.byref(desc), 2, out.ctypes.data_as(C.POINTER(C.c_int32)), out.size) == 1  # VB_ERR_ARG
