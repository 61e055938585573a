"""Host side of the multi-GPU path: one process per GPU, the same
graph on every rank, landmarks (with all of their observations) split over the ranks, and the
reduced camera system summed over ranks by the collective hook of the C ABI
(``vb_problem_create_sharded``).  torch.distributed is plumbing here: it only moves the buffers the
library hands to the hook.

* ``TorchCollective``  — NCCL (or gloo) all-reduce / broadcast on the raw device pointers.
* ``ThreadCollective`` — N ranks as N threads of ONE process on ONE GPU (single-GPU test of the
  N>1 path: same kernels, same exchange points, the sum done with torch ops).
* ``subgraph``         — the shard rule restated on the host graph (which factors a rank evaluates);
  the CPU/gloo tests use it to check the exchange algebra against the oracle.
"""
import copy
import threading

import numpy as np

from . import capi
from . import graph as G


class _RawDeviceArray:
    """__cuda_array_interface__ view of `count` doubles at a raw device pointer."""

    def __init__(self, ptr, count):
        self.__cuda_array_interface__ = {"shape": (int(count),), "typestr": "<f8", "data": (int(ptr), False),
                                         "version": 3, "strides": None}


def alias_tensor(ptr, count, device):
    import torch
    return torch.as_tensor(_RawDeviceArray(ptr, count), device=f"cuda:{device}")


class TorchCollective:
    """In-place sum / broadcast-from-0 over a torch.distributed group, ordered on `stream`
    (a torch.cuda.Stream whose raw handle was given to capi.Context)."""

    def __init__(self, device, stream, group=None):
        import torch.distributed as dist
        self.device, self.stream, self.group, self.dist = device, stream, group, dist
        self.calls, self.bytes = 0, 0
        self._views = {}

    def __call__(self, op, ptr, count, stream_handle):
        import torch
        key = (ptr, count)
        t = self._views.get(key)
        if t is None:
            t = self._views[key] = alias_tensor(ptr, count, self.device)
        with torch.cuda.stream(self.stream):
            if op == capi.COLL_SUM:
                self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM, group=self.group)
            elif op == capi.COLL_BCAST0:
                self.dist.broadcast(t, src=0 if self.group is None else self.dist.get_global_rank(self.group, 0), group=self.group)
            else:
                raise ValueError(f"unknown collective op {op}")
        self.calls += 1
        self.bytes += 8 * count


class ThreadCollective:
    """N ranks as N host threads sharing one GPU.  Each rank calls `for_rank(r)`; the hook
    synchronises the rank's stream, meets the others at a barrier and sums on the device."""

    def __init__(self, nranks, device=0):
        self.n, self.device = nranks, device
        self.barrier = threading.Barrier(nranks)
        self.slots = [None] * nranks
        self.total = None

    def for_rank(self, rank, stream):
        def call(op, ptr, count, stream_handle):
            import torch
            t = alias_tensor(ptr, count, self.device)
            stream.synchronize()
            self.slots[rank] = t
            self.barrier.wait()
            if rank == 0:
                if op == capi.COLL_SUM:
                    tot = self.slots[0].clone()
                    for o in self.slots[1:]:
                        tot += o
                else:
                    tot = self.slots[0].clone()
                self.total = tot
                torch.cuda.synchronize(self.device)
            self.barrier.wait()
            t.copy_(self.total)
            torch.cuda.synchronize(self.device)
            self.barrier.wait()
        return call


def make_sharded_problem(graph, rank, nranks, device, group=None, native=True):
    """(ctx, problem, collective) for one rank of a torch.distributed job.  native: the library's own C++ NCCL hook
    (csrc/nccl_hook.cpp) carries the exchange steps — torch.distributed only hands the 128-byte NCCL id from rank 0 to the
    others; native=False keeps the Python hook (TorchCollective)."""
    import torch
    torch.cuda.set_device(device)
    stream = torch.cuda.Stream(device=device)
    ctx = capi.Context(device, stream=stream.cuda_stream)
    if nranks > 1 and native:
        import torch.distributed as dist
        box = [capi.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(box, src=0 if group is None else dist.get_global_rank(group, 0), group=group)
        coll = capi.NcclComm(rank, nranks, box[0])
        p = capi.Problem(ctx, graph, rank, nranks, nccl_comm=coll)
    else:
        coll = TorchCollective(device, stream, group) if nranks > 1 else None
        p = capi.Problem(ctx, graph, rank, nranks, coll)
    p._stream = stream  # keep the stream alive as long as the problem
    return ctx, p, coll


def run_threaded(graph, nranks, fn, device=0):
    """Run `fn(problem, rank)` on `nranks` shards of `graph` as threads on one GPU; returns the
    list of results (exceptions are re-raised)."""
    import torch
    tc = ThreadCollective(nranks, device)
    out, errs = [None] * nranks, [None] * nranks

    def worker(r):
        try:
            torch.cuda.set_device(device)
            stream = torch.cuda.Stream(device=device)
            ctx = capi.Context(device, stream=stream.cuda_stream)
            p = capi.Problem(ctx, graph, r, nranks, tc.for_rank(r, stream))
            try:
                out[r] = fn(p, r)
            finally:
                stream.synchronize()
                p.close()
                ctx.close()
        except BaseException as ex:
            errs[r] = ex
            tc.barrier.abort()
            import traceback
            errs[r].__notes__ = [f"rank {r}: " + traceback.format_exc()]

    ts = [threading.Thread(target=worker, args=(r,)) for r in range(nranks)]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    # the root cause first: a rank that failed on its own, not one whose hook broke because a peer left
    secondary = lambda e: isinstance(e, threading.BrokenBarrierError) or "collective hook failed" in str(e)
    for e in sorted((e for e in errs if e is not None), key=secondary):
        raise e
    return out


# ---- the shard rule on the host graph -----------------------------------------------------------
def owned_factors(graph, owner, rank, nranks):
    """Per store: indices of the factors rank `rank` evaluates (the library's own plan, host-only):
    visual factors follow the owner of their landmark, the others the rank whose subtree they touch."""
    out = []
    for i, st in enumerate(graph.stores):
        fo, _ = capi.shard_factor_owners(graph, nranks, i)
        if st.type in (G.F_VISUAL, G.F_VISUAL_RS):
            idx = graph.var_index[st.vars[:, 0]]
            assert np.array_equal(fo[idx >= 0], owner[idx[idx >= 0]])  # all observations of a landmark on its owner
        out.append(np.nonzero(fo == rank)[0])
    return out


def subgraph(graph, owner, rank, nranks):
    """The graph restricted to the factors of one rank: same variables and registration indices."""
    g = copy.copy(graph)
    g.stores = []
    for st, keep in zip(graph.stores, owned_factors(graph, owner, rank, nranks)):
        if len(keep) == 0:
            continue
        aux = st.aux
        if st.type in (G.F_INERTIAL, G.F_INERTIAL_SEC, G.F_INERTIAL_SEC_SPLIT) and aux is not None:
            aux = aux.reshape(st.count, G.PREINT_STRIDE)[keep]  # one preintegration record per factor
        g.stores.append(G.FactorStore(st.type, st.vars[keep], st.consts[keep], st.loss, aux))
    return g
