"""torchrun worker of tests/test_gpu_sharded.py::test_nccl_optimize_two_processes: a whole LM run on a
sharded problem over real NCCL; rank 0 prints one JSON line with the summary and a state checksum."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402


def main():
    import torch
    import torch.distributed as dist
    from visual_inertial_bundle_adjustment_b200 import synth, sharding, graph as G
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    os.dup2(2, 1) if rank != 0 else None
    torch.cuda.set_device(local)
    dist.init_process_group("nccl")
    s = synth.make_session(**synth.CONFIGS[sys.argv[1]])
    native = len(sys.argv) < 5 or sys.argv[4] != "python"   # the library's C++ NCCL hook, or the Python / torch.distributed one
    ctx, p, coll = sharding.make_sharded_problem(s.graph, rank, world, local, native=native)
    summ = p.optimize(G.default_settings(int(sys.argv[2])))
    v = p.vars
    t = torch.from_numpy(v.copy()).cuda()
    ref = t.clone()
    dist.broadcast(ref, src=0)
    same = torch.tensor([float(torch.equal(t, ref))], device="cuda")
    dist.all_reduce(same, op=dist.ReduceOp.MIN)
    if rank == 0:
        np.save(sys.argv[3], v)
        sys.stderr.flush()
        print(json.dumps({"iterations": summ.num_iterations, "initial_cost": summ.initial_cost, "final_cost": summ.final_cost,
                          "identical_state_on_all_ranks": bool(same.item()), "collectives": coll.calls,
                          "own_lumps": p.shard_info()["own_lumps"]}), flush=True)
    p.close(); ctx.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
