"""Row-sharded completion fit timing: python -m torch.distributed.run --nproc-per-node N
profiles/bench_complete_sharded.py --rows 20000 --cols 256  (N = 1 runs unsharded)."""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry  # noqa: E402
from profiles.bench_complete import synthetic  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", dest="n", type=int, default=20000)
    ap.add_argument("--cols", dest="d", type=int, default=256)
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    import torch
    import torch.distributed as dist

    pkg = entry.load_package()
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    X = synthetic(args.n, args.d, max(2, args.d // 8), 0.30, 5)
    lo, hi = rank * args.n // world, (rank + 1) * args.n // world
    sch = pkg.Scheduler(local)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    m = pkg.complete_fit_sharded(sch, X[lo:hi], pkg.torch_allreduce()) if world > 1 else pkg.complete_fit(sch, X)
    dt = time.perf_counter() - t0
    t = torch.tensor([dt], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        print(json.dumps({"n": args.n, "d": args.d, "n_gpus": world, "fit_s": float(t.item()), "rounds": m.n_rounds,
                          "steps": m.n_steps}), flush=True)
    m.close()
    sch.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
