"""Node-sharded scoring across the GPUs of one box: one process per GPU, torch.distributed for
the plumbing (NCCL over NVLink on GPUs, gloo in the CPU tests of the host logic).

Scores of different nodes are independent (Logic reads only its own node's UUIDs and
residents, gpu_plugins.go:536-544), so nodes are block-partitioned over the ranks and pods /
prediction tables are replicated.  The one exchange step of the path: NormalizeScore needs the
per-pod minimum and maximum over ALL nodes (gpu_plugins.go:818-827) and host selection needs
the argmax.  Both phases of the C ABI are built for it:

    gs_score_raw_dev   leaves, per pod, hi_arg = max(raw << 32 | 0xFFFFFFFF - global node)
                       and neg_lo = max(100 - raw) over the LOCAL nodes
    all_reduce(MAX)    on the concatenated int64[2, P] buffer makes both global
    gs_normalize_dev   rescales the local columns with the global extrema, decodes the best node
"""
from __future__ import annotations

import numpy as np


def shard_bounds(n_nodes: int, world: int, rank: int) -> tuple[int, int]:
    """Balanced block partition of the node axis: the first n_nodes % world ranks get one more."""
    base, extra = divmod(n_nodes, world)
    begin = rank * base + min(rank, extra)
    return begin, begin + base + (1 if rank < extra else 0)


def decode_extrema(ext: np.ndarray):
    """ext int64[2, P] (after the all-reduce) -> (hi, lo, best_node) as NormalizeScore / selectHost see them."""
    hi = (ext[0] >> 32).astype(np.int64)
    best = (0xFFFFFFFF - (ext[0] & 0xFFFFFFFF)).astype(np.int64)
    lo = 100 - ext[1]
    return hi, lo, best


def normalize_columns(raw: np.ndarray, hi: np.ndarray, lo: np.ndarray) -> np.ndarray:
    """NormalizeScore (gpu_plugins.go:816-841) on a column shard, given the GLOBAL extrema.  Host
    mirror of gs_normalize_dev used by the CPU tests of the sharding protocol."""
    rng = (hi - lo)[:, None]
    r = raw.astype(np.int64)
    out = np.where(rng == 0, 0, (r - lo[:, None]) * 100 // np.where(rng == 0, 1, rng))
    return out.astype(np.uint8)


class ShardedScheduler:
    """This rank's node shard on this rank's GPU.  `group` is a torch.distributed process group
    (None = default); with world size 1 no collective is issued."""

    def __init__(self, pkg, preds, cluster, device_index: int, rank: int = 0, world: int = 1, group=None):
        import torch

        self.torch = torch
        self.rank, self.world, self.group = rank, world, group
        self.begin, self.end = shard_bounds(cluster.n_nodes, world, rank)
        self.local = cluster.shard(self.begin, self.end) if world > 1 else cluster
        self.dev = torch.device("cuda", device_index)
        self.sched = pkg.Scheduler(device_index)
        self.sched.upload_predictions(preds)
        self.sched.upload_cluster(self.local, node_base=self.begin)
        self.n_local = self.end - self.begin

    def alloc_outputs(self, n_pods: int):
        t = self.torch
        return dict(score=t.empty((n_pods, self.n_local), dtype=t.uint8, device=self.dev),
                    sel=t.empty((n_pods, self.n_local), dtype=t.uint8, device=self.dev),
                    ext=t.zeros((2, n_pods), dtype=t.int64, device=self.dev),
                    best=t.empty(n_pods, dtype=t.int32, device=self.dev))

    def score(self, d_app, d_slo, n_pods: int, out: dict, normalize: bool = True):
        """Both phases + the exchange, asynchronous on torch's current stream (must be non-default)."""
        import torch.distributed as dist

        st = self.torch.cuda.current_stream().cuda_stream
        if st == 0:
            raise RuntimeError("run under a non-default torch stream: a NULL stream means the context's own stream")
        self.sched.score_raw_dev(d_app, d_slo, n_pods, out["score"], out["sel"], out["ext"][0], out["ext"][1], stream=st)
        if self.world > 1:
            dist.all_reduce(out["ext"], op=dist.ReduceOp.MAX, group=self.group)
        if normalize:
            self.sched.normalize_dev(n_pods, out["score"], out["ext"][0], out["ext"][1], out["best"], stream=st)
        return out

    def close(self):
        self.sched.close()
