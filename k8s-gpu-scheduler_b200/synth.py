"""Seeded synthetic cluster snapshots, prediction tables and pending-pod batches.

The distributions follow SURVEY.md section 8(d) ("Synthetic snapshot"): value ranges are
those of the reference's own training tables
(pkg/recommender/recommender/configurations_train.ods:2-19, interference_train.ods), the
table shapes are the structure-of-arrays contract of include/gpusched.h.  Everything is a
pure function of (seed, sizes): the oracle, the CUDA path and bench.py all see the same
bytes.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np

SEED = 0xB2005EED
MODEL_NONE, MODEL_A30, MODEL_V100 = 0, 1, 2
APP_UNKNOWN = 0xFFFFFFFF
MAX_GPUS_PER_NODE = 31


@dataclass
class Predictions:
    """Completed recommender tables (what ImputeConfigurations / ImputeInterference serve)."""

    conf: np.ndarray  # f32 [A, K]
    interf: np.ndarray  # f32 [2, A, A]   [model-1, row app, column app]
    col_of: np.ndarray  # i32 [2, MAX_GPUS_PER_NODE+1]  column of "%dP_<model>" or -1
    columns: list = field(default_factory=list)

    @property
    def n_apps(self) -> int:
        return int(self.conf.shape[0])


@dataclass
class Cluster:
    """Node / GPU / resident tables (what Redis + GetSLOs yield, for every node at once)."""

    node_model: np.ndarray  # u8 [N]
    node_ngpu: np.ndarray  # u8 [N]
    gpu_off: np.ndarray  # u32 [N*G+1]
    res_app: np.ndarray  # u32 [R]
    res_slo: np.ndarray  # f32 [R]
    max_gpus: int
    redis_reachable: int = 1

    @property
    def n_nodes(self) -> int:
        return int(self.node_model.shape[0])

    def shard(self, begin: int, end: int) -> "Cluster":
        """Nodes [begin, end) with their GPU/resident slices; offsets rebased to 0."""
        G = self.max_gpus
        lo, hi = int(self.gpu_off[begin * G]), int(self.gpu_off[end * G])
        return Cluster(
            node_model=self.node_model[begin:end].copy(),
            node_ngpu=self.node_ngpu[begin:end].copy(),
            gpu_off=(self.gpu_off[begin * G : end * G + 1] - np.uint32(lo)).astype(np.uint32),
            res_app=self.res_app[lo:hi].copy(),
            res_slo=self.res_slo[lo:hi].copy(),
            max_gpus=G,
            redis_reachable=self.redis_reachable,
        )


@dataclass
class Pods:
    app: np.ndarray  # u32 [P]
    slo: np.ndarray  # f32 [P]
    self_slot: np.ndarray | None = None  # i32 [P] or None (= all -1)

    @property
    def n_pods(self) -> int:
        return int(self.app.shape[0])


def _rng(seed: int, stream: int) -> np.random.Generator:
    return np.random.Generator(np.random.Philox(key=[seed & 0xFFFFFFFFFFFFFFFF, stream]))


def make_predictions(n_apps: int = 1024, with_8p: bool = True, seed: int = SEED) -> Predictions:
    rng = _rng(seed, 1)
    parts = [1, 2, 4, 8] if with_8p else [1, 2, 4]
    columns = [f"{p}P_A30" for p in parts] + [f"{p}P_V100" for p in parts]
    K = len(columns)
    conf = np.exp(rng.uniform(np.log(20.0), np.log(800.0), size=(n_apps, K))).astype(np.float32)
    A = n_apps
    kind = rng.random(size=(2, A, A))
    interf = rng.normal(5.0, 15.0, size=(2, A, A))
    sat = np.empty((2, 1, 1))
    sat[MODEL_A30 - 1] = 400.0
    sat[MODEL_V100 - 1] = 100.0
    interf = np.where(kind < 0.10, sat, interf)
    interf = np.where(kind >= 0.95, -rng.uniform(0.0, 3.0, size=(2, A, A)), interf)
    interf = interf.astype(np.float32)
    col_of = np.full((2, MAX_GPUS_PER_NODE + 1), -1, dtype=np.int32)
    for c, name in enumerate(columns):
        p, model = name.split("P_")
        col_of[(MODEL_A30 if model == "A30" else MODEL_V100) - 1, int(p)] = c
    return Predictions(conf=conf, interf=interf, col_of=col_of, columns=columns)


def make_cluster(
    n_nodes: int,
    n_apps: int = 1024,
    gpus_per_node: int = 8,
    seed: int = SEED,
    p_unknown_app: float = 0.0,
    ragged_gpus: bool = False,
) -> Cluster:
    rng = _rng(seed, 2)
    G = gpus_per_node
    u = rng.random(n_nodes)
    node_model = np.where(u < 0.50, MODEL_A30, np.where(u < 0.95, MODEL_V100, MODEL_NONE)).astype(np.uint8)
    node_ngpu = np.full(n_nodes, G, dtype=np.uint8)
    if ragged_gpus:
        node_ngpu = rng.integers(1, G + 1, size=n_nodes).astype(np.uint8)
    node_ngpu[rng.random(n_nodes) < 0.02] = 0  # no Redis key for the node
    counts = np.clip(rng.poisson(1.5, size=(n_nodes, G)), 0, 4).astype(np.int64)
    counts[np.arange(G)[None, :] >= node_ngpu[:, None]] = 0
    gpu_off = np.zeros(n_nodes * G + 1, dtype=np.uint32)
    gpu_off[1:] = np.cumsum(counts.reshape(-1)).astype(np.uint32)
    R = int(gpu_off[-1])
    res_app = rng.integers(0, n_apps, size=R).astype(np.uint32)
    if p_unknown_app > 0:
        res_app[rng.random(R) < p_unknown_app] = APP_UNKNOWN
    slo = np.round(rng.uniform(10.0, 600.0, size=R), 2)
    slo[rng.random(R) < 0.10] = 0.0
    return Cluster(
        node_model=node_model,
        node_ngpu=node_ngpu,
        gpu_off=gpu_off,
        res_app=res_app,
        res_slo=slo.astype(np.float32),
        max_gpus=G,
    )


def make_pods(n_pods: int, n_apps: int = 1024, seed: int = SEED, p_unknown_app: float = 0.0,
              app_dist: str = "zipf", slo_levels: int = 0) -> Pods:
    """slo_levels > 0: Deployment-like batch — every app has that many distinct SLO values and its
    pods draw from them (pods of one Deployment share app and SLO, deploy/busybox/busybox.yaml:26-28);
    0: the survey's synthetic batch with continuous SLOs."""
    rng = _rng(seed, 3)
    if app_dist == "uniform":  # sensitivity runs: small, equal app buckets
        app = rng.integers(0, n_apps, size=n_pods).astype(np.uint32)
    else:
        w = 1.0 / np.arange(1, n_apps + 1, dtype=np.float64) ** 1.1  # Zipf(1.1) over the app ids
        app = rng.choice(n_apps, size=n_pods, p=w / w.sum()).astype(np.uint32)
    if p_unknown_app > 0:
        app[rng.random(n_pods) < p_unknown_app] = APP_UNKNOWN
    slo = rng.uniform(10.0, 600.0, size=n_pods)
    missing = rng.random(n_pods) < 0.10
    if slo_levels > 0:
        levels = np.round(_rng(seed, 4).uniform(10.0, 600.0, size=(n_apps + 1, slo_levels)), 1)
        row = np.where(app < n_apps, app, n_apps).astype(np.int64)
        slo = levels[row, _rng(seed, 5).integers(0, slo_levels, size=n_pods)]
    slo[missing] = 0.0  # SLO env missing
    return Pods(app=app, slo=slo.astype(np.float32))


def distinct_classes(pods: Pods) -> int:
    """Number of distinct (app, SLO) classes: scores depend on a pending pod only through these."""
    key = pods.app.astype(np.uint64) << np.uint64(32) | pods.slo.view(np.uint32).astype(np.uint64)
    return int(np.unique(key).size)


def make_workload(n_pods: int, n_nodes: int, gpus_per_node: int = 8, n_apps: int = 1024, seed: int = SEED, **kw):
    preds = make_predictions(n_apps, with_8p=kw.pop("with_8p", True), seed=seed)
    cluster = make_cluster(
        n_nodes,
        n_apps,
        gpus_per_node,
        seed=seed,
        p_unknown_app=kw.get("p_unknown_app", 0.0),
        ragged_gpus=kw.get("ragged_gpus", False),
    )
    pods = make_pods(n_pods, n_apps, seed=seed, p_unknown_app=kw.get("p_unknown_app", 0.0))
    return preds, cluster, pods
