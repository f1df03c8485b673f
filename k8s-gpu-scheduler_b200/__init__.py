"""k8s-gpu-scheduler_b200 — B200-native Score/NormalizeScore hot path of k8s-gpu-scheduler.

Python here is a thin ctypes driver over ``libgpusched.so`` (the C ABI in
``include/gpusched.h``), used by the tests and ``bench.py`` in this toolchain-less
container; the production caller is the Go plugin over cgo.

There is no CPU path: importing works anywhere (so that symbols can be checked), but any
compute call needs a CUDA device and raises otherwise.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import synth  # noqa: F401  (re-export)
from .synth import APP_UNKNOWN, MODEL_A30, MODEL_NONE, MODEL_V100, Cluster, Pods, Predictions  # noqa: F401

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("GS_LIB_PATH") or os.path.join(_HERE, "libgpusched.so")

GS_OK = 0
SCORE_RAW, SCORE_NORMALIZE, SCORE_DIRECT, SCORE_NO_DEDUPE, SCORE_LANES_NODES = 0, 1, 2, 4, 8
SEL_NONE = 31
NOALLOC_MAX_PODS, NOALLOC_MAX_APPS, NOALLOC_MAX_RESIDENTS = 65536, 4096, 131072
ERR_NAMES = {-1: "GS_ERR_INVALID", -2: "GS_ERR_CUDA", -3: "GS_ERR_STATE", -4: "GS_ERR_NOMEM",
             -5: "GS_ERR_UNSUPPORTED", -6: "GS_ERR_NOCONV"}

# every symbol include/gpusched.h declares
EXPORTS = [
    "gs_create", "gs_destroy", "gs_last_error", "gs_version", "gs_upload_cluster", "gs_update_node", "gs_upload_predictions",
    "gs_score_batch", "gs_score_decisions", "gs_decisions_dev", "gs_best_dev", "gs_create_multi", "gs_destroy_multi",
    "gs_multi_last_error", "gs_multi_device_count", "gs_mupload_predictions", "gs_mupload_cluster", "gs_mscore_batch",
    "gs_mscore_decisions", "gs_score_raw_dev", "gs_score_raw_dev_noalloc", "gs_normalize_dev", "gs_normalize_scores", "gs_select_mig",
    "gs_set_profiling", "gs_last_score_kernel_ms", "gs_last_stage_a_ms", "gs_selftest", "gs_selftest_filter", "gs_launch_count", "gs_complete_fit", "gs_complete_fit_sharded", "gs_complete_fit_transform", "gs_complete_transform", "gs_model_info", "gs_model_step", "gs_model_destroy", "gs_eig_stats", "gs_complete_last_timing",
]


class GsError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"{ERR_NAMES.get(code, code)}: {msg}")
        self.code = code


class _GsCluster(C.Structure):
    _fields_ = [("n_nodes", C.c_int32), ("max_gpus", C.c_int32), ("node_model", C.c_void_p),
                ("node_ngpu", C.c_void_p), ("gpu_off", C.c_void_p), ("res_app", C.c_void_p),
                ("res_slo", C.c_void_p), ("redis_reachable", C.c_int32)]


class _GsPredictions(C.Structure):
    _fields_ = [("n_apps", C.c_int32), ("n_cols", C.c_int32), ("conf", C.c_void_p), ("interf", C.c_void_p),
                ("col_of", C.c_void_p)]


# int (*gs_allreduce_fn)(double* device_buffer, int64_t count, int op, void* stream, void* user)
ALLREDUCE_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int64, C.c_int, C.c_void_p, C.c_void_p)
REDUCE_SUM, REDUCE_MAX = 0, 1

_lib = None


def load_library():
    """dlopen libgpusched.so (built in-tree by build.py); fails loudly if it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} not built: run `python k8s-gpu-scheduler_b200/build.py` "
                          "(there is no CPU fallback)")
    L = C.CDLL(LIB_PATH)
    vp, i32, u32 = C.c_void_p, C.c_int32, C.c_uint32
    L.gs_create.argtypes = [C.c_int, C.POINTER(vp)]
    L.gs_destroy.argtypes = [vp]
    L.gs_destroy.restype = None
    L.gs_last_error.argtypes = [vp]
    L.gs_last_error.restype = C.c_char_p
    L.gs_version.restype = C.c_char_p
    L.gs_launch_count.restype = C.c_uint64
    L.gs_upload_cluster.argtypes = [vp, C.POINTER(_GsCluster), i32]
    L.gs_upload_predictions.argtypes = [vp, C.POINTER(_GsPredictions)]
    L.gs_update_node.argtypes = [vp, i32, i32, vp, vp, vp]
    L.gs_score_batch.argtypes = [vp, vp, vp, vp, i32, u32, vp, vp, vp]
    L.gs_score_decisions.argtypes = [vp, vp, vp, vp, i32, u32, vp]
    L.gs_decisions_dev.argtypes = [vp, vp, vp, i32, vp, vp, vp, vp, vp]
    L.gs_best_dev.argtypes = [vp, i32, vp, vp, vp, vp]
    L.gs_create_multi.argtypes = [C.POINTER(C.c_int), C.c_int, C.POINTER(vp)]
    L.gs_destroy_multi.argtypes = [vp]
    L.gs_destroy_multi.restype = None
    L.gs_multi_last_error.argtypes = [vp]
    L.gs_multi_last_error.restype = C.c_char_p
    L.gs_multi_device_count.argtypes = [vp]
    L.gs_mupload_predictions.argtypes = [vp, C.POINTER(_GsPredictions)]
    L.gs_mupload_cluster.argtypes = [vp, C.POINTER(_GsCluster)]
    L.gs_mscore_batch.argtypes = [vp, vp, vp, i32, u32, vp, vp, vp]
    L.gs_mscore_decisions.argtypes = [vp, vp, vp, i32, u32, vp]
    L.gs_score_raw_dev.argtypes = [vp, vp, vp, vp, i32, u32, vp, vp, vp, vp, vp]
    L.gs_score_raw_dev_noalloc.argtypes = [vp, vp, vp, i32, vp, vp, vp, vp, vp]
    L.gs_normalize_dev.argtypes = [vp, i32, vp, vp, vp, vp, vp]
    L.gs_normalize_scores.argtypes = [vp, vp, i32]
    L.gs_select_mig.argtypes = [vp, vp, vp, i32, vp]
    L.gs_set_profiling.argtypes = [vp, C.c_int]
    L.gs_last_score_kernel_ms.argtypes = [vp, C.POINTER(C.c_float)]
    L.gs_last_stage_a_ms.argtypes = [vp, C.POINTER(C.c_float)]
    L.gs_selftest.argtypes = [vp, C.c_uint64, i32, C.POINTER(C.c_uint64)]
    L.gs_selftest_filter.argtypes = [vp, C.c_uint64, i32, C.POINTER(C.c_double), C.POINTER(C.c_double),
                                     C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
    L.gs_complete_fit.argtypes = [vp, vp, i32, i32, C.POINTER(vp)]
    L.gs_complete_fit_sharded.argtypes = [vp, vp, i32, i32, ALLREDUCE_FN, vp, C.POINTER(vp)]
    L.gs_complete_fit_transform.argtypes = [vp, vp, i32, i32, ALLREDUCE_FN, vp, vp, C.POINTER(vp)]
    L.gs_complete_transform.argtypes = [vp, vp, vp, i32, vp]
    L.gs_model_info.argtypes = [vp, C.POINTER(i32), C.POINTER(i32), C.POINTER(i32)]
    L.gs_model_step.argtypes = [vp, i32, C.POINTER(i32), vp, C.POINTER(C.c_double)]
    L.gs_model_destroy.argtypes = [vp]
    L.gs_model_destroy.restype = None
    L.gs_complete_last_timing.argtypes = [C.c_void_p, C.POINTER(C.c_double)]
    L.gs_complete_last_timing.restype = C.c_int
    L.gs_eig_stats.argtypes = [C.POINTER(C.c_uint64)]
    L.gs_eig_stats.restype = C.c_uint64
    _lib = L
    return L


def eig_stats():
    """(decompositions, sweeps) the block Jacobi eigensolver has run so far in this process."""
    sw = C.c_uint64(0)
    return int(load_library().gs_eig_stats(C.byref(sw))), int(sw.value)


def launch_count() -> int:
    """Kernels launched by the library so far in this process (gs_launch_count)."""
    return int(load_library().gs_launch_count())


def _ptr(a):
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        return a.ctypes.data_as(C.c_void_p)
    if hasattr(a, "data_ptr"):  # torch tensor (host pinned or device)
        return C.c_void_p(a.data_ptr())
    return C.c_void_p(int(a))


class Scheduler:
    """One context on one CUDA device: tables in HBM + the batched scoring entry points."""

    def __init__(self, device: int = 0):
        self._L = load_library()
        h = C.c_void_p()
        rc = self._L.gs_create(int(device), C.byref(h))
        if rc != GS_OK:
            raise GsError(rc, f"gs_create(device={device}) failed — no usable CUDA device? (no CPU fallback)")
        self._h = h
        self.device = int(device)
        self.n_nodes = 0
        self.n_res = 0

    # -- plumbing ---------------------------------------------------------------------
    def _check(self, rc: int):
        if rc != GS_OK:
            raise GsError(rc, self._L.gs_last_error(self._h).decode())

    def close(self):
        if getattr(self, "_h", None):
            self._L.gs_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- uploads ----------------------------------------------------------------------
    def upload_cluster(self, cl: Cluster, node_base: int = 0):
        keep = [np.ascontiguousarray(cl.node_model, dtype=np.uint8), np.ascontiguousarray(cl.node_ngpu, dtype=np.uint8),
                np.ascontiguousarray(cl.gpu_off, dtype=np.uint32), np.ascontiguousarray(cl.res_app, dtype=np.uint32),
                np.ascontiguousarray(cl.res_slo, dtype=np.float32)]
        s = _GsCluster(int(keep[0].shape[0]), int(cl.max_gpus), *[_ptr(a) for a in keep], int(cl.redis_reachable))
        self._check(self._L.gs_upload_cluster(self._h, C.byref(s), int(node_base)))
        self.n_nodes = int(keep[0].shape[0])
        self.n_res = int(keep[3].shape[0])

    def update_node(self, node: int, gpu_counts, res_app, res_slo):
        """gs_update_node: replace one node's GPU list and residents in place."""
        counts = np.ascontiguousarray(gpu_counts, dtype=np.uint32)
        app = np.ascontiguousarray(res_app, dtype=np.uint32)
        slo = np.ascontiguousarray(res_slo, dtype=np.float32)
        self._check(self._L.gs_update_node(self._h, int(node), int(counts.shape[0]), _ptr(counts), _ptr(app), _ptr(slo)))

    def upload_predictions(self, pr: Predictions):
        conf = np.ascontiguousarray(pr.conf, dtype=np.float32)
        interf = np.ascontiguousarray(pr.interf, dtype=np.float32)
        col_of = np.ascontiguousarray(pr.col_of, dtype=np.int32)
        s = _GsPredictions(int(conf.shape[0]), int(conf.shape[1]) if conf.ndim == 2 else 0, _ptr(conf), _ptr(interf),
                           _ptr(col_of))
        self._check(self._L.gs_upload_predictions(self._h, C.byref(s)))

    # -- scoring, host buffers --------------------------------------------------------
    def score_batch(self, pods: Pods, normalize: bool = True, direct: bool = False, out=None, want_matrices=True,
                    flags: int = 0):
        """Score + (optionally) NormalizeScore for every (pod,node): returns (score, sel, best)."""
        app = np.ascontiguousarray(pods.app, dtype=np.uint32)
        slo = np.ascontiguousarray(pods.slo, dtype=np.float32)
        self_slot = None if pods.self_slot is None else np.ascontiguousarray(pods.self_slot, dtype=np.int32)
        P, N = int(app.shape[0]), self.n_nodes
        if out is not None:
            score, sel, best = out
        elif want_matrices:
            score = np.empty((P, N), dtype=np.uint8)
            sel = np.empty((P, N), dtype=np.uint8)
            best = np.empty(P, dtype=np.int32)
        else:
            score = sel = None
            best = np.empty(P, dtype=np.int32)
        flags = int(flags) | (SCORE_NORMALIZE if normalize else 0) | (SCORE_DIRECT if direct else 0)
        self._check(self._L.gs_score_batch(self._h, _ptr(app), _ptr(slo), _ptr(self_slot), P, flags, _ptr(score),
                                           _ptr(sel), _ptr(best)))
        return score, sel, best

    DECISION_DTYPE = np.dtype([("node", "<i4"), ("sel", "u1"), ("raw_hi", "u1"), ("raw_lo", "u1"), ("mps", "u1")])

    def score_decisions(self, pods: Pods, flags: int = 0, out=None) -> np.ndarray:
        """gs_score_decisions: per pod the chosen node, the selection byte there, the raw score extrema
        and the V100 MPS configuration — 8 bytes per pod."""
        app = np.ascontiguousarray(pods.app, dtype=np.uint32)
        slo = np.ascontiguousarray(pods.slo, dtype=np.float32)
        self_slot = None if pods.self_slot is None else np.ascontiguousarray(pods.self_slot, dtype=np.int32)
        if out is None:
            out = np.empty(app.shape[0], dtype=self.DECISION_DTYPE)
        self._check(self._L.gs_score_decisions(self._h, _ptr(app), _ptr(slo), _ptr(self_slot), int(app.shape[0]),
                                               int(flags), _ptr(out)))
        return out

    # -- scoring, device pointers (torch tensors or raw addresses) ----------------------
    def score_raw_dev(self, d_app, d_slo, n_pods, d_score, d_sel, d_hi_arg, d_neg_lo, d_self=None, direct=False,
                      stream=None, flags: int = 0):
        self._check(self._L.gs_score_raw_dev(self._h, _ptr(d_app), _ptr(d_slo), _ptr(d_self), int(n_pods),
                                             int(flags) | (SCORE_DIRECT if direct else 0), _ptr(d_score), _ptr(d_sel),
                                             _ptr(d_hi_arg), _ptr(d_neg_lo), _ptr(stream)))

    def score_raw_dev_noalloc(self, d_app, d_slo, n_pods, d_score, d_sel, d_hi_arg, d_neg_lo, stream=None):
        """gs_score_raw_dev_noalloc: same results as score_raw_dev, all working memory taken from the
        library's fixed device array (no allocation); refuses batches beyond GS_NOALLOC_MAX_*."""
        self._check(self._L.gs_score_raw_dev_noalloc(self._h, _ptr(d_app), _ptr(d_slo), int(n_pods), _ptr(d_score),
                                                     _ptr(d_sel), _ptr(d_hi_arg), _ptr(d_neg_lo), _ptr(stream)))

    def decisions_dev(self, d_app, d_slo, n_pods, d_sel, d_hi_arg, d_neg_lo, d_out, stream=None):
        """gs_decisions_dev: 8-byte decisions from the selection matrix and the (global) extrema."""
        self._check(self._L.gs_decisions_dev(self._h, _ptr(d_app), _ptr(d_slo), int(n_pods), _ptr(d_sel), _ptr(d_hi_arg),
                                             _ptr(d_neg_lo), _ptr(d_out), _ptr(stream)))

    def normalize_dev(self, n_pods, d_score, d_hi_arg, d_neg_lo, d_best=None, stream=None):
        self._check(self._L.gs_normalize_dev(self._h, int(n_pods), _ptr(d_score), _ptr(d_hi_arg), _ptr(d_neg_lo),
                                             _ptr(d_best), _ptr(stream)))

    def set_profiling(self, enabled: bool):
        self._check(self._L.gs_set_profiling(self._h, int(bool(enabled))))

    def last_score_kernel_ms(self) -> float:
        ms = C.c_float(0)
        self._check(self._L.gs_last_score_kernel_ms(self._h, C.byref(ms)))
        return float(ms.value)

    def last_stage_a_ms(self) -> float:
        ms = C.c_float(0)
        self._check(self._L.gs_last_stage_a_ms(self._h, C.byref(ms)))
        return float(ms.value)

    def complete_last_timing(self) -> dict:
        """Device time of the last completion fit on this context, per estimator fit (sampled with CUDA events)."""
        out = (C.c_double * 4)()
        self._check(self._L.gs_complete_last_timing(self._h, out))
        return {"fits_sampled": int(out[0]), "moments_ms": out[1], "gram_ms": out[2], "estimator_ms": out[3]}

    def selftest(self, seed: int = 1, n_million: int = 100) -> int:
        bad = C.c_uint64(0)
        self._check(self._L.gs_selftest(self._h, int(seed), int(n_million), C.byref(bad)))
        return int(bad.value)

    def selftest_filter(self, seed: int = 1, n_million: int = 100) -> dict:
        """gs_selftest_filter: float32 filter estimate vs the exact float64 tmpScore, on the device."""
        dev, delta = C.c_double(0), C.c_double(0)
        wrong, band, n = C.c_uint64(0), C.c_uint64(0), C.c_uint64(0)
        self._check(self._L.gs_selftest_filter(self._h, int(seed), int(n_million), C.byref(dev), C.byref(delta),
                                               C.byref(wrong), C.byref(band), C.byref(n)))
        return {"max_dev": dev.value, "delta": delta.value, "wrong_floors": int(wrong.value),
                "in_band": int(band.value), "evaluated": int(n.value)}

    # -- the other entry points -------------------------------------------------------
    def normalize_scores(self, scores) -> np.ndarray:
        a = np.ascontiguousarray(scores, dtype=np.int64).copy()
        self._check(self._L.gs_normalize_scores(self._h, _ptr(a), int(a.shape[0])))
        return a

    def select_mig(self, pods: Pods) -> np.ndarray:
        app = np.ascontiguousarray(pods.app, dtype=np.uint32)
        slo = np.ascontiguousarray(pods.slo, dtype=np.float32)
        out = np.empty(app.shape[0], dtype=np.uint8)
        self._check(self._L.gs_select_mig(self._h, _ptr(app), _ptr(slo), int(app.shape[0]), _ptr(out)))
        return out


class MultiScheduler:
    """gs_mctx: node-sharded scoring over several GPUs of one box in this process, through the C ABI alone
    (nodes block-partitioned inside the library; extrema merged over the peer mapping)."""

    def __init__(self, devices):
        self._L = load_library()
        devs = (C.c_int * len(devices))(*[int(d) for d in devices])
        h = C.c_void_p()
        rc = self._L.gs_create_multi(devs, len(devices), C.byref(h))
        if rc != GS_OK:
            raise GsError(rc, f"gs_create_multi({list(devices)}) failed (no peer access, or no such device)")
        self._h = h
        self.n_nodes = 0

    def _check(self, rc):
        if rc != GS_OK:
            raise GsError(rc, self._L.gs_multi_last_error(self._h).decode())

    def close(self):
        if getattr(self, "_h", None):
            self._L.gs_destroy_multi(self._h)
            self._h = None

    def upload_predictions(self, pr: Predictions):
        conf = np.ascontiguousarray(pr.conf, dtype=np.float32)
        interf = np.ascontiguousarray(pr.interf, dtype=np.float32)
        col_of = np.ascontiguousarray(pr.col_of, dtype=np.int32)
        s = _GsPredictions(int(conf.shape[0]), int(conf.shape[1]) if conf.ndim == 2 else 0, _ptr(conf), _ptr(interf),
                           _ptr(col_of))
        self._check(self._L.gs_mupload_predictions(self._h, C.byref(s)))

    def upload_cluster(self, cl: Cluster):
        keep = [np.ascontiguousarray(cl.node_model, dtype=np.uint8), np.ascontiguousarray(cl.node_ngpu, dtype=np.uint8),
                np.ascontiguousarray(cl.gpu_off, dtype=np.uint32), np.ascontiguousarray(cl.res_app, dtype=np.uint32),
                np.ascontiguousarray(cl.res_slo, dtype=np.float32)]
        s = _GsCluster(int(keep[0].shape[0]), int(cl.max_gpus), *[_ptr(a) for a in keep], int(cl.redis_reachable))
        self._check(self._L.gs_mupload_cluster(self._h, C.byref(s)))
        self.n_nodes = int(keep[0].shape[0])

    def score_batch(self, pods: Pods, normalize: bool = True, flags: int = 0, out=None):
        app = np.ascontiguousarray(pods.app, dtype=np.uint32)
        slo = np.ascontiguousarray(pods.slo, dtype=np.float32)
        P, N = int(app.shape[0]), self.n_nodes
        score, sel, best = out if out is not None else (np.empty((P, N), np.uint8), np.empty((P, N), np.uint8),
                                                        np.empty(P, np.int32))
        self._check(self._L.gs_mscore_batch(self._h, _ptr(app), _ptr(slo), P, int(flags) | (SCORE_NORMALIZE if normalize else 0),
                                            _ptr(score), _ptr(sel), _ptr(best)))
        return score, sel, best

    def score_decisions(self, pods: Pods, flags: int = 0) -> np.ndarray:
        app = np.ascontiguousarray(pods.app, dtype=np.uint32)
        slo = np.ascontiguousarray(pods.slo, dtype=np.float32)
        out = np.empty(app.shape[0], dtype=Scheduler.DECISION_DTYPE)
        self._check(self._L.gs_mscore_decisions(self._h, _ptr(app), _ptr(slo), int(app.shape[0]), int(flags), _ptr(out)))
        return out


class CompletionModel:
    """A fitted recommender completion model (gs_model): the reference's Imputer on the GPU."""

    def __init__(self, sched: "Scheduler", handle):
        self._s, self._h = sched, handle
        d, r, n = C.c_int32(), C.c_int32(), C.c_int32()
        sched._L.gs_model_info(handle, C.byref(d), C.byref(r), C.byref(n))
        self.n_features, self.n_rounds, self.n_steps = d.value, r.value, n.value

    def transform(self, X) -> np.ndarray:
        X = np.ascontiguousarray(X, dtype=np.float64)
        if X.ndim != 2 or X.shape[1] != self.n_features:
            raise ValueError(f"expected [m, {self.n_features}]")
        out = np.empty_like(X)
        self._s._check(self._s._L.gs_complete_transform(self._s._h, self._h, _ptr(X), int(X.shape[0]), _ptr(out)))
        return out

    def steps(self):
        """[(feature, coef[d-1], intercept)] in fitting order (IterativeImputer.imputation_sequence_)."""
        out = []
        for s in range(self.n_steps):
            f, ic = C.c_int32(), C.c_double()
            coef = np.empty(max(self.n_features - 1, 0), dtype=np.float64)
            self._s._L.gs_model_step(self._h, s, C.byref(f), _ptr(coef), C.byref(ic))
            out.append((f.value, coef, ic.value))
        return out

    def close(self):
        if self._h:
            self._s._L.gs_model_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def complete_fit(sched: "Scheduler", X) -> CompletionModel:
    """Imputer.fit (pkg/recommender/recommender/recommender.py:15-22) on the GPU."""
    X = np.ascontiguousarray(X, dtype=np.float64)
    h = C.c_void_p()
    sched._check(sched._L.gs_complete_fit(sched._h, _ptr(X), int(X.shape[0]), int(X.shape[1]), C.byref(h)))
    return CompletionModel(sched, h)


def complete_fit_transform(sched: "Scheduler", X):
    """Imputer.fit + the completed training matrix in one pass -> (model, completed)."""
    X = np.ascontiguousarray(X, dtype=np.float64)
    out = np.empty_like(X)
    h = C.c_void_p()
    sched._check(sched._L.gs_complete_fit_transform(sched._h, _ptr(X), int(X.shape[0]), int(X.shape[1]),
                                                    C.cast(None, ALLREDUCE_FN), None, _ptr(out), C.byref(h)))
    return CompletionModel(sched, h), out


def complete_fit_sharded(sched: "Scheduler", X_local, allreduce) -> CompletionModel:
    """Row-sharded Imputer.fit: `allreduce(ptr, count, op, stream)` must all-reduce the float64
    device buffer at `ptr` in place over the ranks (op REDUCE_SUM / REDUCE_MAX) and return 0."""
    X = np.ascontiguousarray(X_local, dtype=np.float64)

    def _cb(ptr, count, op, stream, _user):
        try:
            return int(allreduce(int(ptr), int(count), int(op), int(stream) if stream else 0) or 0)
        except Exception:  # noqa: BLE001  (never let an exception cross the C boundary)
            import traceback

            traceback.print_exc()
            return 1

    cb = ALLREDUCE_FN(_cb)
    h = C.c_void_p()
    sched._check(sched._L.gs_complete_fit_sharded(sched._h, _ptr(X), int(X.shape[0]), int(X.shape[1]), cb, None,
                                                  C.byref(h)))
    return CompletionModel(sched, h)


class DevicePointer:
    """A raw device address as a __cuda_array_interface__ object (zero-copy torch.as_tensor)."""

    def __init__(self, ptr: int, count: int, typestr: str = "<f8"):
        self.__cuda_array_interface__ = {"shape": (count,), "typestr": typestr, "data": (ptr, False), "version": 2}


def torch_allreduce(group=None):
    """An `allreduce` for complete_fit_sharded over torch.distributed (NCCL): the collective is
    enqueued on the library's stream."""
    import torch
    import torch.distributed as dist

    def _ar(ptr, count, op, stream):
        t = torch.as_tensor(DevicePointer(ptr, count), device="cuda")
        with torch.cuda.stream(torch.cuda.ExternalStream(stream)):
            dist.all_reduce(t, op=dist.ReduceOp.SUM if op == REDUCE_SUM else dist.ReduceOp.MAX, group=group)
        return 0

    return _ar


MIG_CONFIGS = ["all-4g.24gb", "all-2g.12gb", "all-1g.6gb"]  # gpu_plugins.go:52
