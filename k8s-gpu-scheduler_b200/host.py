"""ctypes driver of the host-side plugin mirror (include/gpusched_host.h, csrc/host/gpu_plugin.cpp):
the reference plugin's string world -> tables, and Name / PreScore / Score / NormalizeScore."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _GsCluster, _GsPredictions, load_library
from .synth import MAX_GPUS_PER_NODE, Cluster, Predictions


def _lib():
    L = load_library()
    if getattr(L, "_gsp_ready", False):
        return L
    vp, cp = C.c_void_p, C.c_char_p
    L.gsp_builder_create.restype = vp
    L.gsp_builder_destroy.argtypes = [vp]
    L.gsp_builder_destroy.restype = None
    L.gsp_builder_set_configurations.argtypes = [vp, C.POINTER(cp), C.c_int, C.POINTER(cp), C.c_int, vp]
    L.gsp_builder_set_interference.argtypes = [vp, C.POINTER(cp), C.c_int, C.POINTER(cp), C.c_int, vp]
    L.gsp_builder_add_node.argtypes = [vp, cp, cp]
    L.gsp_builder_add_pod.argtypes = [vp, cp, cp, cp, cp, cp]
    L.gsp_builder_set_redis_reachable.argtypes = [vp, C.c_int]
    L.gsp_builder_set_redis_reachable.restype = None
    L.gsp_builder_app_of.argtypes = [vp, cp]
    L.gsp_builder_app_of.restype = C.c_uint32
    L.gsp_parse_slo.argtypes = [cp, C.POINTER(C.c_int)]
    L.gsp_parse_slo.restype = C.c_float
    L.gsp_builder_build.argtypes = [vp]
    L.gsp_builder_cluster.argtypes = [vp]
    L.gsp_builder_cluster.restype = C.POINTER(_GsCluster)
    L.gsp_builder_predictions.argtypes = [vp]
    L.gsp_builder_predictions.restype = C.POINTER(_GsPredictions)
    L.gsp_builder_last_error.argtypes = [vp]
    L.gsp_builder_last_error.restype = cp
    L.gsp_plugin_create.argtypes = [vp, vp]
    L.gsp_plugin_create.restype = vp
    L.gsp_plugin_destroy.argtypes = [vp]
    L.gsp_plugin_destroy.restype = None
    L.gsp_plugin_name.restype = cp
    L.gsp_plugin_last_status.argtypes = [vp]
    L.gsp_plugin_last_status.restype = cp
    L.gsp_plugin_prescore.argtypes = [vp, cp, cp]
    L.gsp_plugin_score.argtypes = [vp, cp, C.POINTER(C.c_int64), C.c_char_p, C.c_int]
    L.gsp_plugin_normalize_score.argtypes = [vp, vp, C.c_int]
    L.gsp_plugin_score_all.argtypes = [vp, vp, C.c_int]
    L.gsp_plugin_postbind.argtypes = [vp, cp, C.c_char_p, C.c_int, C.c_char_p, C.c_int, C.c_char_p, C.c_int]
    L.gsp_builder_add_dcgm_response.argtypes = [vp, cp, cp]
    L.gsp_builder_dcgm_column.argtypes = [vp, cp, C.c_float, vp, C.c_int]
    L._gsp_ready = True
    return L


HOST_EXPORTS = [
    "gsp_builder_create", "gsp_builder_destroy", "gsp_builder_set_configurations", "gsp_builder_set_interference",
    "gsp_builder_add_node", "gsp_builder_add_pod", "gsp_builder_set_redis_reachable", "gsp_builder_app_of",
    "gsp_parse_slo", "gsp_builder_build", "gsp_builder_cluster", "gsp_builder_predictions", "gsp_builder_last_error",
    "gsp_plugin_create", "gsp_plugin_destroy", "gsp_plugin_name", "gsp_plugin_last_status", "gsp_plugin_prescore",
    "gsp_plugin_score", "gsp_plugin_score_all", "gsp_plugin_normalize_score", "gsp_plugin_postbind",
    "gsp_builder_add_dcgm_response", "gsp_builder_dcgm_column",
]


def _enc(s):
    return None if s is None else str(s).encode()


def _strs(items):
    arr = (C.c_char_p * len(items))(*[str(x).encode() for x in items])
    return arr


def parse_slo(value):
    ok = C.c_int(0)
    v = _lib().gsp_parse_slo(_enc(value), C.byref(ok))
    return float(np.float32(v)), bool(ok.value)


class SnapshotBuilder:
    def __init__(self):
        self._L = _lib()
        self._h = C.c_void_p(self._L.gsp_builder_create())

    def set_configurations(self, index, columns, values):
        v = np.ascontiguousarray(values, dtype=np.float32)
        rc = self._L.gsp_builder_set_configurations(self._h, _strs(index), len(index), _strs(columns), len(columns),
                                                    v.ctypes.data_as(C.c_void_p))
        assert rc == 0, rc

    def set_interference(self, index, columns, values):
        v = np.ascontiguousarray(values, dtype=np.float32)
        rc = self._L.gsp_builder_set_interference(self._h, _strs(index), len(index), _strs(columns), len(columns),
                                                  v.ctypes.data_as(C.c_void_p))
        assert rc == 0, rc

    def add_node(self, name, redis_value):
        rc = self._L.gsp_builder_add_node(self._h, _enc(name), _enc(redis_value))
        if rc < 0:
            raise ValueError(self._L.gsp_builder_last_error(self._h).decode())
        return rc

    def add_pod(self, name, phase="Running", cuda_visible_devices=None, slo=None, namespace="default"):
        self._L.gsp_builder_add_pod(self._h, _enc(name), _enc(namespace), _enc(phase), _enc(cuda_visible_devices), _enc(slo))

    def add_dcgm_response(self, node_name, prometheus_json) -> int:
        """One Prometheus instant-query body of the node's DCGM exporter -> number of samples."""
        rc = self._L.gsp_builder_add_dcgm_response(self._h, _enc(node_name), _enc(prometheus_json))
        if rc < 0:
            raise ValueError(self._L.gsp_builder_last_error(self._h).decode())
        return rc

    def dcgm_column(self, metric_name, default, n_nodes, max_gpus) -> np.ndarray:
        """After build(): one DCGM metric as f32[n_nodes, max_gpus] (GPUs in Redis array order)."""
        out = np.empty((n_nodes, max_gpus), dtype=np.float32)
        rc = self._L.gsp_builder_dcgm_column(self._h, _enc(metric_name), float(default), out.ctypes.data_as(C.c_void_p),
                                             int(out.size))
        if rc < 0:
            raise ValueError(self._L.gsp_builder_last_error(self._h).decode())
        return out

    def set_redis_reachable(self, flag):
        self._L.gsp_builder_set_redis_reachable(self._h, int(bool(flag)))

    def app_of(self, pod_name) -> int:
        return int(self._L.gsp_builder_app_of(self._h, _enc(pod_name)))

    def build(self):
        """-> (Predictions, Cluster) as numpy copies of the built tables."""
        assert self._L.gsp_builder_build(self._h) == 0
        cl = self._L.gsp_builder_cluster(self._h).contents
        pr = self._L.gsp_builder_predictions(self._h).contents
        N, G = cl.n_nodes, cl.max_gpus

        def arr(ptr, n, dt):
            if n == 0 or not ptr:
                return np.zeros(0, dtype=dt)
            return np.ctypeslib.as_array(C.cast(ptr, C.POINTER(np.ctypeslib.as_ctypes_type(dt))), shape=(n,)).copy()

        gpu_off = arr(cl.gpu_off, N * G + 1, np.uint32)
        R = int(gpu_off[-1]) if gpu_off.size else 0
        cluster = Cluster(node_model=arr(cl.node_model, N, np.uint8), node_ngpu=arr(cl.node_ngpu, N, np.uint8),
                          gpu_off=gpu_off if gpu_off.size else np.zeros(1, np.uint32),
                          res_app=arr(cl.res_app, R, np.uint32), res_slo=arr(cl.res_slo, R, np.float32), max_gpus=G,
                          redis_reachable=cl.redis_reachable)
        A, K = pr.n_apps, pr.n_cols
        preds = Predictions(conf=arr(pr.conf, A * K, np.float32).reshape(A, K),
                            interf=arr(pr.interf, 2 * A * A, np.float32).reshape(2, A, A),
                            col_of=arr(pr.col_of, 2 * (MAX_GPUS_PER_NODE + 1), np.int32).reshape(2, -1))
        return preds, cluster

    def close(self):
        if self._h:
            self._L.gsp_builder_destroy(self._h)
            self._h = None


class Plugin:
    """The reference plugin's surface: Name / PreScore / Score / NormalizeScore."""

    def __init__(self, sched, builder: SnapshotBuilder):
        self._L = _lib()
        self._keep = (sched, builder)
        h = self._L.gsp_plugin_create(sched._h, builder._h)
        if not h:
            raise RuntimeError("gsp_plugin_create failed: " + sched._L.gs_last_error(sched._h).decode())
        self._h = C.c_void_p(h)

    @staticmethod
    def name():
        return _lib().gsp_plugin_name().decode()

    def pre_score(self, pod_name, slo_env=None):
        rc = self._L.gsp_plugin_prescore(self._h, _enc(pod_name), _enc(slo_env))
        return rc, self._L.gsp_plugin_last_status(self._h).decode()

    def score(self, node_name):
        """-> (score, status code, status message, selected UUID)"""
        s = C.c_int64(0)
        buf = C.create_string_buffer(128)
        rc = self._L.gsp_plugin_score(self._h, _enc(node_name), C.byref(s), buf, 128)
        return int(s.value), rc, self._L.gsp_plugin_last_status(self._h).decode(), buf.value.decode()

    def postbind(self, node_name):
        """PostBind's record for the last PreScore'd pod on node_name: (would_write, CUDA_VISIBLE_DEVICES,
        CUDA_MPS_PINNED_DEVICE_MEM_LIMIT, CUDA_MPS_ACTIVE_THREAD_PERCENTAGE)."""
        u, m, t = C.create_string_buffer(128), C.create_string_buffer(32), C.create_string_buffer(16)
        rc = self._L.gsp_plugin_postbind(self._h, _enc(node_name), u, 128, m, 32, t, 16)
        if rc < 0:
            raise RuntimeError(self._L.gsp_plugin_last_status(self._h).decode())
        return bool(rc), u.value.decode(), m.value.decode(), t.value.decode()

    def score_all(self, n_nodes, out=None):
        a = out if out is not None else np.empty(n_nodes, dtype=np.int64)
        rc = self._L.gsp_plugin_score_all(self._h, a.ctypes.data_as(C.c_void_p), int(n_nodes))
        if rc != 0:
            raise RuntimeError("gsp_plugin_score_all: " + self._L.gsp_plugin_last_status(self._h).decode())
        return a

    def normalize_score(self, scores):
        a = np.ascontiguousarray(scores, dtype=np.int64).copy()
        self._L.gsp_plugin_normalize_score(self._h, a.ctypes.data_as(C.c_void_p), int(a.shape[0]))
        return a

    def close(self):
        if self._h:
            self._L.gsp_plugin_destroy(self._h)
            self._h = None
