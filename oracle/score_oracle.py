"""ctypes front-end of liboracle.so — TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
may import this module.  See score_oracle.c for what the oracle restates and why its
parity with the Go reference is "unpinned".
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


class _Cluster(C.Structure):
    _fields_ = [
        ("n_nodes", C.c_int32),
        ("max_gpus", C.c_int32),
        ("node_model", C.c_void_p),
        ("node_ngpu", C.c_void_p),
        ("gpu_off", C.c_void_p),
        ("res_app", C.c_void_p),
        ("res_slo", C.c_void_p),
        ("redis_reachable", C.c_int32),
    ]


class _Preds(C.Structure):
    _fields_ = [
        ("n_apps", C.c_int32),
        ("n_cols", C.c_int32),
        ("conf", C.c_void_p),
        ("interf", C.c_void_p),
        ("col_of", C.c_void_p),
    ]


def build(force: bool = False) -> str:
    so = os.path.join(_HERE, "liboracle.so")
    src = os.path.join(_HERE, "score_oracle.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.run(["make", "-C", _HERE, "-B", "liboracle.so"], check=True, capture_output=True)
    return so


def lib():
    global _LIB
    if _LIB is None:
        L = C.CDLL(build())
        L.orc_logic.restype = C.c_int64
        L.orc_logic.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_float, C.c_int32, C.c_int32, C.c_void_p]
        L.orc_normalize.restype = None
        L.orc_normalize.argtypes = [C.c_void_p, C.c_int32]
        L.orc_score_batch.restype = None
        L.orc_score_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32,
                                      C.c_int, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        L.orc_select_mig.restype = C.c_int
        L.orc_select_mig.argtypes = [C.c_void_p, C.c_uint32, C.c_float]
        L.orc_select_mig_batch.restype = None
        L.orc_select_mig_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p]
        L.orc_num_threads.restype = C.c_int
        _LIB = L
    return _LIB


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class Oracle:
    """Holds contiguous copies of the tables and evaluates the reference algorithm on the CPU."""

    def __init__(self, preds, cluster):
        self.conf = np.ascontiguousarray(preds.conf, dtype=np.float32)
        self.interf = np.ascontiguousarray(preds.interf, dtype=np.float32)
        self.col_of = np.ascontiguousarray(preds.col_of, dtype=np.int32)
        self._pr = _Preds(self.conf.shape[0], self.conf.shape[1], _p(self.conf), _p(self.interf), _p(self.col_of))
        self.set_cluster(cluster)

    def set_cluster(self, cluster):
        if cluster is None:
            self._cl = None
            return
        self.node_model = np.ascontiguousarray(cluster.node_model, dtype=np.uint8)
        self.node_ngpu = np.ascontiguousarray(cluster.node_ngpu, dtype=np.uint8)
        self.gpu_off = np.ascontiguousarray(cluster.gpu_off, dtype=np.uint32)
        self.res_app = np.ascontiguousarray(cluster.res_app, dtype=np.uint32)
        self.res_slo = np.ascontiguousarray(cluster.res_slo, dtype=np.float32)
        self.n_nodes = int(self.node_model.shape[0])
        self._cl = _Cluster(self.n_nodes, int(cluster.max_gpus), _p(self.node_model), _p(self.node_ngpu),
                            _p(self.gpu_off), _p(self.res_app), _p(self.res_slo), int(cluster.redis_reachable))

    def logic(self, pod_app: int, pod_slo: float, node: int, pod_self: int = -1):
        """Logic() for one (pod, node): returns (raw score, packed selection byte)."""
        sel = C.c_uint8(0)
        s = lib().orc_logic(C.byref(self._cl), C.byref(self._pr), int(pod_app), float(np.float32(pod_slo)),
                            int(pod_self), int(node), C.byref(sel))
        return int(s), int(sel.value)

    def score_batch(self, pods, normalize: bool = True, node_base: int = 0, threads: int = 0):
        app = np.ascontiguousarray(pods.app, dtype=np.uint32)
        slo = np.ascontiguousarray(pods.slo, dtype=np.float32)
        self_slot = None if getattr(pods, "self_slot", None) is None else np.ascontiguousarray(pods.self_slot, dtype=np.int32)
        P, N = app.shape[0], self.n_nodes
        score = np.zeros((P, N), dtype=np.uint8)
        sel = np.zeros((P, N), dtype=np.uint8)
        best = np.zeros(P, dtype=np.int32)
        lib().orc_score_batch(C.byref(self._cl), C.byref(self._pr), _p(app), _p(slo), _p(self_slot), P,
                              int(bool(normalize)), int(node_base), _p(score), _p(sel), _p(best), int(threads))
        return score, sel, best

    def select_mig(self, pods):
        app = np.ascontiguousarray(pods.app, dtype=np.uint32)
        slo = np.ascontiguousarray(pods.slo, dtype=np.float32)
        out = np.zeros(app.shape[0], dtype=np.uint8)
        lib().orc_select_mig_batch(C.byref(self._pr), _p(app), _p(slo), app.shape[0], _p(out))
        return out


def normalize(scores) -> np.ndarray:
    a = np.ascontiguousarray(scores, dtype=np.int64).copy()
    lib().orc_normalize(_p(a), a.shape[0])
    return a


def num_threads() -> int:
    return int(lib().orc_num_threads())
