"""The recommender's serving logic around the GPU completion.

Mirrors pkg/recommender/recom_server.py without its gRPC transport: the two training tables
(TSV with an `index` column, recom_server.py:98-102,218-220) are completed on the GPU
(gs_complete_fit / gs_complete_transform instead of scikit-learn, recommender.py:15-27), rows are
looked up the way the servicer does (`-` -> `_`, first training index contained in the request,
:67-71,:155-156), replies carry float32 results and the column names (utils.py:37-42,
recom.proto:5-17), an unknown app gets `[0]` without columns (:157-160), and a changed file (md5,
:62-64,:84-97) is re-fitted and swapped in under a lock (the reference swaps unsynchronised
globals, :141-148).
"""
from __future__ import annotations

import hashlib
import threading

import numpy as np


def find_index_for_request(request: str, index) -> str:
    """recom_server.py:67-71"""
    for ind in index:
        if ind in request:
            return ind
    return ""


def read_table(path: str):
    """-> (index names, column names, float64 matrix with NaN for empty cells)."""
    with open(path) as f:
        header = f.readline().rstrip("\n").split("\t")
        rows = [ln.rstrip("\n").split("\t") for ln in f if ln.strip()]
    index = [r[0] for r in rows]
    X = np.array([[float(v) if v not in ("", "nan", "NaN") else np.nan for v in r[1:]] for r in rows], dtype=np.float64)
    return index, header[1:], X


class _Table:
    def __init__(self, index, columns, X, model, completed):
        self.index, self.columns, self.X, self.model, self.completed = list(index), list(columns), X, model, completed


class Recommender:
    """ImputeConfigurations / ImputeInterference over GPU-completed tables."""

    def __init__(self, pkg, sched):
        self._pkg, self._sched = pkg, sched
        self._lock = threading.Lock()
        self._tables = {"configurations": None, "interference": None}
        self._paths, self._md5 = {}, {}

    # -- training (recom_server.py:215-235) ------------------------------------------------
    def fit_table(self, which: str, index, columns, X):
        # the server only ever imputes rows of its own training table: fit and completed rows in one pass
        model, completed = self._pkg.complete_fit_transform(self._sched, X)
        table = _Table(index, columns, np.asarray(X, dtype=np.float64), model, completed)
        with self._lock:
            old, self._tables[which] = self._tables[which], table
        if old is not None:
            old.model.close()
        return table

    def fit_file(self, which: str, path: str):
        index, columns, X = read_table(path)
        self._paths[which] = path
        self._md5[which] = hashlib.md5(open(path, "rb").read()).hexdigest()
        return self.fit_table(which, index, columns, X)

    def maybe_reload(self) -> list:
        """The retrain job (recom_server.py:74-134): re-fit every table whose file changed."""
        changed = []
        for which, path in self._paths.items():
            digest = hashlib.md5(open(path, "rb").read()).hexdigest()
            if digest != self._md5.get(which):
                self.fit_file(which, path)
                changed.append(which)
        return changed

    # -- serving (recom_server.py:138-204) -------------------------------------------------
    def _impute(self, which: str, request_index: str):
        with self._lock:
            table = self._tables[which]
        if table is None:
            return np.array([0.0], dtype=np.float32), []
        key = find_index_for_request(request_index.replace("-", "_"), table.index)
        if key == "":
            return np.array([0.0], dtype=np.float32), []
        row = table.completed[table.index.index(key)]
        return row.astype(np.float32), list(table.columns)      # `repeated float result`, recom.proto:15

    def impute_configurations(self, request_index: str):
        return self._impute("configurations", request_index)

    def impute_interference(self, request_index: str):
        return self._impute("interference", request_index)

    # -- bulk export for the scheduler (replaces 2*G*(R+1) RPCs per Score) -------------------
    def fill_builder(self, builder):
        """Hand both completed tables to a host.SnapshotBuilder."""
        with self._lock:
            conf, interf = self._tables["configurations"], self._tables["interference"]
        if conf is None:
            raise RuntimeError("configurations table not fitted")
        builder.set_configurations(conf.index, conf.columns, conf.completed.astype(np.float32))
        if interf is not None:
            builder.set_interference(interf.index, interf.columns, interf.completed.astype(np.float32))
