"""Timing of the recommender completion (gs_complete_fit / gs_complete_transform) on synthetic
rank-r + noise matrices with 30 % missing entries (SURVEY 8(d) config 3 in miniature), next to
scikit-learn's IterativeImputer with the reference's parameters on the host cores.

    python profiles/bench_complete.py [--sizes 2000x64,5000x128] [--sklearn-up-to 2000]
"""
import argparse
import json
import os
import sys
import time
import warnings

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry  # noqa: E402


def synthetic(n, d, rank, p_missing, seed, return_truth=False):
    rng = np.random.default_rng(seed)
    X = rng.normal(size=(n, rank)) @ rng.normal(size=(rank, d)) + rng.normal(scale=0.01, size=(n, d))
    truth = X.copy() if return_truth else None
    mask = rng.random((n, d)) < p_missing
    mask[rng.integers(0, n, size=d), np.arange(d)] = False
    X[mask] = np.nan
    return (X, truth) if return_truth else X


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sizes", default="2000x64,5000x128,10000x256")
    ap.add_argument("--sklearn-up-to", type=int, default=2000)
    args = ap.parse_args()
    pkg = entry.load_package()
    sch = pkg.Scheduler(0)
    for spec in args.sizes.split(","):
        n, d = map(int, spec.split("x"))
        X, truth = synthetic(n, d, max(2, d // 8), 0.30, 5, return_truth=True)
        e0 = pkg.eig_stats()
        t0 = time.perf_counter()
        m = pkg.complete_fit(sch, X)
        t_fit = time.perf_counter() - t0
        e1 = pkg.eig_stats()
        t0 = time.perf_counter()
        out = m.transform(X)
        t_tr = time.perf_counter() - t0
        rec = {"n": n, "d": d, "missing": int(np.isnan(X).sum()), "gpu_fit_s": t_fit, "gpu_transform_s": t_tr,
               "rounds": m.n_rounds, "steps": m.n_steps, "ms_per_fit": 1e3 * t_fit / max(1, m.n_steps),
               "solver": "block_jacobi" if e1[0] > e0[0] else "tridiagonal",
               "eig_sweeps_per_fit": (e1[1] - e0[1]) / max(1, e1[0] - e0[0]),
               # sanity at sizes no CPU reference reaches: the data are low rank + N(0, 0.01^2) noise, so a
               # correct completion recovers the hidden entries to about the noise level
               "rmse_hidden_vs_truth": float(np.sqrt(np.mean((out - truth)[np.isnan(X)] ** 2))),
               "observed_unchanged": bool(np.array_equal(out[~np.isnan(X)], X[~np.isnan(X)]))}
        if n <= args.sklearn_up_to:
            from sklearn.experimental import enable_iterative_imputer  # noqa: F401
            from sklearn.impute import IterativeImputer

            with warnings.catch_warnings():
                warnings.simplefilter("ignore")
                t0 = time.perf_counter()
                imp = IterativeImputer(max_iter=10, random_state=0).fit(X)   # recommender.py:16-17
                rec["sklearn_fit_s"] = time.perf_counter() - t0
                ref = imp.transform(X)
            rec["max_rel_err_vs_sklearn"] = float(np.max(np.abs(out - ref) / np.maximum(np.abs(ref), 1e-9)))
            rec["sklearn_rounds"] = int(imp.n_iter_)
        print(json.dumps(rec), flush=True)
        m.close()
    sch.close()


if __name__ == "__main__":
    main()
