"""Generates tests/golden/impute_*.npz by running the REFERENCE's own Imputer
(pkg/recommender/recommender/recommender.py, imported unchanged from /root/reference) on the
reference's own training tables and on small synthetic matrices.

    PYTHONDONTWRITEBYTECODE=1 python tests/golden/gen_impute_golden.py

Only runs where /root/reference exists (the development container); the fixtures it writes are
committed and are what the tests read.  The reference pins scikit-learn 1.2.1
(pkg/recommender/requirements.txt:15); the version used is recorded in every fixture.
"""
import os
import sys
import warnings

import numpy as np

sys.dont_write_bytecode = True
REF = "/root/reference/pkg/recommender"
HERE = os.path.dirname(os.path.abspath(__file__))


def synthetic(n, d, rank, p_missing, seed):
    """SURVEY 8(d) config-3 shape in miniature: rank-r product + N(0, 0.01^2) noise, uniform NaNs,
    no all-NaN column."""
    rng = np.random.default_rng(seed)
    X = rng.normal(size=(n, rank)) @ rng.normal(size=(rank, d)) + rng.normal(scale=0.01, size=(n, d))
    mask = rng.random((n, d)) < p_missing
    mask[rng.integers(0, n, size=d), np.arange(d)] = False
    X[mask] = np.nan
    return X


def main():
    sys.path.insert(0, REF)
    import pandas as pd
    import sklearn
    from recommender.recommender import Imputer

    cases = {}
    names = {}
    for name in ("configurations", "interference"):
        df = pd.read_csv(f"{REF}/recommender/{name}_train.ods", sep="\t").set_index("index")
        cases[name] = df.to_numpy(dtype=np.float64)
        names[name + "_index"] = np.array(list(df.index), dtype=str)
        names[name + "_columns"] = np.array(list(df.columns), dtype=str)
    # index / column names of the two training tables: the recommender's wire contract keys
    np.savez_compressed(os.path.join(HERE, "recommender_table_names.npz"), **names)
    cases["synth_60x8"] = synthetic(60, 8, 3, 0.30, 1)
    cases["synth_200x12"] = synthetic(200, 12, 4, 0.30, 2)
    cases["synth_40x24_wide"] = synthetic(40, 24, 3, 0.25, 3)   # n_obs < d-1 for most features
    cases["synth_500x32"] = synthetic(500, 32, 6, 0.30, 4)
    cases["synth_400x72"] = synthetic(400, 72, 8, 0.30, 6)      # beyond the Jacobi range: library eigensolver path
    cases["synth_300x136"] = synthetic(300, 136, 8, 0.25, 7)    # wide enough for the 128x128-tile Gram kernel
    cases["synth_320x264"] = synthetic(320, 264, 12, 0.25, 9)   # > 256 columns: block Jacobi on a 17-block pivot grid
    only = [a for a in sys.argv[1:] if not a.startswith("-")]
    if only:
        cases = {k: v for k, v in cases.items() if k in only}
    for name, X in cases.items():
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            imp = Imputer()
            m = imp.fit(X)
            completed = imp.predict(X)                      # what the server serves: transform(train)
            rng = np.random.default_rng(7)
            Xq = X[rng.integers(0, X.shape[0], size=min(16, X.shape[0]))].copy()
            Xq[rng.random(Xq.shape) < 0.2] = np.nan          # unseen missing patterns
            Xq[:, 0] = np.where(np.isnan(Xq).all(axis=1), 1.0, Xq[:, 0])
            q_out = imp.predict(Xq)
        feats = np.array([t.feat_idx for t in m.imputation_sequence_], dtype=np.int32)
        coefs = np.array([t.estimator.coef_ for t in m.imputation_sequence_], dtype=np.float64).reshape(len(feats), -1)
        if X.shape[1] > 200:
            coefs = coefs[-X.shape[1]:]      # the last round only: keeps the fixture small
        inter = np.array([t.estimator.intercept_ for t in m.imputation_sequence_], dtype=np.float64)
        np.savez_compressed(os.path.join(HERE, f"impute_{name}.npz"), X=X, completed=completed, n_iter=m.n_iter_,
                            feats=feats, coefs=coefs, intercepts=inter, means=m.initial_imputer_.statistics_,
                            Xq=Xq, q_out=q_out, sklearn_version=sklearn.__version__)
        print(name, X.shape, "NaN", int(np.isnan(X).sum()), "rounds", m.n_iter_, "steps", len(feats))


if __name__ == "__main__":
    main()
