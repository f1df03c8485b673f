"""GPU parity of the recommender completion (gs_complete_fit / gs_complete_transform) against
golden vectors produced by the reference's own Imputer (tests/golden/gen_impute_golden.py).
Bar (BASELINE north_star): predictions within 1e-5 relative."""
import glob
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLD = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "impute_*.npz")))
RTOL = 1e-5


def rel_err(a, b):
    """Largest |a-b| / max(|b|, 1e-2 * mean|b|): relative error, with entries that happen to lie near
    zero measured against 1 % of the data scale instead of against themselves."""
    if not a.size:
        return 0.0
    floor = 1e-2 * float(np.mean(np.abs(b))) + 1e-300
    return float(np.max(np.abs(a - b) / np.maximum(np.abs(b), floor)))


@pytest.mark.parametrize("path", GOLD, ids=[os.path.basename(p)[7:-4] for p in GOLD])
def test_fit_and_transform_match_reference_imputer(pkg, sched, path):
    g = np.load(path)
    X = g["X"]
    m = pkg.complete_fit(sched, X)
    assert m.n_features == X.shape[1]
    assert m.n_rounds == int(g["n_iter"])
    assert m.n_steps == len(g["feats"])
    steps = m.steps()
    assert [s[0] for s in steps] == g["feats"].tolist()
    completed = m.transform(X)                       # what the recommender serves: transform(train)
    assert rel_err(completed, g["completed"]) < RTOL
    assert np.array_equal(completed[~np.isnan(X)], X[~np.isnan(X)])   # observed entries untouched
    assert rel_err(m.transform(g["Xq"]), g["q_out"]) < RTOL
    m.close()


def test_transform_is_row_independent_and_handles_complete_rows(pkg, sched):
    g = np.load([p for p in GOLD if "synth_200x12" in p][0])
    X = g["X"]
    m = pkg.complete_fit(sched, X)
    whole = m.transform(X)
    for r in (0, 5, 77):
        assert np.array_equal(m.transform(X[r:r + 1]), whole[r:r + 1])
    full = np.nan_to_num(X, nan=1.0)
    assert np.array_equal(m.transform(full), full)
    assert m.transform(np.empty((0, 12))).shape == (0, 12)
    m.close()


def test_fit_errors(pkg, sched):
    X = np.ones((5, 3))
    X[:, 1] = np.nan
    with pytest.raises(pkg.GsError) as e:
        pkg.complete_fit(sched, X)                   # a column with no observed value
    assert e.value.code == -5
    with pytest.raises(pkg.GsError) as e:
        pkg.complete_fit(sched, np.ones((4, 1100)))  # beyond the evidence iteration's shared arrays
    assert e.value.code == -5
    m = pkg.complete_fit(sched, np.arange(6.0).reshape(6, 1))   # a single column: nothing to iterate
    assert m.n_steps == 0 and m.n_rounds == 0
    m.close()
