"""The arithmetic of csrc/tridiag.cu restated in numpy and checked against scikit-learn's BayesianRidge
(what the reference's Imputer runs per feature: pkg/recommender/recommender/recommender.py:15-17 ->
sklearn linear_model/_bayes.py:262-339).  No GPU: this pins the identities the kernels rely on —
  G = Q T Q^T by Householder reflections with LAPACK dsytd2's arithmetic,
  coef = Q (T + mu I)^-1 Q^T b,  gamma = m - mu tr((T + mu I)^-1) with the trace from the LDL^T and UDU^T pivots,
  sse = y.y - x.b~ - mu |x|^2,
and the lowest-mantissa-bit tag of the exchange (a published value loses at most one unit in the last place).
The kernels themselves are compared with the reference's goldens in tests/test_complete_gpu.py."""
import numpy as np
import pytest

sklearn = pytest.importorskip("sklearn")
from sklearn.linear_model import BayesianRidge  # noqa: E402


def householder_tridiagonal(G, b):
    """tri_reduce_kernel, one row set: returns d, e, the reflectors (rows of Vh, tau) and Q^T b."""
    m = G.shape[0]
    A = G.copy()
    bt = b.copy()
    dT, eT, tau = np.zeros(m), np.zeros(m), np.zeros(m)
    Vh = np.zeros((m, m))
    r = A[0].copy()
    for k in range(m - 2):
        x0 = r[k + 1]
        ss = float(np.sum(r[k + 2:] ** 2))
        xb = float(np.sum(r[k + 2:] * bt[k + 2:]))
        if ss > 0.0:
            beta = -np.copysign(np.sqrt(x0 * x0 + ss), x0)
            t = (beta - x0) / beta
            scale = 1.0 / (x0 - beta)
        else:
            beta, t, scale = x0, 0.0, 0.0
        v = np.zeros(m)
        v[k + 1] = 1.0
        v[k + 2:] = r[k + 2:] * scale
        bt[k + 1:] -= t * (scale * xb + bt[k + 1]) * v[k + 1:]
        Vh[k], tau[k], dT[k], eT[k] = v, t, r[k], beta
        p = np.zeros(m)
        p[k + 1:] = t * (A[k + 1:, k + 1:] @ v[k + 1:])
        w = p - 0.5 * t * float(p @ v) * v
        r = A[k + 1] - (w + w[k + 1] * v)                    # the next pivot row, formed by every CTA itself
        A[k + 1:, k + 1:] -= np.outer(v[k + 1:], w[k + 1:]) + np.outer(w[k + 1:], v[k + 1:])
        assert np.allclose(r[k + 1:], A[k + 1, k + 1:], rtol=0, atol=1e-9 * np.abs(G).max())
    dT[m - 2], eT[m - 2], dT[m - 1] = r[m - 2], r[m - 1], A[m - 1, m - 1]
    return dT, eT[:m - 1], Vh, tau, bt


def apply_q(Vh, tau, x):
    x = x.copy()
    for k in range(len(tau) - 3, -1, -1):
        x -= tau[k] * float(Vh[k] @ x) * Vh[k]
    return x


def solve_and_trace(dT, eT, bt, mu):
    """tri_bayes_kernel's solve(): forward pivots + substitution, backward pivots, back substitution."""
    m = len(dT)
    dl, z, rho, x = np.zeros(m), np.zeros(m), np.zeros(m), np.zeros(m)
    dl[0], z[0] = dT[0] + mu, bt[0]
    for i in range(1, m):
        l = eT[i - 1] / dl[i - 1]
        dl[i] = dT[i] + mu - l * eT[i - 1]
        z[i] = bt[i] - l * z[i - 1]
    rho[m - 1] = dT[m - 1] + mu
    for i in range(m - 2, -1, -1):
        rho[i] = dT[i] + mu - eT[i] ** 2 / rho[i + 1]
    x[m - 1] = z[m - 1] / dl[m - 1]
    for i in range(m - 2, -1, -1):
        x[i] = (z[i] - eT[i] * x[i + 1]) / dl[i]
    return x, float(np.sum(1.0 / (dl + rho - (dT + mu))))


def bayes_ridge_tridiagonal(G, b, yy, n):
    m = G.shape[0]
    dT, eT, Vh, tau, bt = householder_tridiagonal(G, b)
    alpha, lam = 1.0 / (yy / n + np.finfo(float).eps), 1.0
    old = np.zeros(m)
    for it in range(300):
        mu = lam / alpha
        x, tr = solve_and_trace(dT, eT, bt, mu)
        xb, sq = float(x @ bt), float(x @ x)
        gamma = m - mu * tr
        sse = max(yy - xb - mu * sq, 0.0)
        lam = (gamma + 2e-6) / (sq + 2e-6)
        alpha = (n - gamma + 2e-6) / (sse + 2e-6)
        if it != 0 and np.abs(apply_q(Vh, tau, old - x)).sum() < 1e-3:
            break
        old = x
    x, _ = solve_and_trace(dT, eT, bt, lam / alpha)
    return apply_q(Vh, tau, x), alpha, lam, it + 1, (dT, eT)


@pytest.mark.parametrize("n,m,rank,seed", [(300, 40, 6, 0), (120, 64, 8, 1), (50, 30, 30, 2), (40, 60, 10, 5)])
def test_tridiagonal_bayesian_ridge_equals_sklearn(n, m, rank, seed):
    rng = np.random.default_rng(seed)
    X = rng.normal(size=(n, rank)) @ rng.normal(size=(rank, m)) + 0.05 * rng.normal(size=(n, m))
    y = X @ rng.normal(size=m) * 0.1 + 0.1 * rng.normal(size=n)
    Xc, yc = X - X.mean(0), y - y.mean()
    G, b, yy = Xc.T @ Xc, Xc.T @ yc, float(yc @ yc)
    coef, alpha, lam, iters, (dT, eT) = bayes_ridge_tridiagonal(G, b, yy, n)
    T = np.diag(dT) + np.diag(eT, 1) + np.diag(eT, -1)
    ev_t, ev_g = np.linalg.eigvalsh(T), np.linalg.eigvalsh(G)
    assert np.max(np.abs(ev_t - ev_g)) < 1e-12 * ev_g.max()          # an orthogonal similarity
    br = BayesianRidge().fit(X, y)
    assert iters == br.n_iter_
    assert np.max(np.abs(coef - br.coef_)) < 1e-8 * max(1.0, np.max(np.abs(br.coef_)))
    assert abs(alpha - br.alpha_) < 1e-8 * br.alpha_ and abs(lam - br.lambda_) < 1e-8 * br.lambda_


def test_trace_of_the_inverse_from_the_two_pivot_sequences():
    rng = np.random.default_rng(3)
    m = 57
    dT, eT, bt = rng.uniform(1.0, 3.0, size=m), rng.normal(size=m - 1) * 0.7, rng.normal(size=m)
    T = np.diag(dT) + np.diag(eT, 1) + np.diag(eT, -1)
    for mu in (1e-6, 0.3, 40.0):
        x, tr = solve_and_trace(dT, eT, bt, mu)
        M = T + mu * np.eye(m)
        assert np.allclose(x, np.linalg.solve(M, bt), rtol=1e-10, atol=0)
        assert abs(tr - np.trace(np.linalg.inv(M))) < 1e-10 * tr


def test_tagged_value_costs_at_most_one_ulp_and_never_matches_the_stale_tag():
    rng = np.random.default_rng(4)
    vals = np.concatenate([rng.normal(size=1000) * 10.0 ** rng.integers(-200, 200, size=1000), [0.0, -0.0, np.inf, 1.0]])
    bits = vals.view(np.uint64)
    for tag in (0, 1):
        tagged = (bits & ~np.uint64(1)) | np.uint64(tag)
        assert np.all((tagged & np.uint64(1)) == tag)
        back = (tagged & ~np.uint64(1)).view(np.float64)
        finite = np.isfinite(vals)
        assert np.all(np.abs(back[finite] - vals[finite]) <= np.spacing(np.abs(vals[finite])))
    # the initial fill (all ones) carries tag 1: the first two steps wait for tag 0
    assert int(np.uint64(0xFFFFFFFFFFFFFFFF) & np.uint64(1)) == 1
