"""Independent checks for the oracle's QP solve (test infrastructure, NOT the product).

Three things that do not share code with the Riccati interior-point method:
  * ``active_set_kkt`` -- for a GIVEN active set, the equality-constrained KKT system solved
    densely (numpy ``solve`` on the condensed Hessian): the exact vertex the interior point
    converges to, with its multipliers;
  * ``condensed_active_set`` -- eliminates the states (dense condensing, numpy) and solves the
    resulting 4N-variable box QP with scipy's BVLS active-set method to machine precision;
  * ``kkt_residuals`` -- stationarity / dynamics / complementarity residuals of a candidate
    solution using a costate recursion.
"""
from __future__ import annotations

import numpy as np

NX, NU = 13, 4


def condense(qp):
    """x_k = G_k U + c_k  (dx_0 = 0).  Returns H (4N x 4N), g, with cost 1/2 U'HU + g'U."""
    A, B, b = qp['A'], qp['B'], qp['b']
    N = A.shape[0]
    G = np.zeros((N + 1, NX, N * NU))
    c = np.zeros((N + 1, NX))
    for k in range(N):
        G[k + 1] = A[k] @ G[k]
        G[k + 1][:, k * NU:(k + 1) * NU] += B[k]
        c[k + 1] = A[k] @ c[k] + b[k]
    H = np.diag(qp['R'].reshape(-1)).astype(float)
    g = qp['r'].reshape(-1).astype(float).copy()
    for k in range(1, N + 1):
        Qk = qp['Q'][k]
        H += G[k].T @ (Qk[:, None] * G[k])
        g += G[k].T @ (Qk * c[k] + qp['q'][k])
    return H, g, G, c


def condensed_active_set(qp, max_iter=None):
    """Exact solution of the box QP on the condensed problem by scipy's BVLS active-set
    method (bounded-variable least squares on the Cholesky factor of the condensed Hessian),
    followed by one Newton polish on the identified free set."""
    from scipy.linalg import cholesky, solve_triangular
    from scipy.optimize import lsq_linear
    H, g, G, c = condense(qp)
    lb, ub = qp['lb'].reshape(-1), qp['ub'].reshape(-1)
    L = cholesky(H, lower=True)
    y = -solve_triangular(L, g, lower=True)
    res = lsq_linear(L.T, y, bounds=(lb, ub), method='bvls', tol=1e-15, max_iter=max_iter)
    U = res.x.copy()
    # polish: exact reduced solve on the free set found by BVLS
    free = res.active_mask == 0
    U[res.active_mask == -1] = lb[res.active_mask == -1]
    U[res.active_mask == 1] = ub[res.active_mask == 1]
    if free.any():
        rhs = -(g[free] + H[np.ix_(free, ~free)] @ U[~free])
        U[free] = np.linalg.solve(H[np.ix_(free, free)], rhs)
    N = qp['A'].shape[0]
    dx = np.einsum('kij,j->ki', G, U) + c
    return dx, U.reshape(N, NU)


def active_set_kkt(qp, act_l, act_u):
    """Exact solution for a given active set: inputs with act_l / act_u (N x 4 boolean) sit on
    their lower / upper bound, the others solve the reduced stationarity system directly.
    Returns (dx, du, lam_l, lam_u); the active set is the optimal one iff du is feasible and the
    multipliers of the active rows are >= 0."""
    H, g, G, c = condense(qp)
    lb, ub = qp['lb'].reshape(-1), qp['ub'].reshape(-1)
    al, au = np.asarray(act_l, bool).reshape(-1), np.asarray(act_u, bool).reshape(-1)
    U = np.zeros_like(g)
    U[al], U[au] = lb[al], ub[au]
    fixed = al | au
    free = ~fixed
    if free.any():
        U[free] = np.linalg.solve(H[np.ix_(free, free)],
                                  -(g[free] + H[np.ix_(free, fixed)] @ U[fixed]))
    grad = H @ U + g                      # = lam_l - lam_u on the active rows, 0 on the free ones
    N = qp['A'].shape[0]
    dx = np.einsum('kij,j->ki', G, U) + c
    return (dx, U.reshape(N, NU), np.where(al, grad, 0.0).reshape(N, NU),
            np.where(au, -grad, 0.0).reshape(N, NU))


def kkt_residuals(qp, dx, du, lam_l, lam_u):
    """max-norm residuals of the QP KKT system for (dx, du, lam)."""
    A, B, b = qp['A'], qp['B'], qp['b']
    N = A.shape[0]
    dyn = 0.0
    for k in range(N):
        dyn = max(dyn, np.max(np.abs(A[k] @ dx[k] + B[k] @ du[k] + b[k] - dx[k + 1])))
    pi = qp['Q'][N] * dx[N] + qp['q'][N]
    stat = 0.0
    for k in range(N - 1, -1, -1):
        su = qp['R'][k] * du[k] + qp['r'][k] + B[k].T @ pi - lam_l[k] + lam_u[k]
        stat = max(stat, np.max(np.abs(su)))
        pi = qp['Q'][k] * dx[k] + qp['q'][k] + A[k].T @ pi
    comp = max(np.max(np.abs(lam_l * (du - qp['lb']))), np.max(np.abs(lam_u * (qp['ub'] - du))))
    feas = max(np.max(qp['lb'] - du), np.max(du - qp['ub']), 0.0)
    return dict(dyn=dyn, stat=stat, comp=comp, feas=feas, x0=np.max(np.abs(dx[0])))
