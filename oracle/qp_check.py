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


def riccati_active_set(qp, act_l, act_u, restart=None):
    """numpy restatement of ONE active-set iteration as the CUDA kernel runs it (round 2):
    backward Riccati recursion with the fixed inputs eliminated, forward roll-out, and the
    multipliers of the fixed inputs from the derivative of the stage's Q-function at the rolled-out
    point,  mu_i = Quu(i,:) du_k + Qux(i,:) dx_k + g_u(i)  -- no costate sweep.
    `restart` = (c, trace of an earlier call): keep the gains of the stages >= c and restart the
    recursion from the cost-to-go (P_c, p_c) -- the kernel's checkpoints; valid when the two active
    sets agree on the stages >= c.  Returns (dx, du, lam_l, lam_u, trace), trace = {stage: (P_k, p_k),
    'K', 'kff', 'keep'}."""
    A, B, b = qp['A'], qp['B'], qp['b']
    Qd, Rd, q, r = qp['Q'], qp['R'], qp['q'], qp['r']
    lb, ub = qp['lb'], qp['ub']
    N, nx, nu = A.shape[0], A.shape[1], B.shape[2]
    al, au = np.asarray(act_l, bool), np.asarray(act_u, bool)
    K, kff = np.zeros((N, nu, nx)), np.zeros((N, nu))
    keep = {}                      # originals of the stages with a fixed input: Quu, Qux, gu
    trace = {N: (np.diag(Qd[N]), q[N].copy())}
    k_hi = N
    if restart is not None:        # trace of an earlier call + the checkpoint stage
        k_hi, old = restart
        P, p = old[k_hi]
        K[k_hi:], kff[k_hi:] = old['K'][k_hi:], old['kff'][k_hi:]
        keep.update({k: v for k, v in old['keep'].items() if k >= k_hi})
        trace[k_hi] = (P.copy(), p.copy())
    else:
        P, p = trace[N]
    for k in range(k_hi - 1, -1, -1):
        wv = P @ b[k] + p
        Quu = B[k].T @ P @ B[k] + np.diag(Rd[k])
        Qux = B[k].T @ P @ A[k]
        Qxx = A[k].T @ P @ A[k] + np.diag(Qd[k])
        gu = B[k].T @ wv + r[k]
        gx = A[k].T @ wv + q[k]
        fixed = al[k] | au[k]
        d = np.where(al[k], lb[k], np.where(au[k], ub[k], 0.0))
        Quu_e, Qux_e, gu_e, gx_e = Quu.copy(), Qux.copy(), gu.copy(), gx.copy()
        if fixed.any():
            keep[k] = (Quu.copy(), Qux.copy(), gu.copy())
            gx_e = gx + Qux[fixed].T @ d[fixed]
            gu_e = gu + Quu[:, fixed] @ d[fixed]
            gu_e[fixed] = -d[fixed]
            Qux_e[fixed] = 0.0
            Quu_e[fixed, :] = 0.0
            Quu_e[:, fixed] = 0.0
            Quu_e[fixed, fixed] = 1.0
        K[k] = -np.linalg.solve(Quu_e, Qux_e)
        kff[k] = -np.linalg.solve(Quu_e, gu_e)
        P = Qxx + Qux_e.T @ K[k]
        P = 0.5 * (P + P.T)
        p = gx_e + Qux_e.T @ kff[k]
        trace[k] = (P.copy(), p.copy())
    dx, du = np.zeros((N + 1, nx)), np.zeros((N, nu))
    mu = np.zeros((N, nu))
    for k in range(N):
        du[k] = K[k] @ dx[k] + kff[k]
        dx[k + 1] = A[k] @ dx[k] + B[k] @ du[k] + b[k]
        if k in keep and (al[k] | au[k]).any():
            Quu, Qux, gu = keep[k]
            f = al[k] | au[k]
            mu[k, f] = (Quu @ du[k] + Qux @ dx[k] + gu)[f]
    trace['K'], trace['kff'], trace['keep'] = K, kff, keep
    return dx, du, np.where(al, mu, 0.0), np.where(au, -mu, 0.0), trace
