"""CPU tests of the oracle (no GPU).  The reference pins nothing for this path (no tests, no
golden vectors), so the oracle is pinned by invariants:
complex-step / finite-difference Jacobians, an independent exact QP solver, KKT residuals,
fixed points, closed-loop convergence, status semantics, and its own committed goldens."""
from pathlib import Path

import numpy as np
import pytest

from tests.problems import aggressive, random_states, setpoint_refs

GOLD = Path(__file__).parent / 'golden' / 'nmpc_golden.npz'


@pytest.fixture(scope='module')
def O():
    from oracle import oracle
    oracle.build()
    return oracle


@pytest.fixture(scope='module')
def cfg(O, params):
    return O.config_from_params(params)


def f_complex(cfg, x, u):
    """Independent numpy restatement of controller.py:286-346 that accepts complex input."""
    qw, qx, qy, qz = x[6:10]
    wx, wy, wz = x[10:13]
    T = u.sum()
    m, g = cfg.mass, cfg.gravity
    Jx, Jy, Jz = cfg.J[0], cfg.J[1], cfg.J[2]
    R3 = np.array([2 * (qx * qz + qw * qy), 2 * (qy * qz - qw * qx), 1 - 2 * (qx * qx + qy * qy)])
    acc = R3 * (T / m) - np.array([0, 0, g])
    qd = 0.5 * np.array([-qx * wx - qy * wy - qz * wz, qw * wx + qy * wz - qz * wy,
                         qw * wy + qz * wx - qx * wz, qw * wz + qx * wy - qy * wx])
    ry, rx, rz = np.array(cfg.rotor_y[:]), np.array(cfg.rotor_x[:]), np.array(cfg.rotor_z[:])
    tau = np.array([u @ ry, -(u @ rx), u @ rz])
    Jw = np.array([Jx * wx, Jy * wy, Jz * wz])
    cr = np.array([wy * Jw[2] - wz * Jw[1], wz * Jw[0] - wx * Jw[2], wx * Jw[1] - wy * Jw[0]])
    wd = (tau - cr) / np.array([Jx, Jy, Jz])
    return np.concatenate([x[3:6], acc, qd, wd])


def test_bounds_and_hover(cfg):
    # controller.py:105-110 with config/params.yaml:23-26 (NOT the dead [4,20] key)
    assert abs(cfg.lbu[0] - 8.54858e-6 * 50.0 ** 2) < 1e-15
    assert abs(cfg.ubu[0] - 8.54858e-6 * 838.0 ** 2) < 1e-12
    assert abs(cfg.mass * cfg.gravity / 4 - 1.6677) < 1e-12


def test_model_matches_complex_restatement(O, cfg):
    rng = np.random.default_rng(0)
    for _ in range(20):
        x, u = rng.normal(size=13), rng.uniform(0, 6, 4)
        assert np.max(np.abs(O.f(cfg, x, u) - f_complex(cfg, x, u).real)) < 1e-13


def test_jacobian_complex_step(O, cfg):
    rng = np.random.default_rng(1)
    h = 1e-30
    for _ in range(10):
        x, u = rng.normal(size=13), rng.uniform(0, 6, 4)
        Fx, Fu = O.jac(cfg, x, u)
        for j in range(13):
            xc = x.astype(complex)
            xc[j] += 1j * h
            assert np.max(np.abs(f_complex(cfg, xc, u.astype(complex)).imag / h - Fx[:, j])) < 1e-12
        for j in range(4):
            uc = u.astype(complex)
            uc[j] += 1j * h
            assert np.max(np.abs(f_complex(cfg, x.astype(complex), uc).imag / h - Fu[:, j])) < 1e-12
        # sparsity stated in SURVEY 5.1: p never appears, dv/dq has 10 non-zeros
        assert np.count_nonzero(Fx[:, :3]) == 0
        assert np.count_nonzero(Fx[3:6, 6:10]) == 10


def _midpoint_complex(cfg, x, u, steps=2):
    h = cfg.dt / steps
    for _ in range(steps):
        k1 = f_complex(cfg, x, u)
        x = x + h * f_complex(cfg, x + 0.5 * h * k1, u)
    return x


def test_discretize_is_midpoint_with_exact_sensitivities(O, cfg):
    rng = np.random.default_rng(2)
    h = 1e-30
    for _ in range(5):
        x, u = rng.normal(size=13), rng.uniform(0, 6, 4)
        xn, A, B = O.discretize(cfg, x, u)
        assert np.max(np.abs(xn - _midpoint_complex(cfg, x.astype(complex), u.astype(complex)).real)) < 1e-13
        for j in range(13):
            xc = x.astype(complex)
            xc[j] += 1j * h
            col = _midpoint_complex(cfg, xc, u.astype(complex)).imag / h
            assert np.max(np.abs(col - A[:, j])) < 1e-12
        for j in range(4):
            uc = u.astype(complex)
            uc[j] += 1j * h
            col = _midpoint_complex(cfg, x.astype(complex), uc).imag / h
            assert np.max(np.abs(col - B[:, j])) < 1e-12
        # structure used by the design notes: dp+/dp = I, dp+/dv = dt I, dv+/dv = I
        assert np.allclose(A[:3, :3], np.eye(3)) and np.allclose(A[:3, 3:6], cfg.dt * np.eye(3))
        assert np.allclose(A[3:6, 3:6], np.eye(3)) and np.count_nonzero(A[3:, :3]) == 0


def test_rk4_order(O, params):
    """RK4 option: halving the step reduces the one-step error ~16x."""
    rng = np.random.default_rng(3)
    x, u = rng.normal(size=13) * 0.3, rng.uniform(1, 3, 4)
    x[6:10] = [1, 0.1, -0.1, 0.05]
    ref = O.discretize(O.config_from_params(params, integ_ns=4, integ_steps=64), x, u)[0]
    e1 = np.abs(O.discretize(O.config_from_params(params, integ_ns=4, integ_steps=1), x, u)[0] - ref).max()
    e2 = np.abs(O.discretize(O.config_from_params(params, integ_ns=4, integ_steps=2), x, u)[0] - ref).max()
    assert 8 < e1 / e2 < 40


def _qp_cases(O, cfg, rng, n, cold):
    N = cfg.N
    hover = cfg.mass * cfg.gravity / 4
    x0, yref, yref_e = aggressive(rng, n, N, hover)
    for b in range(n):
        if cold:
            xt, ut = np.tile(x0[b], (N + 1, 1)), np.zeros((N, 4))
        else:
            xt = np.tile(x0[b], (N + 1, 1)) + rng.normal(0, 0.1, (N + 1, 13))
            xt[0] = x0[b]
            ut = rng.uniform(0, 6, (N, 4))
        yield O.build_qp(cfg, xt, ut, yref[b], yref_e[b])


@pytest.mark.parametrize('cold', [True, False])
def test_qp_against_independent_active_set(O, cfg, cold):
    """Riccati IPM vs two solvers that share no code with it, on 24 bound-active QPs.
    (1) BVLS on the condensed QP finds the optimal active set on its own; it must be the set the
        interior point reports (multiplier > slack).
    (2) For that set the equality-constrained KKT system is solved DENSELY (numpy solve): the
        exact vertex.  The interior point stops at complementarity mu <= ipm_tol, i.e. at distance
        ~mu / lambda* from the vertex: <= 5e-10 at the default ipm_tol = 1e-14 (measured worst case
        2.5e-10), <= 1e-11 at ipm_tol = 1e-18 (worst 4.4e-12); multipliers agree likewise and are
        non-negative on every active row, the vertex is primal feasible."""
    from oracle import qp_check as Q
    rng = np.random.default_rng(4 + cold)
    nact_total = 0
    for qp in _qp_cases(O, cfg, rng, 12, cold):
        for tol, bound in ((1e-14, 5e-10), (1e-18, 1e-11)):
            s = O.qp_solve(qp, tol=tol)
            assert s['status'] == 0 and s['iters'] <= 30
            act_l = s['lam_l'] > (s['du'] - qp['lb'])
            act_u = s['lam_u'] > (qp['ub'] - s['du'])
            dx, du, ll, lu = Q.active_set_kkt(qp, act_l, act_u)
            assert np.max(np.abs(du - s['du'])) <= bound
            assert np.max(np.abs(dx - s['dx'])) <= bound
            assert np.max(np.abs(ll - s['lam_l'])) <= bound and np.max(np.abs(lu - s['lam_u'])) <= bound
            assert ll[act_l].min(initial=1.0) > 0 and lu[act_u].min(initial=1.0) > 0   # dual feasible
            assert np.all(du >= qp['lb'] - 1e-15) and np.all(du <= qp['ub'] + 1e-15)    # primal feasible
        # the active set found independently by BVLS is the same one
        dxb, dub = Q.condensed_active_set(qp)
        assert np.array_equal(dub <= qp['lb'] + 1e-9, act_l) and np.array_equal(dub >= qp['ub'] - 1e-9, act_u)
        assert np.max(np.abs(dub - du)) < 1e-10
        s = O.qp_solve(qp)
        k = Q.kkt_residuals(qp, s['dx'], s['du'], s['lam_l'], s['lam_u'])
        assert k['dyn'] < 1e-11 and k['stat'] < 1e-9 and k['comp'] < 1e-11 and k['feas'] < 1e-12
        nact_total += int(act_l.sum() + act_u.sum())
    assert nact_total > 100


def test_active_set_solve_is_the_exact_kkt_solution(O, cfg):
    """orc_qp_solve_active_set (one Riccati factorisation with the fixed inputs eliminated + costate
    sweep) against the dense KKT solve of oracle/qp_check.py for the SAME active set: primal point,
    multipliers and the accept / reject decision."""
    import ctypes as C
    from oracle import qp_check as Q
    rng = np.random.default_rng(11)
    L = O.lib()
    n_acc = 0
    for qp in _qp_cases(O, cfg, rng, 10, cold=False):
        N = qp['A'].shape[0]
        s = O.qp_solve(qp)
        act_l = s['lam_l'] > (s['du'] - qp['lb'])
        act_u = s['lam_u'] > (qp['ub'] - s['du'])
        for flip in (False, True):     # the optimal set, and a perturbed (wrong) one
            al, au = act_l.copy(), act_u.copy()
            if flip:
                al[0, 0], au[0, 0] = not al[0, 0], False
            act = (al * 1 + au * 2).astype(np.uint8).reshape(-1)
            c_qp = O.OrcQP()
            c_qp.N = N
            keep = {k: np.ascontiguousarray(qp[k], float) for k in ('A', 'B', 'b', 'Q', 'R', 'q', 'r', 'lb', 'ub')}
            for k, a in keep.items():
                setattr(c_qp, k, O._p(a))
            dx, du = np.zeros((N + 1, 13)), np.zeros((N, 4))
            ll, lu = np.zeros((N, 4)), np.zeros((N, 4))
            L.orc_qp_solve_active_set.restype = C.c_int
            rej = L.orc_qp_solve_active_set(C.byref(c_qp), act.ctypes.data_as(C.POINTER(C.c_ubyte)),
                                            O._p(dx), O._p(du), O._p(ll), O._p(lu))
            dxk, duk, llk, luk = Q.active_set_kkt(qp, al, au)
            assert np.max(np.abs(du - duk)) < 1e-10 and np.max(np.abs(dx - dxk)) < 1e-10
            assert np.max(np.abs(ll - llk)) < 1e-9 and np.max(np.abs(lu - luk)) < 1e-9
            feasible = (np.all(duk >= qp['lb'] - 1e-13) and np.all(duk <= qp['ub'] + 1e-13)
                        and llk[al].min(initial=1.0) > -1e-13 and luk[au].min(initial=1.0) > -1e-13)
            assert (rej == 0) == feasible
            n_acc += rej == 0
    assert n_acc >= 10          # every optimal set is accepted (the perturbed ones mostly not)


def test_kernel_form_of_the_active_set_iteration(O, cfg):
    """The CUDA kernel (round 2) runs an active-set iteration differently from the C oracle: the
    multipliers of the fixed inputs are the derivative of the stage's Q-function at the rolled-out
    point (no costate sweep), and a re-solve restarts the backward recursion from a cost-to-go
    checkpoint above the last changed stage.  oracle/qp_check.py::riccati_active_set restates
    that form in numpy; here it is pinned (a) against the dense KKT solve for the same set --
    primal point and multipliers -- and (b) restart vs. full sweep: identical to the bit."""
    from oracle import qp_check as Q
    rng = np.random.default_rng(12)
    n_fixed = 0
    for qp in _qp_cases(O, cfg, rng, 8, cold=False):
        N = qp['A'].shape[0]
        s = O.qp_solve(qp)
        al = s['lam_l'] > (s['du'] - qp['lb'])
        au = s['lam_u'] > (qp['ub'] - s['du'])
        n_fixed += int(al.sum() + au.sum())
        dx, du, ll, lu, tr = Q.riccati_active_set(qp, al, au)
        dxk, duk, llk, luk = Q.active_set_kkt(qp, al, au)
        assert np.max(np.abs(du - duk)) < 1e-10 and np.max(np.abs(dx - dxk)) < 1e-10
        assert np.max(np.abs(ll - llk)) < 1e-9 and np.max(np.abs(lu - luk)) < 1e-9
        # change the set at stages 0..2 only and restart from the checkpoint at stage 4 (N / 5)
        al2, au2 = al.copy(), au.copy()
        al2[0, 1], au2[0, 1] = not al2[0, 1], False
        al2[2, 3], au2[2, 3] = False, not au2[2, 3]
        full = Q.riccati_active_set(qp, al2, au2)
        part = Q.riccati_active_set(qp, al2, au2, restart=(N // 5, tr))
        for a, b in zip(full[:4], part[:4]):
            assert np.array_equal(a, b)
        dxk, duk, llk, luk = Q.active_set_kkt(qp, al2, au2)
        assert np.max(np.abs(part[1] - duk)) < 1e-10
        assert np.max(np.abs(part[2] - llk)) < 1e-9 and np.max(np.abs(part[3] - luk)) < 1e-9
    assert n_fixed > 20


@pytest.mark.parametrize('kind', ['cfg2_sqp', 'aggressive_rti'])
def test_active_set_method_matches_interior_point(O, params, kind):
    """as_shortcut on / off in the oracle: same statuses, same SQP iteration behaviour, u0 within the
    interior point's own accuracy, far fewer factorisations (the strictly convex QP has ONE solution,
    whichever method finds it)."""
    B, N = 96, 20
    rng = np.random.default_rng(12)
    hover = params['vehicle']['mass'] * 9.81 / 4
    if kind == 'cfg2_sqp':
        x0 = random_states(rng, B)
        yref, yref_e = setpoint_refs(rng, x0, N, hover)
        flags = dict(max_sqp_iter=50, sqp_tol=1e-8)
    else:
        x0, yref, yref_e = aggressive(rng, B, N, hover)
        flags = {}
    out = {}
    for on in (0, 1):
        ob = O.OracleBatch(O.config_from_params(params, as_shortcut=on, **flags), B)
        u0, st, _, _ = ob.solve(x0, yref, yref_e)
        out[on] = (u0, st, ob.qp_iters.copy())
    assert np.array_equal(out[0][1], out[1][1]) and (out[1][1] == 0).all()
    assert np.max(np.abs(out[0][0] - out[1][0])) < 5e-9
    assert out[1][2].sum() < 0.6 * out[0][2].sum()
    assert out[1][2].max() <= out[0][2].max()


def test_unconstrained_qp_is_one_factorisation(O, cfg):
    from oracle import qp_check as Q
    rng = np.random.default_rng(6)
    N, hover = cfg.N, cfg.mass * cfg.gravity / 4
    x0 = random_states(rng, 4, tilt_deg=5, v_sigma=0.1, w_sigma=0.1)
    yref, yref_e = setpoint_refs(rng, x0, N, hover, span_xy=0.3, span_z=0.1, yaw=False)
    for b in range(4):
        qp = O.build_qp(cfg, np.tile(x0[b], (N + 1, 1)), np.full((N, 4), hover), yref[b], yref_e[b])
        s = O.qp_solve(qp)
        assert s['status'] == 0 and s['iters'] == 1
        dx, du = Q.condensed_active_set(qp)
        assert np.max(np.abs(du - s['du'])) < 1e-11


def test_fixed_point_at_hover(O, cfg):
    """x = x_ref, u = hover, reference = hover equilibrium  =>  delta = 0 (SURVEY 8c iv)."""
    N, hover = cfg.N, cfg.mass * cfg.gravity / 4
    x = np.array([0.3, -0.2, 1.1, 0, 0, 0, 1, 0, 0, 0, 0, 0, 0.0])
    yref = np.zeros((N, 17))
    yref[:, :13] = x
    yref[:, 13:] = hover
    qp = O.build_qp(cfg, np.tile(x, (N + 1, 1)), np.full((N, 4), hover), yref, x.copy())
    assert np.max(np.abs(qp['b'])) < 1e-14 and np.max(np.abs(qp['q'])) < 1e-14
    s = O.qp_solve(qp)
    assert np.max(np.abs(s['du'])) < 1e-13 and np.max(np.abs(s['dx'])) < 1e-13


def test_closed_loop_cfg1(O, cfg, params):
    """config 1: hover at (0,0,1) -> set-point (2,0,1.5); thrusts stay inside the box."""
    ctl = O.OraclePositionNMPC(params)
    N = cfg.N
    ref = dict(positions=np.tile([2.0, 0.0, 1.5], (N + 1, 1)), velocities=np.zeros((N + 1, 3)),
               quaternions=np.tile([1.0, 0, 0, 0], (N + 1, 1)), body_rates=np.zeros((N + 1, 3)),
               thrusts=np.full((N, 4), ctl.hover_thrust))
    x = np.array([0, 0, 1, 0, 0, 0, 1, 0, 0, 0, 0, 0, 0.0])
    for k in range(160):
        u0, st = ctl.solve(dict(position=x[:3], velocity=x[3:6], quaternion=x[6:10],
                                body_rates=x[10:]), ref)
        assert st == 0
        assert np.all(u0 >= cfg.lbu[0] - 1e-9) and np.all(u0 <= cfg.ubu[0] + 1e-9)
        x = O.plant_step(cfg, x, u0, cfg.dt, 5)
    assert np.linalg.norm(x[:3] - [2.0, 0.0, 1.5]) < 1e-3
    with pytest.raises(ValueError):
        ctl.solve(dict(position=x[:3], velocity=x[3:6], quaternion=np.zeros(4), body_rates=x[10:]), ref)


def test_status_and_warm_start_semantics(O, cfg):
    """controller.py:419-431 (cold = x0 / u = 0 everywhere, warm = un-shifted previous solution),
    :448-461 (failure -> zeros, invalidate; success -> cache) and the node's fail-safe."""
    N, hover = cfg.N, cfg.mass * cfg.gravity / 4
    rng = np.random.default_rng(7)
    x0 = random_states(rng, 3)
    yref, yref_e = setpoint_refs(rng, x0, N, hover)
    ob = O.OracleBatch(cfg, 3)
    assert np.allclose(ob.last_cmd, hover)                 # mpc_controller_node:64
    u0, st, th, sp = ob.solve(x0, yref, yref_e)
    assert np.all(st == 0) and ob.valid.all()
    assert np.allclose(ob.xtraj[:, 0], x0)                 # stage-0 state is x0
    # cold start really linearised at u = 0: redo by hand for problem 0
    qp = O.build_qp(cfg, np.tile(x0[0], (N + 1, 1)), np.zeros((N, 4)), yref[0], yref_e[0])
    s = O.qp_solve(qp)
    assert np.max(np.abs(s['du'][0] - u0[0])) < 1e-12
    # warm start: previous trajectories, stage 0 overwritten with the new x0
    xt_prev, ut_prev = ob.xtraj.copy(), ob.utraj.copy()
    x1 = x0 + 0.01
    u1, st, _, _ = ob.solve(x1, yref, yref_e)
    xt = xt_prev[1].copy()
    xt[0] = x1[1]
    s = O.qp_solve(O.build_qp(cfg, xt, ut_prev[1], yref[1], yref_e[1]))
    assert np.max(np.abs(ut_prev[1, 0] + s['du'][0] - u1[1])) < 1e-12
    # failure: NaN state -> status 4, zeros, invalid, last command republished
    th_before = ob.last_cmd.copy()
    bad = x1.copy()
    bad[2, 5] = np.nan
    u2, st, th, sp = ob.solve(bad, yref, yref_e)
    assert st[2] == 4 and np.all(u2[2] == 0) and not ob.valid[2] and ob.valid[:2].all()
    assert np.allclose(th[2], th_before[2])
    # command map: clip + sqrt(f/kf) clipped to the motor speed range
    kf = cfg.kf
    assert np.allclose(sp[0], np.clip(np.sqrt(np.clip(th[0], cfg.lbu[0], cfg.ubu[0]) / kf),
                                      cfg.w_min, cfg.w_max))


def test_sqp_converges_and_flags(O, params):
    cfg = O.config_from_params(params, max_sqp_iter=50, sqp_tol=1e-8)
    rng = np.random.default_rng(8)
    x0 = random_states(rng, 6)
    yref, yref_e = setpoint_refs(rng, x0, cfg.N, cfg.mass * cfg.gravity / 4)
    ob = O.OracleBatch(cfg, 6)
    u0, st, _, _ = ob.solve(x0, yref, yref_e)
    assert np.all(st == 0)
    for b in range(6):   # SQP fixed point = dynamically consistent trajectory
        for k in range(cfg.N):
            xn, _, _ = O.discretize(cfg, ob.xtraj[b, k], ob.utraj[b, k])
            assert np.max(np.abs(xn - ob.xtraj[b, k + 1])) < 1e-7
    capped = O.OracleBatch(O.config_from_params(params, max_sqp_iter=2, sqp_tol=1e-12), 6)
    _, st2, _, _ = capped.solve(x0, yref, yref_e)
    assert np.all(st2 == 2)                                # ACADOS_MAXITER (extension)


def test_goldens_reproduce(O, params, cfg):
    g = np.load(GOLD)
    for tag, kw in (('mid', {}), ('rk4', dict(integ_ns=4, integ_steps=1))):
        c = O.config_from_params(params, **kw)
        for i in range(g['disc_x'].shape[0]):
            xn, A, B = O.discretize(c, g['disc_x'][i], g['disc_u'][i])
            assert np.max(np.abs(xn - g[f'disc_{tag}_xn'][i])) < 1e-13
            assert np.max(np.abs(A - g[f'disc_{tag}_A'][i])) < 1e-13
            assert np.max(np.abs(B - g[f'disc_{tag}_B'][i])) < 1e-13
    ob = O.OracleBatch(cfg, 16)
    u, s, _, _ = ob.solve(g['b_x0'], g['b_yref'], g['b_yref_e'])
    assert np.array_equal(s, g['b_status_cold']) and np.max(np.abs(u - g['b_u0_cold'])) < 1e-10
    u, s, th, sp = ob.solve(g['b_x0'], g['b_yref'], g['b_yref_e'])
    assert np.max(np.abs(u - g['b_u0_warm'])) < 1e-10
    ob = O.OracleBatch(cfg, 16)
    u, s, _, _ = ob.solve(g['a_x0'], g['a_yref'], g['a_yref_e'])
    assert np.array_equal(s, g['a_status']) and np.max(np.abs(u - g['a_u0'])) < 1e-9
