"""GPU parity tests: the CUDA path (through the C ABI) against the CPU oracle.

Tolerances (FP64): u0 / trajectories within 1e-8 absolute of the oracle (north_star asks 1e-6
relative against acados; both sides here converge the strictly convex QP to complementarity
<= 1e-14, so they agree far tighter); status codes identical.
"""
import numpy as np
import pytest

from tests.problems import aggressive, random_states, setpoint_refs

pytestmark = pytest.mark.gpu

TOL_U = 1e-8


@pytest.fixture(scope='module')
def O(built):
    from oracle import oracle
    return oracle


def _mk(params, B, **solver_over):
    import copy
    from rotors_mpc_controller_b200 import BatchedPositionNMPC
    p = copy.deepcopy(params)
    p['solver'].update(solver_over)
    return BatchedPositionNMPC(p, batch=B), p


def test_discretize_matches_oracle(params, O):
    ctl, _ = _mk(params, 8)
    cfg = O.config_from_params(params)
    rng = np.random.default_rng(0)
    x = rng.normal(size=(32, 13))
    u = rng.uniform(0, 6, (32, 4))
    xn, A, B = ctl.discretize(x, u)
    for i in range(32):
        xr, Ar, Br = O.discretize(cfg, x[i], u[i])
        assert np.max(np.abs(xn[i] - xr)) < 1e-12
        assert np.max(np.abs(A[i] - Ar)) < 1e-12
        assert np.max(np.abs(B[i] - Br)) < 1e-12


@pytest.mark.parametrize('integ,ns,steps', [('rk4', 4, 1), ('euler', 1, 3), ('rk4', 4, 2)])
def test_discretize_other_tableaus(params, O, integ, ns, steps):
    ctl, _ = _mk(params, 4, integrator=integ, integrator_steps=steps)
    cfg = O.config_from_params(params, integ_ns=ns, integ_steps=steps)
    rng = np.random.default_rng(1)
    x = rng.normal(size=(8, 13))
    u = rng.uniform(0, 6, (8, 4))
    xn, A, B = ctl.discretize(x, u)
    for i in range(8):
        xr, Ar, Br = O.discretize(cfg, x[i], u[i])
        assert np.max(np.abs(xn[i] - xr)) < 1e-12
        assert np.max(np.abs(A[i] - Ar)) < 1e-12
        assert np.max(np.abs(B[i] - Br)) < 1e-12


def test_rti_cold_start_batch(params, O):
    """cfg2 distribution, cold start (u = 0 linearisation, controller.py:425-431)."""
    B, N = 512, 20
    ctl, _ = _mk(params, B)
    rng = np.random.default_rng(0)
    x0 = random_states(rng, B)
    yref, yref_e = setpoint_refs(rng, x0, N, ctl.hover_thrust)
    u0, st = ctl.solve(x0, yref, yref_e)
    ob = O.OracleBatch(O.config_from_params(params), B)
    ur, sr, thr, spr = ob.solve(x0, yref, yref_e)
    assert np.array_equal(st, sr)
    assert np.all(st == 0)
    assert np.max(np.abs(u0 - ur)) < TOL_U
    xt, ut, valid = ctl.trajectories()
    assert valid.all()
    assert np.max(np.abs(xt - ob.xtraj)) < TOL_U
    assert np.max(np.abs(ut - ob.utraj)) < TOL_U
    th, sp = ctl.command()
    assert np.max(np.abs(th - thr)) < TOL_U
    assert np.max(np.abs(sp - spr)) < 1e-5  # speeds are ~1e2..1e3 rad/s (sqrt amplifies)
    assert np.array_equal(ctl.stats()['qp_iters'], ob.qp_iters)


def test_rti_warm_start_closed_loop(params, O):
    """Several RTI steps with warm start (un-shifted, controller.py:419-424) and plant roll."""
    B, N = 128, 20
    ctl, _ = _mk(params, B)
    cfg = O.config_from_params(params)
    rng = np.random.default_rng(3)
    x0 = random_states(rng, B)
    yref, yref_e = setpoint_refs(rng, x0, N, ctl.hover_thrust)
    ob = O.OracleBatch(cfg, B)
    for step in range(6):
        u0, st = ctl.solve(x0, yref, yref_e)
        ur, sr, thr, _ = ob.solve(x0, yref, yref_e)
        assert np.array_equal(st, sr), step
        assert np.max(np.abs(u0 - ur)) < TOL_U, step
        th, _ = ctl.command()
        x0 = np.stack([O.plant_step(cfg, x0[b], thr[b], 0.05, 2) for b in range(B)])
    xt, ut, valid = ctl.trajectories()
    assert np.max(np.abs(xt - ob.xtraj)) < TOL_U


def test_active_bounds_batch(params, O):
    """cfg4-like aggressive references: thrust box active on many stages."""
    B, N = 512, 20
    ctl, _ = _mk(params, B)
    rng = np.random.default_rng(4)
    x0, yref, yref_e = aggressive(rng, B, N, ctl.hover_thrust)
    ob = O.OracleBatch(O.config_from_params(params), B)
    for step in range(2):
        u0, st = ctl.solve(x0, yref, yref_e)
        ur, sr, _, _ = ob.solve(x0, yref, yref_e)
        assert np.array_equal(st, sr)
        assert np.max(np.abs(u0 - ur)) < TOL_U
        lo, hi = ctl.input_bounds
        xt, ut, _ = ctl.trajectories()
        nact = np.sum((ut < lo + 1e-7) | (ut > hi - 1e-7))
        assert nact > 100  # the box really is active in this batch
        assert np.all(ut >= lo - 1e-9) and np.all(ut <= hi + 1e-9)
        assert np.max(np.abs(ut - ob.utraj)) < TOL_U


@pytest.mark.parametrize('N,steps', [(1, 2), (2, 2), (3, 2), (4, 2), (5, 2), (7, 2), (19, 2), (21, 2),
                                     (12, 3), (20, 1), (32, 1), (33, 1)])
def test_horizons_and_substeps_with_active_bounds(params, O, N, steps):
    """Round-2 code paths that depend on the horizon and on the integrator sub-steps:
    the checkpoints of the factor sweep sit at stages j N / 5 (several coincide for N < 5, the
    restart logic must skip them), and the linearisation keeps the nominal points of a stage in
    the sweep tile when they fit (N <= 20 at 2 sub-steps, N <= 32 at one; above: column-per-lane
    integration).  Aggressive problems, so the active-set iterations do re-solve."""
    B = 96
    ctl, p = _mk(params, B, horizon_steps=N, integrator_steps=steps)
    rng = np.random.default_rng(100 * N + steps)
    x0, yref, yref_e = aggressive(rng, B, N, ctl.hover_thrust)
    ob = O.OracleBatch(O.config_from_params(p, integ_steps=steps), B)
    for step in range(2):   # cold, then warm from the stored trajectories
        u0, st = ctl.solve(x0, yref, yref_e)
        ur, sr, _, _ = ob.solve(x0, yref, yref_e)
        assert np.array_equal(st, sr)
        assert np.max(np.abs(u0 - ur)) < TOL_U
        assert np.array_equal(ctl.stats()['qp_iters'], ob.qp_iters)
        xt, ut, _ = ctl.trajectories()
        assert np.max(np.abs(xt - ob.xtraj)) < TOL_U and np.max(np.abs(ut - ob.utraj)) < TOL_U
    assert ob.qp_iters.max() >= 2   # some problem needed more than one factorisation
    ctl.close()


def test_explicit_thrust_limits_override(params, O):
    """the YAML's [4,20] N box (dead config in the reference) as an opt-in."""
    B, N = 64, 20
    ctl, p = _mk(params, B, use_thrust_limits=True)
    assert np.allclose(ctl.input_bounds[0], 4.0) and np.allclose(ctl.input_bounds[1], 20.0)
    rng = np.random.default_rng(5)
    x0 = random_states(rng, B)
    yref, yref_e = setpoint_refs(rng, x0, N, ctl.hover_thrust)
    u0, st = ctl.solve(x0, yref, yref_e)
    ob = O.OracleBatch(O.config_from_params(params, thrust_box=(4.0, 20.0)), B)
    ur, sr, _, _ = ob.solve(x0, yref, yref_e)
    assert np.array_equal(st, sr)
    assert np.max(np.abs(u0 - ur)) < TOL_U
    assert np.all(u0 >= 4.0 - 1e-9)


def test_full_sqp(params, O):
    """cfg2: SQP to convergence (extension, max_sqp_iter > 1)."""
    B, N = 128, 20
    ctl, _ = _mk(params, B, max_sqp_iter=50, sqp_tol=1e-8)
    rng = np.random.default_rng(6)
    x0 = random_states(rng, B)
    yref, yref_e = setpoint_refs(rng, x0, N, ctl.hover_thrust)
    u0, st = ctl.solve(x0, yref, yref_e)
    ob = O.OracleBatch(O.config_from_params(params, max_sqp_iter=50, sqp_tol=1e-8), B)
    ur, sr, _, _ = ob.solve(x0, yref, yref_e)
    assert np.array_equal(st, sr)
    ok = st == 0
    assert ok.mean() > 0.95
    assert np.max(np.abs(u0[ok] - ur[ok])) < 1e-6
    # at an SQP fixed point the dynamics defects vanish
    xt, ut, _ = ctl.trajectories()
    cfg = O.config_from_params(params)
    b = int(np.argmax(ok))
    for k in range(N):
        xn, _, _ = O.discretize(cfg, xt[b, k], ut[b, k])
        assert np.max(np.abs(xn - xt[b, k + 1])) < 1e-7


def test_status_and_failsafe(params, O):
    """NaN state -> status 4, u0 = 0, warm start invalidated, last-valid command republished
    (controller.py:448-450, mpc_controller_node:124-129); the other problems are unaffected."""
    B, N = 32, 20
    ctl, _ = _mk(params, B)
    rng = np.random.default_rng(7)
    x0 = random_states(rng, B)
    yref, yref_e = setpoint_refs(rng, x0, N, ctl.hover_thrust)
    ob = O.OracleBatch(O.config_from_params(params), B)
    ctl.solve(x0, yref, yref_e)
    ob.solve(x0, yref, yref_e)
    th_prev = ctl.command()[0].copy()
    bad = x0.copy()
    bad[3, 4] = np.nan
    bad[9, 0] = np.inf
    u0, st = ctl.solve(bad, yref, yref_e)
    ur, sr, thr, spr = ob.solve(bad, yref, yref_e)
    assert np.array_equal(st, sr)
    assert st[3] == 4 and st[9] == 4 and np.sum(st != 0) == 2
    assert np.all(u0[3] == 0) and np.all(u0[9] == 0)
    th, sp = ctl.command()
    assert np.allclose(th[3], th_prev[3]) and np.allclose(th[9], th_prev[9])
    assert np.max(np.abs(th - thr)) < TOL_U
    _, _, valid = ctl.trajectories()
    assert not valid[3] and not valid[9] and valid.sum() == B - 2
    # next solve cold-starts the failed problems and succeeds again
    u0, st = ctl.solve(x0, yref, yref_e)
    ur, sr, _, _ = ob.solve(x0, yref, yref_e)
    assert np.all(st == 0) and np.array_equal(st, sr)
    assert np.max(np.abs(u0 - ur)) < TOL_U


def test_single_vehicle_dropin_closed_loop(params, O):
    """cfg1: PositionNMPC API, hover at (0,0,1) stepping to (2,0,1.5), nominal plant."""
    from rotors_mpc_controller_b200 import PositionNMPC, ReferenceGenerator
    ctl = PositionNMPC(params)
    orc = O.OraclePositionNMPC(params)
    cfg = O.config_from_params(params)
    ref = ReferenceGenerator(params['reference'])
    ref.update_hover_thrust(ctl.hover_thrust)
    ref.set_target(position=[2.0, 0.0, 1.5], yaw=0.0)
    x = np.array([0, 0, 1, 0, 0, 0, 1, 0, 0, 0, 0, 0, 0.0])
    lo, hi = ctl.input_bounds
    for k in range(120):
        state = dict(position=x[:3], velocity=x[3:6], quaternion=x[6:10], body_rates=x[10:])
        horizon = ref.build_horizon(ctl.horizon, ctl.dt)
        u0, st = ctl.solve(state, horizon)
        ur, sr = orc.solve(state, horizon)
        assert st == sr == 0
        assert np.max(np.abs(u0 - ur)) < TOL_U
        assert np.all(u0 >= lo - 1e-9) and np.all(u0 <= hi + 1e-9)
        thrust, speed = ctl.command()
        x = O.plant_step(cfg, x, thrust, ctl.dt, 5)
    assert np.linalg.norm(x[:3] - [2.0, 0.0, 1.5]) < 2e-2
    with pytest.raises(ValueError):
        ctl.solve(dict(position=x[:3], velocity=x[3:6], quaternion=np.zeros(4),
                       body_rates=x[10:]), horizon)


def test_device_tensor_path(params, O):
    import torch
    B, N = 256, 20
    ctl, _ = _mk(params, B)
    rng = np.random.default_rng(8)
    x0 = random_states(rng, B)
    yref, yref_e = setpoint_refs(rng, x0, N, ctl.hover_thrust)
    dev = torch.device('cuda', 0)
    u0, st = ctl.solve(torch.as_tensor(x0, device=dev), torch.as_tensor(yref, device=dev),
                       torch.as_tensor(yref_e, device=dev))
    torch.cuda.synchronize()
    ob = O.OracleBatch(O.config_from_params(params), B)
    ur, sr, _, _ = ob.solve(x0, yref, yref_e)
    assert np.array_equal(st.cpu().numpy(), sr)
    assert np.max(np.abs(u0.cpu().numpy() - ur)) < TOL_U


def test_long_horizon(params, O):
    """cfg5 (reduced batch): N = 600 runs through the same kernel with a global workspace."""
    B, N = 8, 600
    ctl, p = _mk(params, B, horizon_steps=N)
    rng = np.random.default_rng(9)
    x0 = random_states(rng, B)
    yref, yref_e = setpoint_refs(rng, x0, N, ctl.hover_thrust)
    u0, st = ctl.solve(x0, yref, yref_e)
    ob = O.OracleBatch(O.config_from_params(p), B)
    ur, sr, _, _ = ob.solve(x0, yref, yref_e)
    assert np.array_equal(st, sr)
    assert np.max(np.abs(u0 - ur)) < 1e-7


def test_committed_goldens(params):
    """CUDA path against tests/golden/nmpc_golden.npz (made by tests/golden/make_golden.py)."""
    from pathlib import Path
    g = np.load(Path(__file__).parent / 'golden' / 'nmpc_golden.npz')
    ctl, _ = _mk(params, 6)
    xn, A, B = ctl.discretize(g['disc_x'], g['disc_u'])
    assert np.max(np.abs(xn - g['disc_mid_xn'])) < 1e-12
    assert np.max(np.abs(A - g['disc_mid_A'])) < 1e-12 and np.max(np.abs(B - g['disc_mid_B'])) < 1e-12
    ctl, _ = _mk(params, 16)
    u, s = ctl.solve(g['b_x0'], g['b_yref'], g['b_yref_e'])
    assert np.array_equal(s, g['b_status_cold']) and np.max(np.abs(u - g['b_u0_cold'])) < TOL_U
    assert np.max(np.abs(ctl.trajectories()[0] - g['b_xtraj_cold'])) < TOL_U
    u, s = ctl.solve(g['b_x0'], g['b_yref'], g['b_yref_e'])
    assert np.max(np.abs(u - g['b_u0_warm'])) < TOL_U
    th, sp = ctl.command()
    assert np.max(np.abs(th - g['b_thrust_warm'])) < TOL_U
    assert np.max(np.abs(sp - g['b_speed_warm'])) < 1e-5
    ctl, _ = _mk(params, 16)
    u, s = ctl.solve(g['a_x0'], g['a_yref'], g['a_yref_e'])
    assert np.array_equal(s, g['a_status']) and np.max(np.abs(u - g['a_u0'])) < TOL_U
    assert np.max(np.abs(ctl.trajectories()[1] - g['a_utraj'])) < TOL_U
    assert np.array_equal(ctl.stats()['qp_iters'], g['a_iters'])
    ctl, _ = _mk(params, 8, max_sqp_iter=50, sqp_tol=1e-8)
    u, s = ctl.solve(g['s_x0'], g['s_yref'], g['s_yref_e'])
    assert np.array_equal(s, g['s_status']) and np.max(np.abs(u - g['s_u0'])) < 1e-6
    # cfg1 closed loop replayed open-loop on the recorded states (warm start carried along)
    from rotors_mpc_controller_b200 import BatchedPositionNMPC
    one = BatchedPositionNMPC(params, batch=1)
    for k in range(g['cfg1_x'].shape[0]):
        u, s = one.solve(g['cfg1_x'][k][None], g['cfg1_yref'][None], g['cfg1_yref_e'][None])
        assert s[0] == 0 and np.max(np.abs(u[0] - g['cfg1_u0'][k])) < TOL_U


def test_large_batch_properties(params):
    """Size-independent properties at BASELINE scale (65536 problems, no oracle): all statuses
    0, thrusts inside the box, solving twice from the same warm start is deterministic, a
    solve from the converged SQP point leaves it unchanged (idempotence)."""
    B, N = 65536, 20
    ctl, _ = _mk(params, B)
    rng = np.random.default_rng(11)
    x0, yref, yref_e = aggressive(rng, B, N, ctl.hover_thrust)
    u0, st = ctl.solve(x0, yref, yref_e)
    u0 = u0.copy()
    assert np.all(st == 0)
    lo, hi = ctl.input_bounds
    assert np.all(u0 >= lo - 1e-9) and np.all(u0 <= hi + 1e-9)
    it = ctl.stats()['qp_iters']
    assert it.max() <= 40 and it.min() >= 1
    ctl.invalidate()
    u1, st1 = ctl.solve(x0, yref, yref_e)
    assert np.array_equal(u0, u1)          # bit-for-bit repeatable
    sq, _ = _mk(params, 4096, max_sqp_iter=80, sqp_tol=1e-9)
    xs = random_states(rng, 4096)
    ys, es = setpoint_refs(rng, xs, N, ctl.hover_thrust)
    ua, sa = sq.solve(xs, ys, es)
    ua = ua.copy()
    ub, sb = sq.solve(xs, ys, es)          # warm start = fixed point -> no movement
    ok = (sa == 0) & (sb == 0)
    assert ok.mean() > 0.9
    assert np.max(np.abs(ua[ok] - ub[ok])) < 1e-7


@pytest.mark.parametrize('kind', ['cfg2', 'aggressive'])
def test_fp32_variant(params, O, kind):
    """FP32 compute variant (north_star: u0 within 1e-3 relative of the FP64 reference; measured
    against the FP64 oracle with the thrust scale max(|u0|, hover) as denominator), identical
    status codes.  I/O stays FP64."""
    B, N = 512, 20
    ctl, p = _mk(params, B, precision='f32')
    rng = np.random.default_rng(21)
    if kind == 'cfg2':
        x0 = random_states(rng, B)
        yref, yref_e = setpoint_refs(rng, x0, N, ctl.hover_thrust)
    else:
        x0, yref, yref_e = aggressive(rng, B, N, ctl.hover_thrust)
    u0, st = ctl.solve(x0, yref, yref_e)
    ob = O.OracleBatch(O.config_from_params(params), B)
    ur, sr, _, _ = ob.solve(x0, yref, yref_e)
    assert np.array_equal(st, sr)
    rel = np.abs(u0 - ur) / np.maximum(np.abs(ur), ctl.hover_thrust)
    assert rel.max() < 1e-3, rel.max()
    # warm-started second solve stays inside the tolerance as well
    u1, st1 = ctl.solve(x0, yref, yref_e)
    ur1, sr1, _, _ = ob.solve(x0, yref, yref_e)
    assert np.array_equal(st1, sr1)
    rel1 = np.abs(u1 - ur1) / np.maximum(np.abs(ur1), ctl.hover_thrust)
    assert rel1.max() < 1e-3, rel1.max()
