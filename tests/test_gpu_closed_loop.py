"""GPU parity of the closed-loop kernels against the oracle restatements.

FP64 throughout; tolerances: references / odometry 1e-12 absolute (libm vs CUDA sin/cos/atan2
differ in the last ulp), plant step 1e-12, whole closed loops 1e-7 on states after tens of steps
(differences of ~1e-13 per solve are carried through the closed loop), statuses identical."""
import math

import numpy as np
import pytest

from tests.problems import random_states

pytestmark = pytest.mark.gpu


@pytest.fixture(scope='module')
def O(built):
    from oracle import oracle
    return oracle


def _mk(params, B, plant=None, **solver_over):
    import copy
    from rotors_mpc_controller_b200 import BatchedPositionNMPC, ClosedLoop, Plant
    p = copy.deepcopy(params)
    p['solver'].update(solver_over)
    ctl = BatchedPositionNMPC(p, batch=B)
    return ClosedLoop(ctl, Plant(dt=ctl.dt, **(plant or {}))), ctl, p


def test_reference_kernels(params, O):
    B = 37
    cl, ctl, p = _mk(params, B)
    cfg = O.config_from_params(p)
    rng = np.random.default_rng(0)
    pos, yaw = rng.uniform(-3, 3, (B, 3)), rng.uniform(-math.pi, math.pi, B)
    yr, ye = cl.reference_setpoint(pos, yaw)
    orr, oe = O.setpoint_reference(cfg, pos, yaw)
    assert np.max(np.abs(yr.cpu().numpy() - orr)) < 1e-15
    assert np.max(np.abs(ye.cpu().numpy() - oe)) < 1e-15
    yr, ye = cl.reference_setpoint(pos)                       # yaw = None
    orr, oe = O.setpoint_reference(cfg, pos)
    assert np.array_equal(yr.cpu().numpy(), orr) and np.array_equal(ye.cpu().numpy(), oe)
    t0, A, per = rng.uniform(0, 5, B), rng.uniform(1, 3, B), rng.uniform(2, 4, B)
    yr, ye = cl.reference_figure8(t0, A, per, centre=(0.5, -0.25, 1.5), t_offset=1.3)
    orr, oe = O.figure8_reference(cfg, t0, A, per, centre=(0.5, -0.25, 1.5), t_offset=1.3)
    assert np.max(np.abs(yr.cpu().numpy() - orr)) < 1e-12
    assert np.max(np.abs(ye.cpu().numpy() - oe)) < 1e-12


def test_odometry_to_state(params, O):
    B = 64
    cl, ctl, _ = _mk(params, B)
    rng = np.random.default_rng(1)
    odom = rng.normal(size=(B, 13))
    odom[3, 3:7] = 0.0                                        # zero quaternion -> flagged
    odom[5, 3:7] = [0.0, math.sqrt(0.5), 0.0, math.sqrt(0.5)]  # pitch = +90 deg (gimbal branch)
    x0, bad = cl.odometry_to_state(odom)
    x0, bad = x0.cpu().numpy(), bad.cpu().numpy()
    for b in range(B):
        xr, br = O.odometry_to_state(odom[b])
        assert bool(bad[b]) == bool(br)
        assert np.max(np.abs(x0[b] - xr)) < 1e-12, b


@pytest.mark.parametrize('plant', [dict(substeps=1), dict(substeps=50),
                                   dict(substeps=10, motor_model=1, rotor_drag=8.06428e-5,
                                        mass=0.716),
                                   # the RotorS-faithful "model-mismatch" plant
                                   dict(substeps=10, motor_model=1, rotor_drag=8.06428e-5,
                                        mass=0.716, rolling_moment=1e-6, rotors_yaw_sign=True),
                                   dict(substeps=4, rolling_moment=1e-3)])
def test_plant_step(params, O, plant):
    B = 48
    cl, ctl, p = _mk(params, B, plant=plant)
    cfg = O.config_from_params(p)
    rng = np.random.default_rng(2)
    x = random_states(rng, B, tilt_deg=40.0)
    x[:, 6:10] *= rng.uniform(0.9, 1.1, (B, 1))               # plant quaternion need not be unit
    u = rng.uniform(0.02, 6.0, (B, 4))
    cl.reset(x)
    w0 = rng.uniform(200.0, 800.0, (B, 4))
    cl.rotor_speed.copy_(cl._dev(w0, (B, 4)))
    x0 = cl.plant_step(u).cpu().numpy()
    xn = cl.x.cpu().numpy()
    wn = cl.rotor_speed.cpu().numpy()
    for b in range(B):
        xr, wr = O.plant_step_ext(cfg, x[b], u[b], cfg.dt, rotor=w0[b], **plant)
        assert np.max(np.abs(xn[b] - xr)) < 1e-12
        if plant.get('motor_model'):
            assert np.max(np.abs(wn[b] - wr)) < 1e-9
        xq = xr.copy()
        xq[6:10] /= np.linalg.norm(xq[6:10])
        assert np.max(np.abs(x0[b] - xq)) < 1e-12


def test_rollout_setpoint_matches_oracle_loop(params, O):
    """cfg1 + random set-point steps: 40 closed-loop steps, nominal plant (5 RK4 sub-steps)."""
    B, steps = 24, 40
    cl, ctl, p = _mk(params, B, plant=dict(substeps=5))
    cfg = O.config_from_params(p)
    rng = np.random.default_rng(3)
    x = random_states(rng, B)
    x[0] = 0.0
    x[0, 2], x[0, 6] = 1.0, 1.0
    sp = x[:, 0:3] + rng.uniform(-2, 2, (B, 3))
    sp[0] = [2.0, 0.0, 1.5]
    yaw = rng.uniform(-1, 1, B)
    cl.reset(x)
    cl.run_setpoint(steps, sp, yaw)
    xg, sg = cl.x.cpu().numpy(), cl.stats.cpu().numpy()
    u0, st, th, spd = (t.cpu().numpy() for t in cl.last_outputs())
    ref = O.setpoint_reference(cfg, sp, yaw)
    xo, so, _, out = O.closed_loop(cfg, x, steps, lambda k: ref, lambda k: sp,
                                   plant=dict(substeps=5))
    assert np.max(np.abs(xg - xo)) < 1e-7
    assert np.array_equal(st, out[1]) and np.max(np.abs(th - out[2])) < 1e-7
    assert np.array_equal(sg[:, [2, 3, 4, 7]], so[:, [2, 3, 4, 7]])
    assert np.max(np.abs(sg - so)) < 1e-6
    assert ctl.kernel_launches >= 2 * steps


def test_rollout_figure8_motor_lag_matches_oracle_loop(params, O):
    """cfg4-like figure-8 tracking with the RotorS motor lag + drag plant; bounds become active."""
    B, steps = 16, 30
    plant = dict(substeps=10, motor_model=1, rotor_drag=8.06428e-5, mass=0.716)
    cl, ctl, p = _mk(params, B, plant=plant)
    cfg = O.config_from_params(p)
    rng = np.random.default_rng(4)
    x = random_states(rng, B)
    x[:, 0:3] = [0.0, 0.0, 1.5]
    t0, A, per = rng.uniform(0, 1, B), rng.uniform(1, 3, B), rng.uniform(2, 4, B)
    cl.reset(x)
    w0 = cl.rotor_speed.cpu().numpy().copy()
    cl.run_figure8(steps, t0, A, per)
    xg, sg = cl.x.cpu().numpy(), cl.stats.cpu().numpy()

    def ref_fn(k):
        return O.figure8_reference(cfg, t0, A, per, t_offset=k * cfg.dt)

    def pref_fn(k):
        return ref_fn(k)[0][:, 1, 0:3]

    xo, so, wo, out = O.closed_loop(cfg, x, steps, ref_fn, pref_fn, plant=plant, rotor_init=w0)
    assert np.max(np.abs(xg - xo)) < 1e-6
    assert np.array_equal(sg[:, 2], so[:, 2]) and np.array_equal(sg[:, 7], so[:, 7])
    assert np.max(np.abs(sg[:, [0, 1, 5, 6]] - so[:, [0, 1, 5, 6]])) < 1e-5
    assert np.max(np.abs(cl.rotor_speed.cpu().numpy() - wo)) < 1e-5
    s = cl.summary()
    assert s['problems'] == B and s['nonfinite'] == 0


def test_rollout_large_batch_properties(params):
    """cfg3 at reduced length: 16384 vehicles x 60 steps; every vehicle reaches its set-point,
    no failed solve, no non-finite state; a second identical run is bit-for-bit repeatable."""
    B, steps = 16384, 60
    cl, ctl, _ = _mk(params, B, plant=dict(substeps=5))
    rng = np.random.default_rng(5)
    x = random_states(rng, B)
    sp = x[:, 0:3] + np.concatenate([rng.uniform(-2, 2, (B, 2)), rng.uniform(-0.5, 0.5, (B, 1))], 1)
    cl.reset(x)
    cl.run_setpoint(steps, sp)
    x1, s1 = cl.x.cpu().numpy().copy(), cl.summary()
    assert s1['failed_solves'] == 0 and s1['nonfinite'] == 0
    assert s1['max_final_pos_err'] < 0.35 and s1['mean_final_pos_err'] < 0.1
    cl.reset(x)
    cl.run_setpoint(steps, sp)
    assert np.array_equal(cl.x.cpu().numpy(), x1)
