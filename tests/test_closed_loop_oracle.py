"""CPU checks of the oracle restatements used to verify the closed-loop kernels (f1-f3)."""
import math

import numpy as np

from tests.problems import random_states


def _cfg(params):
    from oracle import oracle as O
    return O, O.config_from_params(params)


def test_setpoint_reference_matches_reference_generator(built, params):
    """oracle.setpoint_reference == ReferenceGenerator.build_horizon + stack_reference
    (reference.py:75-91, controller.py:433-445)."""
    O, cfg = _cfg(params)
    from rotors_mpc_controller_b200 import ReferenceGenerator, stack_reference
    rng = np.random.default_rng(0)
    pos, yaw = rng.uniform(-3, 3, (5, 3)), rng.uniform(-math.pi, math.pi, 5)
    yref, yref_e = O.setpoint_reference(cfg, pos, yaw)
    hover = cfg.mass * cfg.gravity / 4.0
    for b in range(5):
        g = ReferenceGenerator({})
        g.set_target(pos[b], yaw=yaw[b])
        g.update_hover_thrust(hover)
        yr, ye = stack_reference(g.build_horizon(cfg.N, cfg.dt), cfg.N)
        assert np.array_equal(yr, yref[b]) and np.array_equal(ye, yref_e[b])


def test_figure8_reference_matches_batch_reference(built, params):
    O, cfg = _cfg(params)
    from rotors_mpc_controller_b200 import BatchReference
    rng = np.random.default_rng(1)
    t0, A, per = rng.uniform(0, 5, 6), rng.uniform(1, 3, 6), rng.uniform(2, 4, 6)
    yref, yref_e = O.figure8_reference(cfg, t0, A, per, t_offset=0.35)
    br = BatchReference(cfg.N, cfg.dt, cfg.mass * cfg.gravity / 4.0)
    yr, ye = br.figure8(t0 + 0.35, A, per)
    assert np.max(np.abs(yr - yref)) < 1e-12 and np.max(np.abs(ye - yref_e)) < 1e-12
    # velocity rows are the time derivative of the position rows
    y2, _ = O.figure8_reference(cfg, t0, A, per, t_offset=0.35 + 1e-6)
    fd = (y2[:, :, 0:3] - yref[:, :, 0:3]) / 1e-6
    assert np.max(np.abs(fd - yref[:, :, 3:6])) < 1e-4


def test_odometry_to_state_is_rotation_by_quaternion(built, params):
    """The node goes quaternion -> euler -> R(rpy); away from gimbal lock that is R(q)."""
    O, _ = _cfg(params)
    rng = np.random.default_rng(2)
    for _ in range(20):
        q = rng.normal(size=4)
        q /= np.linalg.norm(q)
        w, x, y, z = q
        R = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
                      [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
                      [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]])
        vb, p, om = rng.normal(size=3), rng.normal(size=3), rng.normal(size=3)
        scale = rng.uniform(0.5, 2.0)           # un-normalised message quaternion
        odom = np.concatenate([p, scale * np.array([x, y, z, w]), vb, om])
        x0, bad = O.odometry_to_state(odom)
        assert not bad
        assert np.allclose(x0[0:3], p) and np.allclose(x0[10:13], om)
        assert np.allclose(x0[3:6], R @ vb, atol=1e-12)
        assert np.allclose(x0[6:10], q, atol=1e-15)
    x0, bad = O.odometry_to_state(np.zeros(13))
    assert bad


def test_plant_ext_reduces_to_nominal_and_lags(built, params):
    O, cfg = _cfg(params)
    rng = np.random.default_rng(3)
    x = random_states(rng, 1)[0]
    u = rng.uniform(0.5, 3.0, 4)
    xa = O.plant_step(cfg, x, u, cfg.dt, 5)
    xb, _ = O.plant_step_ext(cfg, x, u, cfg.dt, 5)
    assert np.max(np.abs(xa - xb)) < 1e-14
    # motor lag: rotor speed approaches the command with exp(-h/tau) per sub-step
    w_cmd = np.sqrt(u / cfg.kf)
    w0 = 0.5 * w_cmd
    _, w1 = O.plant_step_ext(cfg, x, u, cfg.dt, 50, motor_model=1, rotor=w0)
    expect = w_cmd + (w0 - w_cmd) * math.exp(-cfg.dt / 0.0125)
    assert np.allclose(w1, expect, rtol=1e-12)
    _, w2 = O.plant_step_ext(cfg, x, u, cfg.dt, 50, motor_model=1, rotor=2.0 * w_cmd)
    assert np.allclose(w2, w_cmd + w_cmd * math.exp(-cfg.dt / 0.025), rtol=1e-12)
    # rotor drag only acts against the velocity component perpendicular to the rotor axis
    xs = x.copy()
    xs[6:10] = [1.0, 0.0, 0.0, 0.0]
    xs[3:6] = [1.0, 0.0, 0.7]
    xs[10:13] = 0.0
    hov = np.full(4, cfg.mass * cfg.gravity / 4.0)
    xn0, _ = O.plant_step_ext(cfg, xs, hov, 0.01, 1)
    xn1, _ = O.plant_step_ext(cfg, xs, hov, 0.01, 1, rotor_drag=8.06428e-5)
    assert xn1[3] < xn0[3] and abs(xn1[5] - xn0[5]) < 1e-12


def test_closed_loop_oracle_converges(built, params):
    """cfg1-like: hover at (0,0,1) -> set-point (2,0,1.5); statistics are self-consistent."""
    O, cfg = _cfg(params)
    B = 3
    x = np.zeros((B, 13))
    x[:, 2], x[:, 6] = 1.0, 1.0
    sp = np.tile([2.0, 0.0, 1.5], (B, 1))
    ref = O.setpoint_reference(cfg, sp)
    xf, stats, _, out = O.closed_loop(cfg, x, 120, lambda k: ref, lambda k: sp,
                                      plant=dict(substeps=5))
    assert np.all(stats[:, 2] == 0) and np.all(stats[:, 7] == 0)
    assert np.all(stats[:, 1] < 0.05) and np.all(stats[:, 5] <= 2.07)
    assert np.allclose(np.linalg.norm(xf[:, 0:3] - sp, axis=1), stats[:, 1])


def test_plant_rolling_moment_and_rotors_yaw_sign(built, params):
    """f1 model-mismatch terms against an independent numpy restatement of
    gazebo_motor_model.cpp:458-474: the body-rate derivative gains J^-1 R' (-sum|w| mu v_perp) and
    the yaw drag torque flips sign; everything else of the step is unchanged."""
    O, cfg = _cfg(params)
    rng = np.random.default_rng(5)
    J = np.array(cfg.J[:])
    for _ in range(10):
        x = random_states(rng, 1, tilt_deg=40.0, v_sigma=2.0)[0]
        u = rng.uniform(0.5, 5.0, 4)
        h = 1e-7                                   # one RK4 step of 0.1 us: x+ = x + h f to O(h^2)
        base, _ = O.plant_step_ext(cfg, x, u, h)
        mu = 1e-3
        roll, _ = O.plant_step_ext(cfg, x, u, h, rolling_moment=mu)
        yaw, _ = O.plant_step_ext(cfg, x, u, h, rotors_yaw_sign=True)
        w, qx, qy, qz = x[6:10]
        R = np.array([[1 - 2 * (qy * qy + qz * qz), 2 * (qx * qy - w * qz), 2 * (qx * qz + w * qy)],
                      [2 * (qx * qy + w * qz), 1 - 2 * (qx * qx + qz * qz), 2 * (qy * qz - w * qx)],
                      [2 * (qx * qz - w * qy), 2 * (qy * qz + w * qx), 1 - 2 * (qx * qx + qy * qy)]])
        v = x[3:6]
        vperp = v - (v @ R[:, 2]) * R[:, 2]
        saw = np.sum(np.sqrt(u / cfg.kf))
        dw = (R.T @ (-saw * mu * vperp)) / J
        assert np.max(np.abs((roll - base)[10:13] / h - dw)) < 1e-5 * max(1.0, np.abs(dw).max())
        assert np.max(np.abs((roll - base)[:10])) < 1e-10          # second order in h
        tz = u @ np.array(cfg.rotor_z[:])
        assert abs((yaw - base)[12] / h - (-2.0 * tz / J[2])) < 1e-5 * max(1.0, abs(tz / J[2]))
        assert np.max(np.abs((yaw - base)[:12])) < 1e-10
