"""ctypes front-end of the CPU ORACLE (test infrastructure, NOT the product).

PARITY UNPINNED: the reference has no golden vectors and acados cannot be installed here;
see ``oracle/nmpc_oracle.h``.  Only ``tests/``, ``__graft_entry__.smoke()`` and
``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs import this module.

``OraclePositionNMPC`` mirrors ``PositionNMPC`` of
``/root/reference/src/rotors_mpc_controller/controller.py`` (constructor dict :55-60,
``solve(state, reference)`` :385-463) so one test-suite can drive oracle and product.
"""
from __future__ import annotations

import ctypes as C
import math
import os
import subprocess
from pathlib import Path
from typing import Dict, Tuple

import numpy as np

_HERE = Path(__file__).resolve().parent
NX, NU, NZ = 13, 4, 17


class OrcConfig(C.Structure):
    _fields_ = [
        ('N', C.c_int), ('dt', C.c_double), ('W', C.c_double * 17), ('We', C.c_double * 13),
        ('lm', C.c_double), ('qp_iter_max', C.c_int), ('mass', C.c_double),
        ('J', C.c_double * 3), ('gravity', C.c_double), ('rotor_x', C.c_double * 4),
        ('rotor_y', C.c_double * 4), ('rotor_z', C.c_double * 4), ('lbu', C.c_double * 4),
        ('ubu', C.c_double * 4), ('kf', C.c_double), ('w_min', C.c_double),
        ('w_max', C.c_double), ('integ_ns', C.c_int), ('integ_steps', C.c_int),
        ('cost_scale_dt', C.c_int), ('lm_scaled_by_dt', C.c_int), ('max_sqp_iter', C.c_int),
        ('sqp_tol', C.c_double), ('ipm_tol', C.c_double), ('nan_status', C.c_int),
        ('ipm_warm', C.c_int), ('as_shortcut', C.c_int), ('keep_state_box', C.c_int),
        ('state_box', C.c_double),
    ]


class OrcQP(C.Structure):
    _fields_ = ([('N', C.c_int)] + [(n, C.POINTER(C.c_double))
                                    for n in ('A', 'B', 'b', 'Q', 'R', 'q', 'r', 'lb', 'ub',
                                              'lbx', 'ubx')] + [('xbox', C.c_int)])


def build(force: bool = False) -> Path:
    """Compile oracle/liboracle.so with the committed Makefile (gcc only)."""
    so = _HERE / 'liboracle.so'
    srcs = [_HERE / n for n in ('nmpc_oracle.c', 'nmpc_oracle_qp.c', 'nmpc_oracle_solve.c',
                                'nmpc_oracle.h', 'Makefile')]
    if force or not so.exists() or any(s.stat().st_mtime > so.stat().st_mtime for s in srcs):
        try:
            subprocess.run(['make', '-C', str(_HERE), '-B', 'liboracle.so'], check=True,
                           capture_output=True, text=True)
        except (subprocess.CalledProcessError, FileNotFoundError):
            if not so.exists():
                raise
    return so


_lib = None
_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)
_bp = C.POINTER(C.c_ubyte)


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(str(build()))
        _lib.orc_qp_alloc.restype = C.POINTER(OrcQP)
        _lib.orc_qp_alloc.argtypes = [C.c_int]
        _lib.orc_qp_free.argtypes = [C.POINTER(OrcQP)]
        _lib.orc_qp_solve.restype = C.c_int
        _lib.orc_solve.restype = C.c_int
    return _lib


def _p(a, t=_dp):
    return a.ctypes.data_as(t)


def config_from_params(params: Dict[str, Dict[str, object]], **flags) -> OrcConfig:
    """params dict -> solver constants, following controller.py:63-155."""
    s, v = params['solver'], params['vehicle']
    w = params.get('world', {})
    c = OrcConfig()
    c.N = int(s['horizon_steps'])
    c.dt = float(s['dt'])
    W = np.concatenate([np.asarray(s.get('position_weight', [10.0, 10.0, 8.0]), float),
                        np.asarray(s.get('velocity_weight', [1.0, 1.0, 0.2]), float),
                        np.asarray(s.get('quaternion_weight', [3.2] * 4), float),
                        np.asarray(s.get('rate_weight', [1.4, 1.4, 0.4]), float),
                        np.asarray(s.get('control_weight', [1.75] * 4), float)])
    We = np.asarray(s.get('terminal_weight', [5.0, 5.0, 3.0, 2.0, 2.0, 2.0, 12.0, 12.0, 12.0,
                                              18.5, 2.0, 2.0, 1.8]), float)
    c.W[:] = W.tolist()
    c.We[:] = We.tolist()
    c.lm = float(s.get('regularization', 7.0e-3))
    c.qp_iter_max = int(s.get('iter_max', 600))
    c.mass = float(v['mass'])
    inertia = np.asarray(v.get('inertia', [0.007, 0, 0, 0, 0.007, 0, 0, 0, 0.012]),
                         float).reshape(3, 3)
    c.J[:] = [inertia[0, 0], inertia[1, 1], inertia[2, 2]]
    c.gravity = float(w.get('gravity', 9.81))
    L = float(v.get('arm_length', 0.17))
    kf = float(v.get('rotor_force_constant', 8.54858e-6))
    km = float(v.get('rotor_moment_constant', 0.016))
    wmin = float(v.get('motor_min_speed', 0.0))
    wmax = float(v.get('motor_max_speed', 2000.0))
    if str(v.get('rotor_configuration', '+')).strip() != '+':
        raise ValueError('only "+" lay-out')
    c.rotor_x[:] = [L, 0.0, -L, 0.0]
    c.rotor_y[:] = [0.0, L, 0.0, -L]
    c.rotor_z[:] = [-km, km, -km, km]
    lo, hi = max(0.0, kf * wmin ** 2), kf * wmax ** 2
    if 'thrust_box' in flags and flags['thrust_box'] is not None:
        lo, hi = flags['thrust_box']
    c.lbu[:] = [lo] * 4
    c.ubu[:] = [hi] * 4
    c.kf, c.w_min, c.w_max = kf, wmin, wmax
    c.integ_ns = int(flags.get('integ_ns', 2))
    c.integ_steps = int(flags.get('integ_steps', 2))
    c.cost_scale_dt = int(flags.get('cost_scale_dt', 1))
    c.lm_scaled_by_dt = int(flags.get('lm_scaled_by_dt', 1))
    c.max_sqp_iter = int(flags.get('max_sqp_iter', 1))
    c.sqp_tol = float(flags.get('sqp_tol', 1e-8))
    c.ipm_tol = float(flags.get('ipm_tol', 1e-14))
    c.nan_status = int(flags.get('nan_status', 4))
    c.ipm_warm = int(bool(flags.get('ipm_warm', s.get('ipm_warm_start', True))))
    c.as_shortcut = int(bool(flags.get('as_shortcut', s.get('active_set_shortcut', True))))
    c.keep_state_box = int(flags.get('keep_state_box', 0))   # controller.py:257-261 (oracle-only)
    c.state_box = float(flags.get('state_box', 1.0e6))
    return c


def f(cfg, x, u):
    out = np.zeros(NX)
    lib().orc_f(C.byref(cfg), _p(np.ascontiguousarray(x, float)),
                _p(np.ascontiguousarray(u, float)), _p(out))
    return out


def jac(cfg, x, u):
    Fx, Fu = np.zeros((NX, NX)), np.zeros((NX, NU))
    lib().orc_jac(C.byref(cfg), _p(np.ascontiguousarray(x, float)),
                  _p(np.ascontiguousarray(u, float)), _p(Fx), _p(Fu))
    return Fx, Fu


def discretize(cfg, x, u):
    xn, A, B = np.zeros(NX), np.zeros((NX, NX)), np.zeros((NX, NU))
    lib().orc_discretize(C.byref(cfg), _p(np.ascontiguousarray(x, float)),
                         _p(np.ascontiguousarray(u, float)), _p(xn), _p(A), _p(B))
    return xn, A, B


def plant_step(cfg, x, u, dt, substeps=1):
    xn = np.zeros(NX)
    lib().orc_plant_step(C.byref(cfg), _p(np.ascontiguousarray(x, float)),
                         _p(np.ascontiguousarray(u, float)), C.c_double(dt), C.c_int(substeps),
                         _p(xn))
    return xn


def build_qp(cfg, xtraj, utraj, yref, yref_e) -> Dict[str, np.ndarray]:
    N = cfg.N
    qp = lib().orc_qp_alloc(N)
    xtraj, utraj = np.ascontiguousarray(xtraj, float), np.ascontiguousarray(utraj, float)
    yref, yref_e = np.ascontiguousarray(yref, float), np.ascontiguousarray(yref_e, float)
    lib().orc_build_qp(C.byref(cfg), _p(xtraj), _p(utraj), _p(yref), _p(yref_e), qp)
    shp = dict(A=(N, NX, NX), B=(N, NX, NU), b=(N, NX), Q=(N + 1, NX), R=(N, NU), q=(N + 1, NX),
               r=(N, NU), lb=(N, NU), ub=(N, NU))
    out = {k: np.ctypeslib.as_array(getattr(qp.contents, k), shape=s).copy()
           for k, s in shp.items()}
    if qp.contents.xbox:
        for k in ('lbx', 'ubx'):
            out[k] = np.ctypeslib.as_array(getattr(qp.contents, k), shape=(N + 1, NX)).copy()
    lib().orc_qp_free(qp)
    return out


def qp_solve(qpd: Dict[str, np.ndarray], iter_max=600, tol=1e-14):
    N = qpd['A'].shape[0]
    qp = OrcQP()
    qp.N = N
    keep = {k: np.ascontiguousarray(qpd[k], float) for k in ('A', 'B', 'b', 'Q', 'R', 'q', 'r',
                                                              'lb', 'ub')}
    for k, a in keep.items():
        setattr(qp, k, _p(a))
    if 'lbx' in qpd:      # state-box rows (keep_state_box)
        keep['lbx'] = np.ascontiguousarray(qpd['lbx'], float)
        keep['ubx'] = np.ascontiguousarray(qpd['ubx'], float)
        qp.lbx, qp.ubx, qp.xbox = _p(keep['lbx']), _p(keep['ubx']), 1
    dx, du = np.zeros((N + 1, NX)), np.zeros((N, NU))
    ll, lu = np.zeros((N, NU)), np.zeros((N, NU))
    it = C.c_int(0)
    st = lib().orc_qp_solve(C.byref(qp), C.c_int(iter_max), C.c_double(tol), _p(dx), _p(du),
                            _p(ll), _p(lu), C.byref(it))
    return dict(status=st, iters=it.value, dx=dx, du=du, lam_l=ll, lam_u=lu)


class OracleBatch:
    """Batch state (warm-start store, valid mask, last command) + orc_solve_batch."""

    def __init__(self, cfg: OrcConfig, batch: int):
        self.cfg, self.B, N = cfg, batch, cfg.N
        self.xtraj = np.zeros((batch, N + 1, NX))
        self.utraj = np.zeros((batch, N, NU))
        self.valid = np.zeros(batch, np.uint8)
        hover = cfg.mass * cfg.gravity / 4.0
        self.last_cmd = np.full((batch, NU), hover)
        self.qp_iters = np.zeros(batch, np.int32)

    def solve(self, x0, yref, yref_e):
        B = self.B
        x0 = np.ascontiguousarray(x0, float).reshape(B, NX)
        yref = np.ascontiguousarray(yref, float).reshape(B, self.cfg.N, NZ)
        yref_e = np.ascontiguousarray(yref_e, float).reshape(B, NX)
        u0, status = np.zeros((B, NU)), np.zeros(B, np.int32)
        thrust, speed = np.zeros((B, NU)), np.zeros((B, NU))
        lib().orc_solve_batch(C.byref(self.cfg), C.c_int(B), _p(x0), _p(yref), _p(yref_e),
                              _p(self.xtraj), _p(self.utraj), _p(self.valid, _bp),
                              _p(self.last_cmd), _p(u0), _p(status, _ip),
                              _p(self.qp_iters, _ip), _p(thrust), _p(speed))
        return u0, status, thrust, speed


def stack_reference(reference: Dict[str, np.ndarray], N: int) -> Tuple[np.ndarray, np.ndarray]:
    """controller.py:433-445: per-stage yref = [p,v,q,w,thrust], terminal = [p,v,q,w]."""
    yref = np.hstack([np.asarray(reference['positions'], float)[:N],
                      np.asarray(reference['velocities'], float)[:N],
                      np.asarray(reference['quaternions'], float)[:N],
                      np.asarray(reference['body_rates'], float)[:N],
                      np.asarray(reference['thrusts'], float)[:N]])
    yref_e = np.hstack([np.asarray(reference['positions'], float)[-1],
                        np.asarray(reference['velocities'], float)[-1],
                        np.asarray(reference['quaternions'], float)[-1],
                        np.asarray(reference['body_rates'], float)[-1]])
    return yref, yref_e


class OraclePositionNMPC:
    """Same call surface as the reference's PositionNMPC, backed by the C oracle."""

    def __init__(self, params, **flags):
        self.cfg = config_from_params(params, **flags)
        self._batch = OracleBatch(self.cfg, 1)
        self.nx, self.nu = NX, NU

    @property
    def horizon(self):
        return self.cfg.N

    @property
    def hover_thrust(self):
        return self.cfg.mass * self.cfg.gravity / 4.0

    def solve(self, state, reference):
        q = np.asarray(state['quaternion'], float).reshape(4)
        n = np.linalg.norm(q)
        if n == 0.0:
            raise ValueError('Quaternion norm must be non-zero.')
        x0 = np.concatenate([np.asarray(state['position'], float).reshape(3),
                             np.asarray(state['velocity'], float).reshape(3), q / n,
                             np.asarray(state['body_rates'], float).reshape(3)])
        yref, yref_e = stack_reference(reference, self.cfg.N)
        u0, status, _, _ = self._batch.solve(x0[None], yref[None], yref_e[None])
        return u0[0].copy(), int(status[0])


# ------------------------------------------------------------------ closed loop (f1-f3 checkers)
def plant_step_ext(cfg, x, u, dt, substeps=1, motor_model=0, tau_up=0.0125, tau_down=0.025,
                   rotor_drag=0.0, mass=0.0, rotor=None, rolling_moment=0.0, rotors_yaw_sign=False):
    """orc_plant_step_full: returns (xn, rotor_speed_new)."""
    xn = np.zeros(NX)
    rot = np.zeros(NU) if rotor is None else np.ascontiguousarray(rotor, float).copy()
    lib().orc_plant_step_full.restype = None
    lib().orc_plant_step_full(C.byref(cfg), _p(np.ascontiguousarray(x, float)),
                              _p(np.ascontiguousarray(u, float)), C.c_double(dt),
                              C.c_int(substeps), C.c_int(motor_model), C.c_double(tau_up),
                              C.c_double(tau_down), C.c_double(rotor_drag), C.c_double(mass),
                              C.c_double(rolling_moment), C.c_int(int(bool(rotors_yaw_sign))),
                              _p(rot), _p(xn))
    return xn, rot


def setpoint_reference(cfg, positions, yaws=None):
    """B x (ReferenceGenerator.build_horizon, reference.py:75-91, with update_hover_thrust :70-72)
    stacked as controller.py:433-445."""
    p = np.asarray(positions, float).reshape(-1, 3)
    B, N = p.shape[0], cfg.N
    yaw = np.zeros(B) if yaws is None else np.asarray(yaws, float).reshape(B)
    hover = cfg.mass * cfg.gravity / 4.0
    yref, yref_e = np.zeros((B, N, NZ)), np.zeros((B, NX))
    for b in range(B):
        half = 0.5 * yaw[b]                       # reference.py:11-13
        q = np.array([math.cos(half), 0.0, 0.0, math.sin(half)])
        row = np.concatenate([p[b], np.zeros(3), q, np.zeros(3)])
        yref[b, :, :NX] = row
        yref[b, :, NX:] = hover
        yref_e[b] = row
    return yref, yref_e


def figure8_reference(cfg, t0, amplitude, period, centre=(0.0, 0.0, 1.5), t_offset=0.0):
    """p(t) = c + (A sin wt, A/2 sin 2wt, 0) and its velocity at t0 + t_offset + k dt."""
    t0 = np.asarray(t0, float).reshape(-1)
    B, N = t0.shape[0], cfg.N
    A = np.broadcast_to(np.asarray(amplitude, float), (B,))
    per = np.broadcast_to(np.asarray(period, float), (B,))
    hover = cfg.mass * cfg.gravity / 4.0
    yref, yref_e = np.zeros((B, N, NZ)), np.zeros((B, NX))
    for b in range(B):
        w = 2.0 * math.pi / per[b]
        for k in range(N + 1):
            wt = w * (t0[b] + t_offset + cfg.dt * k)
            row = np.zeros(NX)
            row[0] = centre[0] + A[b] * math.sin(wt)
            row[1] = centre[1] + 0.5 * A[b] * math.sin(2.0 * wt)
            row[2] = centre[2]
            row[3] = A[b] * w * math.cos(wt)
            row[4] = A[b] * w * math.cos(2.0 * wt)
            row[6] = 1.0
            if k < N:
                yref[b, k, :NX] = row
                yref[b, k, NX:] = hover
            else:
                yref_e[b] = row
    return yref, yref_e


def odometry_to_state(odom):
    """nodes/mpc_controller_node:88-113 (+ quaternion_to_euler :25-44, _rotation_matrix :139-150)
    followed by controller.py:401-411.  odom = [p(3), q_xyzw(4), v_body(3), w(3)].
    Returns (x0[13], zero_quaternion flag)."""
    o = np.asarray(odom, float).reshape(13)
    qx, qy, qz, qw = o[3], o[4], o[5], o[6]
    norm = math.sqrt(qx * qx + qy * qy + qz * qz + qw * qw)
    if norm == 0.0:
        roll = pitch = yaw = 0.0
    else:
        ex, ey, ez, ew = qx / norm, qy / norm, qz / norm, qw / norm
        roll = math.atan2(2.0 * (ew * ex + ey * ez), 1.0 - 2.0 * (ex * ex + ey * ey))
        sinp = 2.0 * (ew * ey - ez * ex)
        pitch = math.copysign(math.pi / 2.0, sinp) if abs(sinp) >= 1.0 else math.asin(sinp)
        yaw = math.atan2(2.0 * (ew * ez + ex * ey), 1.0 - 2.0 * (ey * ey + ez * ez))
    cr, sr, cp, sp = math.cos(roll), math.sin(roll), math.cos(pitch), math.sin(pitch)
    cy, sy = math.cos(yaw), math.sin(yaw)
    rot = np.array([[cp * cy, cy * sp * sr - sy * cr, cy * sp * cr + sy * sr],
                    [cp * sy, sy * sp * sr + cy * cr, sy * sp * cr - cy * sr],
                    [-sp, cp * sr, cp * cr]])
    q = np.array([qw, qx, qy, qz])
    nq = np.linalg.norm(q)
    x0 = np.concatenate([o[0:3], rot @ o[7:10], q / nq if nq != 0.0 else q, o[10:13]])
    return x0, nq == 0.0


def closed_loop(cfg, x_init, steps, ref_fn, pref_fn, plant=None, rotor_init=None):
    """Reference closed loop for a batch: per step reference -> orc_solve_batch (warm start,
    command map, last-valid fallback) -> plant step -> statistics, exactly the order of
    nmpc_rollout_device.  ref_fn(k) -> (yref, yref_e); pref_fn(k) -> p_ref[B,3] the error of the
    NEW state is measured against.  Returns (x, stats[B,8], rotor, last outputs)."""
    plant = {**dict(dt=cfg.dt, substeps=1, motor_model=0, tau_up=0.0125, tau_down=0.025,
                    rotor_drag=0.0, mass=0.0, rolling_moment=0.0, rotors_yaw_sign=False),
             **(plant or {})}
    x = np.array(x_init, float).reshape(-1, NX).copy()
    B = x.shape[0]
    ob = OracleBatch(cfg, B)
    rotor = np.zeros((B, NU)) if rotor_init is None else np.array(rotor_init, float).copy()
    stats = np.zeros((B, 8))
    out = None
    for k in range(steps):
        yref, yref_e = ref_fn(k)
        x0 = x.copy()
        nq = np.linalg.norm(x0[:, 6:10], axis=1)
        x0[:, 6:10] = x0[:, 6:10] / np.where(nq != 0.0, nq, 1.0)[:, None]
        u0, status, thrust, speed = ob.solve(x0, yref, yref_e)
        out = (u0, status, thrust, speed)
        pref = pref_fn(k)
        for b in range(B):
            x[b], rot = plant_step_ext(cfg, x[b], thrust[b], plant['dt'], plant['substeps'],
                                       plant['motor_model'], plant['tau_up'], plant['tau_down'],
                                       plant['rotor_drag'], plant['mass'], rotor[b],
                                       plant['rolling_moment'], plant['rotors_yaw_sign'])
            if plant['motor_model']:
                rotor[b] = rot
        e = np.linalg.norm(x[:, 0:3] - pref, axis=1)
        stats[:, 0] += e * e
        stats[:, 1] = e
        stats[:, 2] += (status != 0)
        stats[:, 3] += ob.qp_iters
        lo, hi = np.array(cfg.lbu[:]), np.array(cfg.ubu[:])
        tol = 1e-9 * (hi - lo)
        stats[:, 4] += np.any((thrust <= lo + tol) | (thrust >= hi - tol), axis=1)
        stats[:, 5] = np.maximum(stats[:, 5], e)
        stats[:, 6] += thrust.sum(axis=1)
        stats[:, 7] += ~np.all(np.isfinite(x), axis=1)
    return x, stats, rotor, out
