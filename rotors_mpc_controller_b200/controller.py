"""NMPC controller front-end backed by the sm_100a CUDA solver (``libnmpc_b200.so``).

Drop-in for ``PositionNMPC`` of
``/root/reference/src/rotors_mpc_controller/controller.py`` (constructor :55-60,
``reconfigure`` :63-172, ``solve`` :385-463, accessors :358-382): same parameter dict, same
``solve(state, reference) -> (u0, status)`` contract, same warm-start / fail-safe semantics and
acados-style status codes.  What changes is everything below that API: instead of describing an
``AcadosOcp``, generating C and compiling it, the constants are packed into an ``nmpc_config_t``
and handed to a CUDA handle that owns the warm-start store and the last-valid command on the
device.  There is no CPU fallback: without the shared library or a CUDA device construction
fails.

``BatchedPositionNMPC`` exposes the same solver for B independent vehicles at once.
"""
from __future__ import annotations

import ctypes as C
import hashlib
from dataclasses import dataclass
from pathlib import Path
from typing import Dict, Optional, Tuple

import numpy as np

from . import _abi
from .reference import stack_reference

_INTEGRATORS = {'euler': 1, 'midpoint': 2, 'rk4': 4}


@dataclass
class ControllerParams:
    """Same fields as the reference dataclass (controller.py:24-49)."""
    horizon_steps: int
    dt: float
    position_weight: np.ndarray
    velocity_weight: np.ndarray
    quaternion_weight: np.ndarray
    rate_weight: np.ndarray
    control_weight: np.ndarray
    terminal_weight: np.ndarray
    regularization: float
    solver_iter_max: int
    codegen_directory: Path
    mass: float
    inertia: np.ndarray
    gravity: float
    rotor_force_constant: float
    rotor_moment_constant: float
    rotor_x_offsets: np.ndarray
    rotor_y_offsets: np.ndarray
    rotor_z_torque: np.ndarray
    input_lower_bounds: np.ndarray
    input_upper_bounds: np.ndarray
    hover_thrust_per_motor: float
    motor_min_speed: float
    motor_max_speed: float


def _vec(cfg, key, default):
    return np.asarray(cfg.get(key, default), dtype=float)


def controller_params_from_dict(params: Dict[str, Dict[str, object]]) -> ControllerParams:
    """params dict -> ControllerParams, following controller.py:69-155 (incl. its defaults,
    KeyError on missing solver/vehicle/horizon_steps/dt/mass, ValueError on a non-'+' frame)."""
    solver_cfg = params['solver']
    vehicle_cfg = params['vehicle']
    world_cfg = params.get('world', {})
    mass = float(vehicle_cfg['mass'])
    gravity = float(world_cfg.get('gravity', 9.81))
    inertia = _vec(vehicle_cfg, 'inertia', [0.007, 0, 0, 0, 0.007, 0, 0, 0, 0.012]).reshape(3, 3)
    arm = float(vehicle_cfg.get('arm_length', 0.17))
    kf = float(vehicle_cfg.get('rotor_force_constant', 8.54858e-6))
    km = float(vehicle_cfg.get('rotor_moment_constant', 0.016))
    w_min = float(vehicle_cfg.get('motor_min_speed', 0.0))
    w_max = float(vehicle_cfg.get('motor_max_speed', 2000.0))
    layout = str(vehicle_cfg.get('rotor_configuration', '+')).strip()
    if layout != '+':
        raise ValueError(
            f'rotors_mpc_controller only supports a "+" rotor lay-out, got "{layout}".')
    # rotor 0 front, 1 left, 2 back, 3 right; spin -1,+1,-1,+1 (controller.py:98-103)
    spin = np.array([-1.0, 1.0, -1.0, 1.0])
    # the thrust box comes from the motor speed range, NOT from controller.thrust_limits
    # (controller.py:105-110; the YAML key is parsed and never used by the reference)
    lo = max(0.0, kf * w_min ** 2)
    hi = kf * w_max ** 2
    if bool(solver_cfg.get('use_thrust_limits', False)):  # opt-in extension
        lo, hi = (float(v) for v in params['controller']['thrust_limits'])
    codegen = Path(solver_cfg.get('codegen_directory',
                                  Path.home() / '.cache' / 'rotors_mpc_controller')).expanduser()
    return ControllerParams(
        horizon_steps=int(solver_cfg['horizon_steps']),
        dt=float(solver_cfg['dt']),
        position_weight=_vec(solver_cfg, 'position_weight', [10.0, 10.0, 8.0]),
        velocity_weight=_vec(solver_cfg, 'velocity_weight', [1.0, 1.0, 0.2]),
        quaternion_weight=_vec(solver_cfg, 'quaternion_weight', [3.2] * 4),
        rate_weight=_vec(solver_cfg, 'rate_weight', [1.4, 1.4, 0.4]),
        control_weight=_vec(solver_cfg, 'control_weight', [1.75] * 4),
        terminal_weight=_vec(solver_cfg, 'terminal_weight',
                             [5.0, 5.0, 3.0, 2.0, 2.0, 2.0, 12.0, 12.0, 12.0, 18.5, 2.0, 2.0, 1.8]),
        regularization=float(solver_cfg.get('regularization', 7.0e-3)),
        solver_iter_max=int(solver_cfg.get('iter_max', 600)),
        codegen_directory=codegen,
        mass=mass,
        inertia=np.array([inertia[0, 0], inertia[1, 1], inertia[2, 2]]),
        gravity=gravity,
        rotor_force_constant=kf,
        rotor_moment_constant=km,
        rotor_x_offsets=np.array([arm, 0.0, -arm, 0.0]),
        rotor_y_offsets=np.array([0.0, arm, 0.0, -arm]),
        rotor_z_torque=spin * km,
        input_lower_bounds=np.full(4, lo),
        input_upper_bounds=np.full(4, hi),
        hover_thrust_per_motor=mass * gravity / 4.0,
        motor_min_speed=w_min,
        motor_max_speed=w_max,
    )


def abi_config(cp: ControllerParams, solver_cfg: Dict[str, object]) -> _abi.NmpcConfig:
    """ControllerParams (+ optional extension keys of the solver section) -> nmpc_config_t.
    Defaults reproduce the acados options of controller.py:179-190."""
    c = _abi.NmpcConfig()
    c.N = cp.horizon_steps
    c.qp_iter_max = cp.solver_iter_max
    integ = str(solver_cfg.get('integrator', 'midpoint')).lower()
    if integ not in _INTEGRATORS:
        raise ValueError(f'unknown integrator "{integ}" (euler | midpoint | rk4)')
    c.integ_ns = _INTEGRATORS[integ]                       # ERK, num_stages=2 (:183,:187)
    c.integ_steps = int(solver_cfg.get('integrator_steps', 2))  # num_steps=2 (:188)
    c.cost_scale_dt = int(bool(solver_cfg.get('cost_scale_dt', True)))
    c.lm_scaled_by_dt = int(bool(solver_cfg.get('lm_scaled_by_dt', True)))
    c.max_sqp_iter = int(solver_cfg.get('max_sqp_iter', 1))     # SQP_RTI (nlp_solver_type unset)
    c.nan_status = int(solver_cfg.get('nan_status', 4))
    c.ipm_warm_start = int(bool(solver_cfg.get('ipm_warm_start', True)))
    c.active_set_shortcut = int(bool(solver_cfg.get('active_set_shortcut', True)))
    disp = str(solver_cfg.get('dispatch', 'auto')).lower()
    if disp not in ('auto', 'fifo', 'lpt'):
        raise ValueError('dispatch must be "auto", "fifo" or "lpt"')
    c.dispatch = {'auto': 0, 'fifo': 1, 'lpt': 2}[disp]
    prec = str(solver_cfg.get('precision', 'f64')).lower()
    if prec not in ('f64', 'f32'):
        raise ValueError('precision must be "f64" or "f32"')
    c.precision = _abi.PRECISION_F64 if prec == 'f64' else _abi.PRECISION_F32
    c.dt = cp.dt
    stage_w = np.concatenate([cp.position_weight, cp.velocity_weight, cp.quaternion_weight,
                              cp.rate_weight, cp.control_weight])      # :234-240
    if stage_w.shape != (17,) or cp.terminal_weight.shape != (13,):
        raise ValueError('weight vectors must have 3+3+4+3+4 and 13 entries')
    c.W[:] = stage_w.tolist()
    c.We[:] = cp.terminal_weight.tolist()
    c.lm = cp.regularization
    c.mass = cp.mass
    c.J[:] = cp.inertia.tolist()
    c.gravity = cp.gravity
    c.rotor_x[:] = cp.rotor_x_offsets.tolist()
    c.rotor_y[:] = cp.rotor_y_offsets.tolist()
    c.rotor_z[:] = cp.rotor_z_torque.tolist()
    c.lbu[:] = cp.input_lower_bounds.tolist()
    c.ubu[:] = cp.input_upper_bounds.tolist()
    c.kf = cp.rotor_force_constant
    c.w_min = cp.motor_min_speed
    c.w_max = cp.motor_max_speed
    c.sqp_tol = float(solver_cfg.get('sqp_tol', 1e-8))
    c.ipm_tol = float(solver_cfg.get('ipm_tol', 1e-14))
    return c


def config_signature(cp: ControllerParams) -> str:
    """sha1 identity of all constants, as the reference keys its codegen dir (:192-213)."""
    sig = repr((cp.horizon_steps, cp.dt, tuple(cp.position_weight), tuple(cp.velocity_weight),
                tuple(cp.quaternion_weight), tuple(cp.rate_weight), tuple(cp.control_weight),
                tuple(cp.terminal_weight), cp.mass, tuple(cp.inertia), cp.gravity,
                cp.rotor_force_constant, cp.rotor_moment_constant, tuple(cp.rotor_x_offsets),
                tuple(cp.rotor_y_offsets), tuple(cp.rotor_z_torque),
                tuple(cp.input_lower_bounds), tuple(cp.input_upper_bounds), cp.regularization,
                cp.solver_iter_max))
    return hashlib.sha1(sig.encode('utf-8')).hexdigest()[:10]


class _Handle:
    """Owner of one ``nmpc_handle_t*``."""

    def __init__(self, cfg: _abi.NmpcConfig, batch: int, device: int):
        self._lib = _abi.lib()
        self.ptr = C.c_void_p()
        _abi.check(self._lib.nmpc_create(C.byref(cfg), batch, device, C.byref(self.ptr)),
                   'nmpc_create')

    def close(self):
        if getattr(self, 'ptr', None) is not None and self.ptr:
            self._lib.nmpc_destroy(self.ptr)
            self.ptr = None

    def __del__(self):  # pragma: no cover
        try:
            self.close()
        except Exception:
            pass


class BatchedPositionNMPC:
    """B independent hummingbird NMPC problems solved by one kernel launch.

    ``solve(x0[B,13], yref[B,N,17], yref_e[B,13])`` takes numpy arrays (host path: the copies
    are part of the call) or torch CUDA float64 tensors (device path, asynchronous on the
    current torch stream) and returns ``(u0[B,4], status[B])`` of the same kind;
    ``command()`` returns the clipped thrusts and rotor speeds of the last solve with the
    node's last-valid fallback applied (nodes/mpc_controller_node:124-131,152-164).
    """

    nx, nu, ny, ny_e = 13, 4, 17, 13

    def __init__(self, params: Dict[str, Dict[str, object]], batch: int, device: int = 0) -> None:
        self.batch = int(batch)
        self.device = int(device)
        self._h: Optional[_Handle] = None
        self.reconfigure(params)

    # -- construction ----------------------------------------------------------------
    def reconfigure(self, params: Dict[str, Dict[str, object]]) -> None:
        cp = controller_params_from_dict(params)
        cfg = abi_config(cp, params['solver'])
        same_shape = (self._h is not None and cfg.N == self._cfg.N
                      and cfg.precision == self._cfg.precision)
        if same_shape:   # constants swap only: no allocation, no compilation
            _abi.check(_abi.lib().nmpc_reconfigure(self._h.ptr, C.byref(cfg)), 'nmpc_reconfigure')
        else:
            new = _Handle(cfg, self.batch, self.device)
            if self._h is not None:
                self._h.close()
            self._h = new
        self.config, self._cfg = cp, cfg
        self.mass, self.gravity = cp.mass, cp.gravity
        self.codegen_path = cp.codegen_directory / (
            f'h{cp.horizon_steps}_dt{int(cp.dt * 1e6)}_{config_signature(cp)}')
        self._out = None

    def close(self) -> None:
        if self._h is not None:
            self._h.close()
            self._h = None

    # -- accessors (controller.py:358-382) ----------------------------------------------
    @property
    def horizon(self) -> int:
        return self.config.horizon_steps

    @property
    def dt(self) -> float:
        return self.config.dt

    @property
    def hover_thrust(self) -> float:
        return self.config.hover_thrust_per_motor

    @property
    def rotor_force_constant(self) -> float:
        return self.config.rotor_force_constant

    @property
    def motor_speed_limits(self) -> Tuple[float, float]:
        return self.config.motor_min_speed, self.config.motor_max_speed

    @property
    def input_bounds(self) -> Tuple[np.ndarray, np.ndarray]:
        return self.config.input_lower_bounds, self.config.input_upper_bounds

    @property
    def kernel_launches(self) -> int:
        return int(_abi.lib().nmpc_kernel_launches(self._h.ptr))

    @property
    def workspace_bytes(self) -> int:
        return int(_abi.lib().nmpc_workspace_bytes(self._h.ptr))

    def set_bounds(self, lower, upper) -> None:
        lo = (C.c_double * 4)(*np.broadcast_to(np.asarray(lower, float), (4,)).tolist())
        hi = (C.c_double * 4)(*np.broadcast_to(np.asarray(upper, float), (4,)).tolist())
        _abi.check(_abi.lib().nmpc_set_bounds(self._h.ptr, lo, hi), 'nmpc_set_bounds')
        self.config.input_lower_bounds = np.array(list(lo))
        self.config.input_upper_bounds = np.array(list(hi))
        self._cfg.lbu[:] = list(lo)
        self._cfg.ubu[:] = list(hi)

    def invalidate(self, mask=None) -> None:
        """Drop the warm start (all problems, or those with mask != 0)."""
        if mask is not None and hasattr(mask, 'data_ptr'):   # device mask: one async kernel
            import torch
            if (not mask.is_cuda or mask.dtype != torch.uint8 or not mask.is_contiguous()
                    or mask.numel() != self.batch or mask.device.index != self.device):
                raise ValueError('expected a contiguous CUDA uint8 mask of length batch on the '
                                 "handle's device")
            stream = torch.cuda.current_stream(mask.device).cuda_stream
            _abi.check(_abi.lib().nmpc_invalidate_device(self._h.ptr, mask.data_ptr(), stream),
                       'nmpc_invalidate_device')
            return
        m = None if mask is None else np.ascontiguousarray(mask, dtype=np.uint8)
        if m is not None and m.shape != (self.batch,):
            raise ValueError('mask must have one entry per problem')
        _abi.check(_abi.lib().nmpc_invalidate(self._h.ptr, _abi.ptr(m)), 'nmpc_invalidate')

    def debug_poison(self, on: bool = True) -> None:
        """NaN-fill workspace and shared tile before use (see nmpc_debug_poison)."""
        _abi.check(_abi.lib().nmpc_debug_poison(self._h.ptr, int(bool(on))), 'nmpc_debug_poison')

    def reset_last_command(self) -> None:
        _abi.check(_abi.lib().nmpc_reset_last_command(self._h.ptr), 'nmpc_reset_last_command')

    # -- solve --------------------------------------------------------------------------
    def solve(self, x0, yref, yref_e):
        B, N = self.batch, self.horizon
        lib = _abi.lib()
        if hasattr(x0, 'data_ptr'):  # torch tensors on the device
            import torch
            for t, shape in ((x0, (B, 13)), (yref, (B, N, 17)), (yref_e, (B, 13))):
                if (not t.is_cuda or t.dtype != torch.float64 or not t.is_contiguous()
                        or tuple(t.shape) != shape):
                    raise ValueError(f'expected contiguous CUDA float64 tensor of shape {shape}')
                if t.device.index != self.device:
                    raise ValueError(f'tensor on {t.device}, but this solver was built for '
                                     f'cuda:{self.device}')
            if self._out is None or not hasattr(self._out[0], 'data_ptr'):
                dev = x0.device
                self._out = (torch.empty((B, 4), dtype=torch.float64, device=dev),
                             torch.empty((B,), dtype=torch.int32, device=dev),
                             torch.empty((B, 4), dtype=torch.float64, device=dev),
                             torch.empty((B, 4), dtype=torch.float64, device=dev))
            u0, status, thrust, speed = self._out
            stream = torch.cuda.current_stream(x0.device).cuda_stream
            _abi.check(lib.nmpc_solve_device(self._h.ptr, x0.data_ptr(), yref.data_ptr(),
                                             yref_e.data_ptr(), u0.data_ptr(), status.data_ptr(),
                                             thrust.data_ptr(), speed.data_ptr(), stream),
                       'nmpc_solve_device')
            return u0, status
        x0 = np.ascontiguousarray(x0, dtype=np.float64).reshape(B, 13)
        yref = np.ascontiguousarray(yref, dtype=np.float64).reshape(B, N, 17)
        yref_e = np.ascontiguousarray(yref_e, dtype=np.float64).reshape(B, 13)
        if self._out is None or hasattr(self._out[0], 'data_ptr'):
            self._out = (np.empty((B, 4)), np.empty((B,), np.int32), np.empty((B, 4)),
                         np.empty((B, 4)))
        u0, status, thrust, speed = self._out
        _abi.check(lib.nmpc_solve_host(self._h.ptr, _abi.ptr(x0), _abi.ptr(yref),
                                       _abi.ptr(yref_e), _abi.ptr(u0), _abi.ptr(status),
                                       _abi.ptr(thrust), _abi.ptr(speed)), 'nmpc_solve_host')
        return u0, status

    def command(self):
        """(thrust[B,4], speed[B,4]) of the last solve (clipped; last-valid on failure)."""
        if self._out is None:
            raise RuntimeError('command() called before solve()')
        return self._out[2], self._out[3]

    # -- state of the warm-start store ---------------------------------------------------
    def trajectories(self):
        B, N = self.batch, self.horizon
        x = np.empty((B, N + 1, 13))
        u = np.empty((B, N, 4))
        v = np.empty((B,), np.uint8)
        _abi.check(_abi.lib().nmpc_get_trajectories(self._h.ptr, _abi.ptr(x), _abi.ptr(u),
                                                    _abi.ptr(v), 0), 'nmpc_get_trajectories')
        return x, u, v.astype(bool)

    def set_trajectories(self, x, u, valid) -> None:
        B, N = self.batch, self.horizon
        x = np.ascontiguousarray(x, np.float64).reshape(B, N + 1, 13)
        u = np.ascontiguousarray(u, np.float64).reshape(B, N, 4)
        v = np.ascontiguousarray(valid, np.uint8).reshape(B)
        _abi.check(_abi.lib().nmpc_set_trajectories(self._h.ptr, _abi.ptr(x), _abi.ptr(u),
                                                    _abi.ptr(v)), 'nmpc_set_trajectories')

    def stats(self):
        q = np.empty((self.batch,), np.int32)
        s = np.empty((self.batch,), np.int32)
        _abi.check(_abi.lib().nmpc_get_stats(self._h.ptr, _abi.ptr(q), _abi.ptr(s), 0),
                   'nmpc_get_stats')
        return {'qp_iters': q, 'sqp_iters': s}

    def discretize(self, x, u):
        """Test hook: (xn, A, B) of the integrator + sensitivities for n (x,u) pairs."""
        import torch
        dev = torch.device('cuda', self.device)
        xt = torch.as_tensor(np.ascontiguousarray(x, np.float64).reshape(-1, 13), device=dev)
        ut = torch.as_tensor(np.ascontiguousarray(u, np.float64).reshape(-1, 4), device=dev)
        n = xt.shape[0]
        xn = torch.empty((n, 13), dtype=torch.float64, device=dev)
        A = torch.empty((n, 13, 13), dtype=torch.float64, device=dev)
        Bm = torch.empty((n, 13, 4), dtype=torch.float64, device=dev)
        stream = torch.cuda.current_stream(dev).cuda_stream
        _abi.check(_abi.lib().nmpc_discretize_device(self._h.ptr, n, xt.data_ptr(), ut.data_ptr(),
                                                     xn.data_ptr(), A.data_ptr(), Bm.data_ptr(),
                                                     stream), 'nmpc_discretize_device')
        torch.cuda.synchronize(dev)
        return xn.cpu().numpy(), A.cpu().numpy(), Bm.cpu().numpy()


class PositionNMPC(BatchedPositionNMPC):
    """Single-vehicle controller with the reference's exact call surface."""

    def __init__(self, params: Dict[str, Dict[str, object]], device: int = 0) -> None:
        self._prev_solution_valid = False
        self._io = None
        super().__init__(params, batch=1, device=device)

    def reconfigure(self, params: Dict[str, Dict[str, object]]) -> None:
        super().reconfigure(params)
        self._prev_solution_valid = False      # controller.py:167
        self._io = None

    def solve(self, state: Dict[str, np.ndarray], reference: Dict[str, np.ndarray]
              ) -> Tuple[np.ndarray, int]:
        """``(u0[4], status)``; ``u0`` is zeros when ``status != 0`` (controller.py:448-450).
        Same marshalling as controller.py:401-445, written into preallocated buffers (this call is
        the latency path: every numpy temporary costs as much as a Riccati stage)."""
        N = self.horizon
        buf = self._io
        if buf is None or buf[1].shape[1] != N:
            x0 = np.empty((1, 13))
            yref = np.empty((1, N, 17))
            yref_e = np.empty((1, 13))
            out = (np.empty((1, 4)), np.empty((1,), np.int32), np.empty((1, 4)), np.empty((1, 4)))
            ptrs = [_abi.ptr(a) for a in (x0, yref, yref_e) + out]
            buf = self._io = (x0, yref, yref_e, out, ptrs)
        x0, yref, yref_e, out, ptrs = buf
        quaternion = np.asarray(state['quaternion'], dtype=float).reshape(4)
        norm_q = np.linalg.norm(quaternion)
        if norm_q == 0.0:
            raise ValueError('Quaternion norm must be non-zero.')
        x = x0[0]
        x[0:3] = np.asarray(state['position'], dtype=float).reshape(3)
        x[3:6] = np.asarray(state['velocity'], dtype=float).reshape(3)
        x[6:10] = quaternion / norm_q
        x[10:13] = np.asarray(state['body_rates'], dtype=float).reshape(3)
        pos, vel = reference['positions'], reference['velocities']       # controller.py:433-445
        quat, rate = reference['quaternions'], reference['body_rates']
        y, ye = yref[0], yref_e[0]
        y[:, 0:3] = pos[:N]
        y[:, 3:6] = vel[:N]
        y[:, 6:10] = quat[:N]
        y[:, 10:13] = rate[:N]
        y[:, 13:17] = reference['thrusts'][:N]
        ye[0:3] = pos[-1]
        ye[3:6] = vel[-1]
        ye[6:10] = quat[-1]
        ye[10:13] = rate[-1]
        _abi.check(_abi.lib().nmpc_solve_host(self._h.ptr, *ptrs), 'nmpc_solve_host')
        self._out = out
        st = int(out[1][0])
        self._prev_solution_valid = (st == 0)
        return (np.zeros(self.nu) if st != 0 else out[0][0].copy()), st

    def command(self) -> Tuple[np.ndarray, np.ndarray]:
        thrust, speed = super().command()
        return thrust[0].copy(), speed[0].copy()

    @property
    def _prev_solution(self) -> Dict[str, np.ndarray]:
        x, u, _ = self.trajectories()
        return {'x': x[0], 'u': u[0]}
