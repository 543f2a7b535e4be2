"""Closed-loop Monte-Carlo runs that stay on the device.

What the reference does one vehicle at a time through ROS + Gazebo --
``nodes/mpc_controller_node``: odometry -> state (:88-113), reference horizon (:119-120), solve
(:122-123), command map + last-valid fail-safe (:124-131, :152-164) -- with the RotorS plant
integrating in between, is run here for B vehicles per GPU by ``nmpc_rollout_device``: per control
step one launch of the solve kernel and one of the plant kernel (plus the reference kernel for a
moving track); states, references, warm starts and commands never leave HBM.

The plant is the controller's own model (``controller.py:286-346``) integrated with RK4; the
first-order rotor-speed lag and rotor drag of RotorS' motor model
(``rotors_gazebo_plugins/src/gazebo_motor_model.cpp:415-483``, ``common.h:147-183``) are optional.
the rolling moment (``:470-474``) and RotorS' own yaw-torque sign (``:458-459``) as well -- the
opt-in "model-mismatch" plant.  Gazebo's rigid-body engine and sensor noise are not restated.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Dict, Optional

import numpy as np

from . import _abi
from .controller import BatchedPositionNMPC

STAT_NAMES = ('sum_sq_pos_err', 'last_pos_err', 'n_failed_solves', 'sum_qp_iters',
              'n_saturated_steps', 'max_pos_err', 'sum_thrust', 'n_nonfinite')


class _DeviceArray:
    """Handle-owned device memory seen through ``__cuda_array_interface__`` (zero-copy)."""

    def __init__(self, ptr: int, shape, typestr: str) -> None:
        self.__cuda_array_interface__ = {'shape': tuple(shape), 'typestr': typestr,
                                         'data': (int(ptr), False), 'version': 2}


@dataclass
class Plant:
    """``nmpc_plant_t``.  Defaults: nominal plant, 1 ms RK4 steps over the control interval are
    obtained with ``substeps = round(dt / 0.001)`` (Gazebo's max_step_size,
    ``launch/hummingbird_mpc.launch:9-10``)."""
    dt: float
    substeps: int = 1
    motor_model: int = 0
    tau_up: float = 0.0125          # rotors_description/urdf/hummingbird.xacro:29-46
    tau_down: float = 0.025
    rotor_drag: float = 0.0         # 8.06428e-05 in hummingbird.xacro
    mass: float = 0.0               # <= 0: the controller's mass (RotorS: 0.716 = 0.68 + 4 x 0.009)
    rolling_moment: float = 0.0     # 1e-06 in hummingbird.xacro:42
    rotors_yaw_sign: bool = False   # True: RotorS' yaw-torque sign (opposite to the controller's)

    def to_c(self) -> _abi.NmpcPlant:
        p = _abi.NmpcPlant()
        p.substeps, p.motor_model, p.dt = int(self.substeps), int(self.motor_model), float(self.dt)
        p.tau_up, p.tau_down = float(self.tau_up), float(self.tau_down)
        p.rotor_drag, p.mass = float(self.rotor_drag), float(self.mass)
        p.rolling_moment = float(self.rolling_moment)
        p.rotors_yaw_sign, p.reserved0 = int(bool(self.rotors_yaw_sign)), 0
        return p


class ClosedLoop:
    """Device-resident closed loop around a :class:`BatchedPositionNMPC`.

    All array arguments are torch CUDA float64 tensors (or numpy arrays, which are uploaded once).
    """

    def __init__(self, controller: BatchedPositionNMPC, plant: Optional[Plant] = None) -> None:
        import torch
        self.ctl = controller
        self.plant = plant or Plant(dt=controller.dt)
        self.dev = torch.device('cuda', controller.device)
        self._torch = torch
        B = controller.batch
        self.x = torch.zeros((B, 13), dtype=torch.float64, device=self.dev)
        self.rotor_speed = torch.zeros((B, 4), dtype=torch.float64, device=self.dev)
        self.stats = torch.zeros((B, _abi.NSTATS), dtype=torch.float64, device=self.dev)
        self.t = 0.0
        self._keep = []

    # -- helpers ---------------------------------------------------------------------------
    def _dev(self, a, shape):
        torch = self._torch
        if a is None:
            return None
        t = torch.as_tensor(np.asarray(a, dtype=np.float64) if not hasattr(a, 'data_ptr') else a,
                            dtype=torch.float64, device=self.dev).contiguous()
        if tuple(t.shape) != tuple(shape):
            raise ValueError(f'expected shape {shape}, got {tuple(t.shape)}')
        return t

    def _stream(self):
        return self._torch.cuda.current_stream(self.dev).cuda_stream

    def reset(self, x_init, hover_rotors: bool = True) -> None:
        """Set the plant state (quaternion as given), drop warm starts / last commands / stats."""
        B = self.ctl.batch
        self.x.copy_(self._dev(x_init, (B, 13)))
        w_hover = float(np.sqrt(self.ctl.hover_thrust / max(self.ctl.rotor_force_constant, 1e-9)))
        self.rotor_speed.fill_(w_hover if hover_rotors else 0.0)
        self.stats.zero_()
        self.t = 0.0
        self._torch.cuda.synchronize(self.dev)
        self.ctl.invalidate()
        self.ctl.reset_last_command()

    # -- single pieces (also used by the parity tests) ----------------------------------------
    def reference_setpoint(self, positions, yaws=None):
        torch, B, N = self._torch, self.ctl.batch, self.ctl.horizon
        pos = self._dev(positions, (B, 3))
        yaw = self._dev(yaws, (B,))
        yref = torch.empty((B, N, 17), dtype=torch.float64, device=self.dev)
        yref_e = torch.empty((B, 13), dtype=torch.float64, device=self.dev)
        _abi.check(_abi.lib().nmpc_reference_setpoint_device(
            self.ctl._h.ptr, pos.data_ptr(), None if yaw is None else yaw.data_ptr(),
            yref.data_ptr(), yref_e.data_ptr(), self._stream()), 'nmpc_reference_setpoint_device')
        return yref, yref_e

    def reference_figure8(self, t0, amplitude, period, centre=(0.0, 0.0, 1.5), t_offset=0.0):
        torch, B, N = self._torch, self.ctl.batch, self.ctl.horizon
        t0, amp, per = (self._dev(a, (B,)) for a in (t0, amplitude, period))
        yref = torch.empty((B, N, 17), dtype=torch.float64, device=self.dev)
        yref_e = torch.empty((B, 13), dtype=torch.float64, device=self.dev)
        c = (C.c_double * 3)(*[float(v) for v in centre])
        _abi.check(_abi.lib().nmpc_reference_figure8_device(
            self.ctl._h.ptr, t0.data_ptr(), amp.data_ptr(), per.data_ptr(), c, float(t_offset),
            yref.data_ptr(), yref_e.data_ptr(), self._stream()), 'nmpc_reference_figure8_device')
        return yref, yref_e

    def odometry_to_state(self, odom):
        """``odom[B,13] = [p, q_xyzw, v_body, w]`` -> ``(x0[B,13], zero_quaternion[B])``."""
        torch, B = self._torch, self.ctl.batch
        od = self._dev(odom, (B, 13))
        x0 = torch.empty((B, 13), dtype=torch.float64, device=self.dev)
        bad = torch.empty((B,), dtype=torch.int32, device=self.dev)
        _abi.check(_abi.lib().nmpc_odometry_to_state_device(
            self.ctl._h.ptr, od.data_ptr(), x0.data_ptr(), bad.data_ptr(), self._stream()),
            'nmpc_odometry_to_state_device')
        return x0, bad

    def plant_step(self, thrust):
        """Advance ``self.x`` one control interval under ``thrust[B,4]``; returns the state the
        solver is fed next (quaternion normalised)."""
        torch, B = self._torch, self.ctl.batch
        th = self._dev(thrust, (B, 4))
        x0 = torch.empty((B, 13), dtype=torch.float64, device=self.dev)
        p = self.plant.to_c()
        _abi.check(_abi.lib().nmpc_plant_step_device(
            self.ctl._h.ptr, C.byref(p), self.x.data_ptr(), x0.data_ptr(), th.data_ptr(),
            self.rotor_speed.data_ptr() if self.plant.motor_model else None, self._stream()),
            'nmpc_plant_step_device')
        self.t += self.plant.dt
        return x0

    # -- whole runs -----------------------------------------------------------------------------
    def _run(self, r: _abi.NmpcRollout, steps: int) -> None:
        r.steps = int(steps)
        r.plant = self.plant.to_c()
        r.x = self.x.data_ptr()
        r.rotor_speed = self.rotor_speed.data_ptr() if self.plant.motor_model else None
        r.t_start = float(self.t)
        r.stats = self.stats.data_ptr()
        _abi.check(_abi.lib().nmpc_rollout_device(self.ctl._h.ptr, C.byref(r), self._stream()),
                   'nmpc_rollout_device')
        self.t += steps * self.plant.dt

    def run_setpoint(self, steps: int, positions, yaws=None) -> None:
        """``steps`` control steps towards constant set-points (asynchronous on the current stream)."""
        B = self.ctl.batch
        pos, yaw = self._dev(positions, (B, 3)), self._dev(yaws, (B,))
        self._keep = [pos, yaw]
        r = _abi.NmpcRollout()
        r.ref_mode = _abi.REF_SETPOINT
        r.sp_pos = pos.data_ptr()
        r.sp_yaw = None if yaw is None else yaw.data_ptr()
        self._run(r, steps)

    def run_figure8(self, steps: int, t0, amplitude, period, centre=(0.0, 0.0, 1.5)) -> None:
        B = self.ctl.batch
        t0, amp, per = (self._dev(a, (B,)) for a in (t0, amplitude, period))
        self._keep = [t0, amp, per]
        r = _abi.NmpcRollout()
        r.ref_mode = _abi.REF_FIGURE8
        r.f8_t0, r.f8_amplitude, r.f8_period = t0.data_ptr(), amp.data_ptr(), per.data_ptr()
        r.f8_centre[:] = [float(v) for v in centre]
        self._run(r, steps)

    def last_outputs(self):
        """(u0, status, thrust, speed) of the last solve of a run (torch CUDA tensors, copies)."""
        torch, B = self._torch, self.ctl.batch
        ptrs = [C.c_void_p() for _ in range(4)]
        _abi.check(_abi.lib().nmpc_last_outputs(self.ctl._h.ptr, *[C.byref(p) for p in ptrs]),
                   'nmpc_last_outputs')
        out = []
        for p, (shape, ts) in zip(ptrs, (((B, 4), '<f8'), ((B,), '<i4'), ((B, 4), '<f8'),
                                         ((B, 4), '<f8'))):
            out.append(torch.as_tensor(_DeviceArray(p.value, shape, ts), device=self.dev).clone())
        return tuple(out)

    def summary(self, group=None, reduce: bool = True) -> Dict[str, float]:
        """Reduce the per-problem statistics on the device (and over ranks when a process group
        is initialised): the <=64-entry vector north_star lets NCCL carry."""
        torch = self._torch
        s = self.stats
        vec = torch.stack([
            torch.tensor(float(s.shape[0]), dtype=torch.float64, device=self.dev),
            s[:, 0].sum(), s[:, 1].sum(), s[:, 2].sum(), s[:, 3].sum(), s[:, 4].sum(),
            s[:, 6].sum(), s[:, 7].sum(), (s[:, 2] > 0).sum().to(torch.float64)])
        mx = torch.stack([s[:, 5].max(), s[:, 1].max()])
        import torch.distributed as dist
        if (reduce and dist.is_available() and dist.is_initialized()
                and dist.get_world_size(group) > 1):
            dist.all_reduce(vec, op=dist.ReduceOp.SUM, group=group)
            dist.all_reduce(mx, op=dist.ReduceOp.MAX, group=group)
        v, m = vec.cpu().numpy(), mx.cpu().numpy()
        n = v[0]
        return {'problems': n, 'mean_sq_pos_err_sum': v[1] / n, 'mean_final_pos_err': v[2] / n,
                'failed_solves': v[3], 'qp_iters': v[4], 'saturated_steps': v[5],
                'thrust_sum': v[6], 'nonfinite': v[7], 'problems_with_failures': v[8],
                'max_pos_err': m[0], 'max_final_pos_err': m[1]}
