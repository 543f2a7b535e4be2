"""Reference horizons for the NMPC (host side).

``ReferenceGenerator`` keeps the call surface of the reference's
``/root/reference/src/rotors_mpc_controller/reference.py`` (:16-91): a constant set-point
(position, velocity, yaw/quaternion, body rates, per-rotor thrust) tiled over the horizon into
the dict that ``PositionNMPC.solve`` consumes.  ``stack_reference`` turns such a dict into the
``yref[N,17]`` / ``yref_e[13]`` arrays of controller.py:433-445, and ``BatchReference`` builds
whole batches of them (set-points and figure-8 tracks) for the batched solver.
"""
from __future__ import annotations

import threading
from typing import Dict, Optional, Tuple

import numpy as np


def yaw_quaternion(yaw) -> np.ndarray:
    """(w,x,y,z) of a pure yaw rotation (reference.py:11-13)."""
    half = 0.5 * np.asarray(yaw, dtype=float)
    zero = np.zeros_like(half)
    return np.stack([np.cos(half), zero, zero, np.sin(half)], axis=-1)


class ReferenceGenerator:
    """Constant set-point reference, thread-safe like the reference implementation."""

    def __init__(self, config: Dict[str, object]) -> None:
        self.frame = config.get('frame', 'world')
        self._lock = threading.Lock()
        self._position = np.asarray(config.get('default_position', [0.0, 0.0, 1.0]), dtype=float)
        self._velocity = np.asarray(config.get('default_velocity', [0.0, 0.0, 0.0]), dtype=float)
        self._yaw = float(config.get('default_yaw', 0.0))
        self._quaternion = yaw_quaternion(self._yaw)
        self._body_rates = np.zeros(3)
        self._thrust = np.zeros(4)  # a bare generator has zero thrust reference (:28)

    def set_target(self, position, velocity=None, yaw: Optional[float] = None, quaternion=None,
                   body_rates=None, thrust=None) -> None:
        with self._lock:
            self._position = np.asarray(position, dtype=float).reshape(3)
            if velocity is not None:
                self._velocity = np.asarray(velocity, dtype=float).reshape(3)
            if quaternion is not None:
                q = np.asarray(quaternion, dtype=float).reshape(4)
                n = np.linalg.norm(q)
                self._quaternion = q / n if n != 0.0 else q
                if yaw is not None:
                    self._yaw = yaw
            elif yaw is not None:
                self._yaw = float(yaw)
                self._quaternion = yaw_quaternion(self._yaw)
            if body_rates is not None:
                self._body_rates = np.asarray(body_rates, dtype=float).reshape(3)
            if thrust is not None:
                t = np.asarray(thrust, dtype=float).reshape(-1)
                self._thrust = t
                if t.shape[0] != 4:
                    raise ValueError('Thrust reference must have four components.')

    def update_defaults(self, position, velocity, yaw: float, frame: Optional[str] = None) -> None:
        with self._lock:
            self._position = np.asarray(position, dtype=float).reshape(3)
            self._velocity = np.asarray(velocity, dtype=float).reshape(3)
            self._yaw = float(yaw)
            self._quaternion = yaw_quaternion(self._yaw)
            self._body_rates = np.zeros(3)
            if frame is not None:
                self.frame = frame

    def update_hover_thrust(self, thrust_per_motor: float) -> None:
        with self._lock:
            self._thrust = np.full(4, float(thrust_per_motor))

    def build_horizon(self, horizon: int, dt: float) -> Dict[str, np.ndarray]:
        """Tile the set-point: (N+1) states, N thrusts (reference.py:75-91; dt unused there)."""
        with self._lock:
            n1 = horizon + 1
            return {
                'positions': np.tile(self._position, (n1, 1)),
                'velocities': np.tile(self._velocity, (n1, 1)),
                'quaternions': np.tile(self._quaternion, (n1, 1)),
                'body_rates': np.tile(self._body_rates, (n1, 1)),
                'thrusts': np.tile(self._thrust, (horizon, 1)),
                'yaws': np.full((n1,), self._yaw, dtype=float),
            }


def stack_reference(reference: Dict[str, np.ndarray], horizon: int
                    ) -> Tuple[np.ndarray, np.ndarray]:
    """Reference dict -> (yref[N,17], yref_e[13]) exactly as controller.py:433-445 stacks it."""
    N = horizon
    yref = np.empty((N, 17))
    yref[:, 0:3] = np.asarray(reference['positions'], dtype=float)[:N]
    yref[:, 3:6] = np.asarray(reference['velocities'], dtype=float)[:N]
    yref[:, 6:10] = np.asarray(reference['quaternions'], dtype=float)[:N]
    yref[:, 10:13] = np.asarray(reference['body_rates'], dtype=float)[:N]
    yref[:, 13:17] = np.asarray(reference['thrusts'], dtype=float)[:N]
    yref_e = np.concatenate([np.asarray(reference['positions'], dtype=float)[-1],
                             np.asarray(reference['velocities'], dtype=float)[-1],
                             np.asarray(reference['quaternions'], dtype=float)[-1],
                             np.asarray(reference['body_rates'], dtype=float)[-1]])
    return yref, yref_e


class BatchReference:
    """Batched reference builder producing ``yref[B,N,17]`` and ``yref_e[B,13]`` (numpy)."""

    def __init__(self, horizon: int, dt: float, hover_thrust: float) -> None:
        self.N, self.dt, self.hover = int(horizon), float(dt), float(hover_thrust)

    def setpoints(self, positions, yaws=None) -> Tuple[np.ndarray, np.ndarray]:
        """Constant set-points: what B ReferenceGenerators (+update_hover_thrust) would build."""
        p = np.asarray(positions, dtype=float).reshape(-1, 3)
        B = p.shape[0]
        q = yaw_quaternion(np.zeros(B) if yaws is None else np.asarray(yaws, dtype=float))
        yref = np.zeros((B, self.N, 17))
        yref[:, :, 0:3] = p[:, None, :]
        yref[:, :, 6:10] = q[:, None, :]
        yref[:, :, 13:17] = self.hover
        yref_e = np.zeros((B, 13))
        yref_e[:, 0:3] = p
        yref_e[:, 6:10] = q
        return yref, yref_e

    def figure8(self, t0, amplitude, period, centre=(0.0, 0.0, 1.5)
                ) -> Tuple[np.ndarray, np.ndarray]:
        """p(t) = c + (A sin wt, A/2 sin 2wt, 0) sampled at the stage times, with velocities."""
        t0 = np.asarray(t0, dtype=float).reshape(-1)
        A = np.broadcast_to(np.asarray(amplitude, dtype=float), t0.shape)
        w = 2.0 * np.pi / np.broadcast_to(np.asarray(period, dtype=float), t0.shape)
        B = t0.shape[0]
        t = t0[:, None] + self.dt * np.arange(self.N + 1)[None, :]
        wt = w[:, None] * t
        pos = np.zeros((B, self.N + 1, 3))
        vel = np.zeros((B, self.N + 1, 3))
        pos[:, :, 0] = centre[0] + A[:, None] * np.sin(wt)
        pos[:, :, 1] = centre[1] + 0.5 * A[:, None] * np.sin(2.0 * wt)
        pos[:, :, 2] = centre[2]
        vel[:, :, 0] = A[:, None] * w[:, None] * np.cos(wt)
        vel[:, :, 1] = A[:, None] * w[:, None] * np.cos(2.0 * wt)
        yref = np.zeros((B, self.N, 17))
        yref[:, :, 0:3] = pos[:, :self.N]
        yref[:, :, 3:6] = vel[:, :self.N]
        yref[:, :, 6] = 1.0
        yref[:, :, 13:17] = self.hover
        yref_e = np.zeros((B, 13))
        yref_e[:, 0:3] = pos[:, -1]
        yref_e[:, 3:6] = vel[:, -1]
        yref_e[:, 6] = 1.0
        return yref, yref_e
