"""Seeded synthetic problem generators shared by tests and bench."""
import numpy as np


def random_states(rng, B, tilt_deg=20.0, v_sigma=0.5, w_sigma=0.5):
    """cfg2 distribution: hover at (0,0,1) +- 0.5 m, random velocity / attitude / rates."""
    x0 = np.zeros((B, 13))
    x0[:, 0:3] = np.array([0.0, 0.0, 1.0]) + rng.uniform(-0.5, 0.5, (B, 3))
    x0[:, 3:6] = rng.normal(0.0, v_sigma, (B, 3))
    axis = rng.normal(size=(B, 3))
    axis /= np.linalg.norm(axis, axis=1, keepdims=True)
    ang = np.deg2rad(rng.uniform(0.0, tilt_deg, B))
    x0[:, 6] = np.cos(ang / 2)
    x0[:, 7:10] = np.sin(ang / 2)[:, None] * axis
    x0[:, 10:13] = rng.normal(0.0, w_sigma, (B, 3))
    return x0


def setpoint_refs(rng, x0, N, hover, span_xy=2.0, span_z=0.5, yaw=True):
    B = x0.shape[0]
    sp = x0[:, 0:3] + np.concatenate([rng.uniform(-span_xy, span_xy, (B, 2)),
                                      rng.uniform(-span_z, span_z, (B, 1))], axis=1)
    yw = rng.uniform(-np.pi, np.pi, B) if yaw else np.zeros(B)
    yref = np.zeros((B, N, 17))
    yref[:, :, 0:3] = sp[:, None, :]
    yref[:, :, 6] = np.cos(yw / 2)[:, None]
    yref[:, :, 9] = np.sin(yw / 2)[:, None]
    yref[:, :, 13:17] = hover
    yref_e = yref[:, 0, :13].copy()
    return yref, yref_e


def aggressive(rng, B, N, hover):
    """cfg4-like: large tilts / rates and 3-6 m set-point steps that saturate the thrust box."""
    x0 = random_states(rng, B, tilt_deg=60.0, v_sigma=1.5, w_sigma=2.0)
    yref, yref_e = setpoint_refs(rng, x0, N, hover, span_xy=6.0, span_z=6.0, yaw=False)
    return x0, yref, yref_e
