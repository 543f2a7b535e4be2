"""ctypes binding of ``libnmpc_b200.so`` (C ABI declared in ``include/nmpc_b200.h``).

This is the only way the Python host reaches the solver: there is no CPU fallback and no
alternative backend.  If the shared library is missing or no CUDA device is present the
functions here raise ``NmpcError`` / ``OSError`` -- loudly, by design.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

_PKG = Path(__file__).resolve().parent
# NMPC_B200_LIB points at another build of the SAME library (kernel tuning experiments)
LIB_PATH = Path(os.environ.get('NMPC_B200_LIB', _PKG / 'libnmpc_b200.so'))

NX, NU, NY = 13, 4, 17
PRECISION_F64, PRECISION_F32 = 0, 1

E_INVALID, E_CUDA, E_NOMEM, E_NODEVICE = -1, -2, -3, -4


class NmpcError(RuntimeError):
    def __init__(self, code: int, what: str, detail: str = ''):
        self.code = code
        super().__init__(f'{what}: error {code} ({_strerror(code)}){": " + detail if detail else ""}')


class NmpcConfig(C.Structure):
    """Mirror of ``nmpc_config_t``."""
    _fields_ = [
        ('N', C.c_int32), ('qp_iter_max', C.c_int32), ('integ_ns', C.c_int32),
        ('integ_steps', C.c_int32), ('cost_scale_dt', C.c_int32), ('lm_scaled_by_dt', C.c_int32),
        ('max_sqp_iter', C.c_int32), ('nan_status', C.c_int32), ('precision', C.c_int32),
        ('ipm_warm_start', C.c_int32), ('active_set_shortcut', C.c_int32), ('dispatch', C.c_int32),
        ('dt', C.c_double), ('W', C.c_double * 17), ('We', C.c_double * 13), ('lm', C.c_double),
        ('mass', C.c_double), ('J', C.c_double * 3), ('gravity', C.c_double),
        ('rotor_x', C.c_double * 4), ('rotor_y', C.c_double * 4), ('rotor_z', C.c_double * 4),
        ('lbu', C.c_double * 4), ('ubu', C.c_double * 4), ('kf', C.c_double),
        ('w_min', C.c_double), ('w_max', C.c_double), ('sqp_tol', C.c_double),
        ('ipm_tol', C.c_double),
    ]


class NmpcPlant(C.Structure):
    """Mirror of ``nmpc_plant_t``."""
    _fields_ = [('substeps', C.c_int32), ('motor_model', C.c_int32), ('dt', C.c_double),
                ('tau_up', C.c_double), ('tau_down', C.c_double), ('rotor_drag', C.c_double),
                ('mass', C.c_double), ('rolling_moment', C.c_double),
                ('rotors_yaw_sign', C.c_int32), ('reserved0', C.c_int32)]


class NmpcRollout(C.Structure):
    """Mirror of ``nmpc_rollout_t`` (all pointers are device addresses)."""
    _fields_ = [('steps', C.c_int32), ('ref_mode', C.c_int32), ('plant', NmpcPlant),
                ('x', C.c_void_p), ('rotor_speed', C.c_void_p), ('sp_pos', C.c_void_p),
                ('sp_yaw', C.c_void_p), ('f8_t0', C.c_void_p), ('f8_amplitude', C.c_void_p),
                ('f8_period', C.c_void_p), ('f8_centre', C.c_double * 3),
                ('t_start', C.c_double), ('stats', C.c_void_p)]


REF_SETPOINT, REF_FIGURE8, NSTATS = 0, 1, 8
STATIC_SCRATCH_BYTES, STATIC_MAX_BATCH = 40 << 20, 65536   # nmpc_solve_static limits

# every symbol include/nmpc_b200.h declares: (restype, argtypes)
_vp, _dp, _ip, _bp = C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p
SYMBOLS = {
    'nmpc_create': (C.c_int, [C.POINTER(NmpcConfig), C.c_int32, C.c_int32, C.POINTER(_vp)]),
    'nmpc_destroy': (C.c_int, [_vp]),
    'nmpc_reconfigure': (C.c_int, [_vp, C.POINTER(NmpcConfig)]),
    'nmpc_set_bounds': (C.c_int, [_vp, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    'nmpc_invalidate': (C.c_int, [_vp, _bp]),
    'nmpc_invalidate_device': (C.c_int, [_vp, _bp, _vp]),
    'nmpc_reset_last_command': (C.c_int, [_vp]),
    'nmpc_solve_device': (C.c_int, [_vp, _dp, _dp, _dp, _dp, _ip, _dp, _dp, _vp]),
    'nmpc_solve_host': (C.c_int, [_vp, _dp, _dp, _dp, _dp, _ip, _dp, _dp]),
    # allocation-free variant (scratch in a static device array): cfg, batch, device, x0, yref, yref_e,
    # xtraj, utraj, valid, last_cmd, u0, status, thrust, speed, qp_iters, sqp_iters, stream
    'nmpc_solve_static': (C.c_int, [C.POINTER(NmpcConfig), C.c_int32, C.c_int32, _dp, _dp, _dp, _dp,
                                    _dp, _bp, _dp, _dp, _ip, _dp, _dp, _ip, _ip, _vp]),
    'nmpc_get_trajectories': (C.c_int, [_vp, _dp, _dp, _bp, C.c_int32]),
    'nmpc_get_stats': (C.c_int, [_vp, _ip, _ip, C.c_int32]),
    'nmpc_set_trajectories': (C.c_int, [_vp, _dp, _dp, _bp]),
    'nmpc_discretize_device': (C.c_int, [_vp, C.c_int32, _dp, _dp, _dp, _dp, _dp, _vp]),
    'nmpc_reference_setpoint_device': (C.c_int, [_vp, _dp, _dp, _dp, _dp, _vp]),
    'nmpc_reference_figure8_device': (C.c_int, [_vp, _dp, _dp, _dp, C.POINTER(C.c_double),
                                                C.c_double, _dp, _dp, _vp]),
    'nmpc_odometry_to_state_device': (C.c_int, [_vp, _dp, _dp, _ip, _vp]),
    'nmpc_plant_step_device': (C.c_int, [_vp, C.POINTER(NmpcPlant), _dp, _dp, _dp, _dp, _vp]),
    'nmpc_rollout_device': (C.c_int, [_vp, C.POINTER(NmpcRollout), _vp]),
    'nmpc_last_outputs': (C.c_int, [_vp, C.POINTER(_vp), C.POINTER(_vp), C.POINTER(_vp),
                                    C.POINTER(_vp)]),
    'nmpc_fp64_peak': (C.c_int, [C.c_int32, C.POINTER(C.c_double)]),
    'nmpc_fp32_peak': (C.c_int, [C.c_int32, C.POINTER(C.c_double)]),
    'nmpc_debug_poison': (C.c_int, [_vp, C.c_int32]),
    'nmpc_batch': (C.c_int, [_vp]),
    'nmpc_horizon': (C.c_int, [_vp]),
    'nmpc_workspace_bytes': (C.c_int64, [_vp]),
    'nmpc_kernel_launches': (C.c_int, [_vp]),
    'nmpc_strerror': (C.c_char_p, [C.c_int]),
    'nmpc_last_error': (C.c_char_p, []),
    'nmpc_version': (C.c_char_p, []),
}

_lib = None


def lib():
    """Load the shared library (once).  Raises OSError if it has not been built."""
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            raise OSError(f'{LIB_PATH} not found: build it with `python -c "import __graft_entry__ '
                          f'as g; g.build()"` or `make -C rotors_mpc_controller_b200/csrc` '
                          f'(there is no CPU fallback)')
        handle = C.CDLL(str(LIB_PATH))
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(handle, name)
            fn.restype, fn.argtypes = res, args
        _lib = handle
    return _lib


def _strerror(code: int) -> str:
    try:
        return lib().nmpc_strerror(code).decode()
    except Exception:  # pragma: no cover
        return '?'


def check(code: int, what: str) -> None:
    if code != 0:
        raise NmpcError(code, what, lib().nmpc_last_error().decode())


def ptr(arr) -> int:
    """Raw address of a numpy array or torch tensor (or None)."""
    if arr is None:
        return None
    if hasattr(arr, 'data_ptr'):
        return arr.data_ptr()
    return arr.ctypes.data
