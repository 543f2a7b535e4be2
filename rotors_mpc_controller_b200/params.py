"""Configuration loading with the semantics of the reference's
``/root/reference/src/rotors_mpc_controller/params.py`` (``load_params`` :153-183,
``apply_dynamic_configuration`` :186-294), ROS optional.

Search order of ``load_params`` (reference :34-58): ``$ROTORS_MPC_PARAMS``, the package default
``config/params.yaml``, then the ROS package path when ``rospkg`` is importable.  Every section
is coerced to plain floats / ints / strings with the reference's defaults, and the seven
top-level sections are required (reference :123-127).
"""
from __future__ import annotations

import os
from pathlib import Path
from typing import Any, Dict, List, Tuple

REQUIRED_SECTIONS = ('solver', 'vehicle', 'controller', 'world', 'reference', 'topics', 'node')

_TERMINAL_DEFAULT = [5.0, 5.0, 3.0, 2.0, 2.0, 2.0, 12.0, 12.0, 12.0, 18.5, 2.0, 2.0, 1.8]

# section -> [(key, kind, default)]; kind: 'i' int, 'f' float, 's' str, ('v', n) float vector
_SCHEMA = {
    'solver': [
        ('horizon_steps', 'i', 20), ('dt', 'f', 0.05),
        ('position_weight', ('v', 3), [10.0, 10.0, 8.0]),
        ('velocity_weight', ('v', 3), [1.0, 1.0, 0.2]),
        ('quaternion_weight', ('v', 4), [3.2, 3.2, 3.2, 3.2]),
        ('rate_weight', ('v', 3), [1.4, 1.4, 0.4]),
        ('control_weight', ('v', 4), [1.75, 1.75, 1.75, 1.75]),
        ('terminal_weight', ('v', 13), _TERMINAL_DEFAULT),
        ('regularization', 'f', 7.0e-3), ('iter_max', 'i', 600),
    ],
    'vehicle': [
        ('mass', 'f', 0.68),
        ('inertia', ('v', 9), [0.007, 0.0, 0.0, 0.0, 0.007, 0.0, 0.0, 0.0, 0.012]),
        ('arm_length', 'f', 0.17), ('rotor_force_constant', 'f', 8.54858e-6),
        ('rotor_moment_constant', 'f', 0.016), ('motor_min_speed', 'f', 0.0),
        ('motor_max_speed', 'f', 2000.0), ('drag_coefficients', ('v', 3), [0.0, 0.0, 0.0]),
        ('rotor_configuration', 's', '+'),
    ],
    'controller': [('thrust_limits', ('v', 2), [4.0, 20.0])],
    'world': [('gravity', 'f', 9.81)],
    'reference': [
        ('frame', 'raw', 'world'), ('default_position', ('v', 3), [1.0, 1.0, 1.0]),
        ('default_velocity', ('v', 3), [0.0, 0.0, 0.0]),
        ('default_acceleration', ('v', 3), [0.0, 0.0, 0.0]), ('default_yaw', 'f', 0.0),
    ],
    'node': [('rate', 'f', 50.0), ('log_interval', 'f', 3.0)],
}
_LENGTH_MESSAGES = {
    ('vehicle', 'inertia'): 'vehicle.inertia must contain 9 values (row-major 3x3).',
    ('vehicle', 'drag_coefficients'): 'vehicle.drag_coefficients must contain 3 values.',
    ('controller', 'thrust_limits'): 'controller.thrust_limits must contain [min, max].',
}
# stale keys the reference silently drops (:114-116, :147-149)
_DROPPED = {'controller': ('attitude_gains', 'max_tilt_deg', 'max_tilt_angle'),
            'node': ('max_tilt_deg', 'yaw_rate_gain', 'yaw_rate_limit')}


def _read_yaml(path: Path) -> Dict[str, Any]:
    try:
        import yaml
    except Exception as exc:  # pragma: no cover
        raise RuntimeError('PyYAML is required to load controller parameters.') from exc
    if not path.is_file():
        raise FileNotFoundError(f'Parameter file not found: {path}')
    with path.open('r', encoding='utf-8') as fh:
        data = yaml.safe_load(fh) or {}
    if not isinstance(data, dict):
        raise ValueError(f'Root of {path} must be a mapping.')
    return data


def default_params_path() -> Path:
    return Path(__file__).resolve().parent / 'config' / 'params.yaml'


def _search_paths() -> List[Path]:
    found: List[Path] = []
    env = os.environ.get('ROTORS_MPC_PARAMS')
    if env:
        found.append(Path(env).expanduser())
    found.append(default_params_path())
    try:
        import rospkg  # type: ignore
        found.append(Path(rospkg.RosPack().get_path('rotors_mpc_controller')) / 'config'
                     / 'params.yaml')
    except Exception:
        pass
    return list(dict.fromkeys(found))


def merge_overrides(base: Dict[str, Any], extra: Dict[str, Any]) -> Dict[str, Any]:
    """Recursive dict update (reference :61-67): nested mappings merge, leaves replace."""
    for key, val in extra.items():
        if isinstance(val, dict) and isinstance(base.get(key), dict):
            merge_overrides(base[key], val)
        else:
            base[key] = val
    return base


def _coerce_section(name: str, cfg: Dict[str, Any]) -> None:
    for key, kind, default in _SCHEMA[name]:
        val = cfg.get(key, default)
        if kind == 'i':
            cfg[key] = int(val)
        elif kind == 'f':
            cfg[key] = float(val)
        elif kind == 's':
            cfg[key] = str(val).strip()
        elif kind == 'raw':
            cfg[key] = val
        else:
            n = kind[1]
            if (name, key) in _LENGTH_MESSAGES and len(val) != n:
                raise ValueError(_LENGTH_MESSAGES[(name, key)])
            cfg[key] = [float(v) for v in val]
    for stale in _DROPPED.get(name, ()):
        cfg.pop(stale, None)
    if name == 'solver' and 'codegen_directory' in cfg:
        cfg['codegen_directory'] = str(Path(cfg['codegen_directory']).expanduser())


def coerce_params(cfg: Dict[str, Any]) -> Dict[str, Any]:
    """Validate + coerce a raw parameter dict in place (reference :70-150)."""
    missing = set(REQUIRED_SECTIONS) - cfg.keys()
    if missing:
        raise ValueError(f'Missing required top-level sections: {sorted(missing)}')
    for section in _SCHEMA:
        _coerce_section(section, cfg[section])
    topics = cfg['topics']
    for key in ('state', 'motor', 'reference'):
        if key not in topics:
            raise ValueError(f"Missing topic configuration '{key}'")
        topics[key] = str(topics[key])
    return cfg


def load_params() -> Dict[str, Any]:
    """Load, override (ROS private params when a node is initialised) and coerce."""
    cfg = None
    for cand in _search_paths():
        if cand.is_file():
            cfg = _read_yaml(cand)
            cfg['params_yaml'] = str(cand)
            break
    if cfg is None:
        raise FileNotFoundError('No configuration file found for rotors_mpc_controller.')
    try:
        import rospy  # type: ignore
        if rospy.core.is_initialized():
            extra = rospy.get_param('~')
            if isinstance(extra, dict):
                merge_overrides(cfg, extra)
    except Exception:
        pass
    return coerce_params(cfg)


# dynamic_reconfigure field names (cfg/rotors_mpc.cfg) -> (section, key, component suffixes)
_VECTOR_FIELDS = [
    ('solver', 'position_weight', 'solver_position_weight_', ('x', 'y', 'z'), True),
    ('solver', 'velocity_weight', 'solver_velocity_weight_', ('x', 'y', 'z'), True),
    ('solver', 'quaternion_weight', 'solver_quat_weight_', ('w', 'x', 'y', 'z'), False),
    ('solver', 'rate_weight', 'solver_rate_weight_', ('x', 'y', 'z'), False),
    ('solver', 'control_weight', 'solver_control_weight_', ('f1', 'f2', 'f3', 'f4'), False),
    ('solver', 'terminal_weight', 'solver_terminal_weight_',
     ('px', 'py', 'pz', 'vx', 'vy', 'vz', 'qw', 'qx', 'qy', 'qz', 'wx', 'wy', 'wz'), False),
    ('vehicle', 'inertia', 'vehicle_inertia_',
     ('xx', 'xy', 'xz', 'yx', 'yy', 'yz', 'zx', 'zy', 'zz'), True),
    ('vehicle', 'drag_coefficients', 'vehicle_drag_', ('x', 'y', 'z'), True),
    ('reference', 'default_position', 'reference_position_', ('x', 'y', 'z'), True),
    ('reference', 'default_velocity', 'reference_velocity_', ('x', 'y', 'z'), True),
]
_SCALAR_FIELDS = [
    ('solver', 'horizon_steps', 'solver_horizon_steps', int, True),
    ('solver', 'dt', 'solver_dt', float, True),
    ('solver', 'regularization', 'solver_regularization', float, True),
    ('solver', 'iter_max', 'solver_iter_max', int, False),
    ('solver', 'codegen_directory', 'solver_codegen_directory', str, True),
    ('vehicle', 'mass', 'vehicle_mass', float, True),
    ('vehicle', 'arm_length', 'vehicle_arm_length', float, True),
    ('vehicle', 'rotor_force_constant', 'vehicle_rotor_force_constant', float, True),
    ('vehicle', 'rotor_moment_constant', 'vehicle_rotor_moment_constant', float, True),
    ('vehicle', 'motor_min_speed', 'vehicle_motor_min_speed', float, True),
    ('vehicle', 'motor_max_speed', 'vehicle_motor_max_speed', float, True),
    ('world', 'gravity', 'world_gravity', float, True),
    ('reference', 'frame', 'reference_frame', str, True),
    ('reference', 'default_yaw', 'reference_yaw', float, True),
    ('node', 'rate', 'node_rate', float, True),
    ('node', 'log_interval', 'node_log_interval', float, True),
]


def apply_dynamic_configuration(params: Dict[str, Any], config: Any
                                ) -> Tuple[Dict[str, str], Dict[str, list], Dict[str, float]]:
    """Fold a dynamic_reconfigure ``config`` object into ``params`` (reference :186-294).

    Fields the reference reads with ``getattr(config, name, current)`` stay optional here
    (``required=False``); the others raise ``AttributeError`` when absent, as there."""
    for section, key, prefix, parts, required in _VECTOR_FIELDS:
        current = params[section].get(key)
        out = []
        for i, part in enumerate(parts):
            name = prefix + part
            if required:
                out.append(float(getattr(config, name)))
            else:
                out.append(float(getattr(config, name, current[i])))
        params[section][key] = out
    for section, key, name, conv, required in _SCALAR_FIELDS:
        if required:
            params[section][key] = conv(getattr(config, name))
        else:
            params[section][key] = conv(getattr(config, name, params[section][key]))
    params['controller']['thrust_limits'] = [float(config.controller_thrust_min),
                                             float(config.controller_thrust_max)]
    topics = {'state': str(config.topic_state), 'motor': str(config.topic_motor),
              'reference': str(config.topic_reference)}
    params['topics'] = topics
    ref = params['reference']
    reference_defaults = {'position': ref['default_position'], 'velocity': ref['default_velocity'],
                          'yaw': ref['default_yaw'], 'frame': ref['frame']}
    node_meta = {'rate': params['node']['rate'], 'log_interval': params['node']['log_interval']}
    return topics, reference_defaults, node_meta


# ------------------------------------------------------------------ slider ranges (cfg/rotors_mpc.cfg:8-92)
def _ranges() -> Dict[str, Tuple[float, float]]:
    r: Dict[str, Tuple[float, float]] = {
        'solver_horizon_steps': (1, 600), 'solver_dt': (0.01, 0.2),
        'solver_regularization': (1e-6, 0.1), 'solver_iter_max': (1, 600),
        'vehicle_mass': (0.2, 5.0), 'vehicle_arm_length': (0.05, 1.0),
        'vehicle_rotor_force_constant': (1e-7, 1e-4),
        'vehicle_rotor_moment_constant': (1e-3, 0.1),
        'vehicle_motor_min_speed': (0.0, 500.0), 'vehicle_motor_max_speed': (100.0, 2000.0),
        'controller_thrust_min': (0.0, 20.0), 'controller_thrust_max': (0.0, 40.0),
        'world_gravity': (5.0, 15.0), 'reference_position_z': (0.0, 10.0),
        'reference_yaw': (-3.1416, 3.1416), 'node_rate': (10.0, 200.0),
        'node_log_interval': (0.1, 10.0),
    }
    for a in 'xyz':
        r[f'solver_position_weight_{a}'] = (0.0, 50.0)
        r[f'solver_velocity_weight_{a}'] = (0.0, 10.0)
        r[f'solver_rate_weight_{a}'] = (0.0, 20.0)
        r[f'solver_terminal_weight_p{a}'] = (0.0, 50.0)
        r[f'solver_terminal_weight_v{a}'] = (0.0, 20.0)
        r[f'solver_terminal_weight_w{a}'] = (0.0, 20.0)
        r[f'vehicle_drag_{a}'] = (-0.5, 0.5)
        r[f'reference_velocity_{a}'] = (-5.0, 5.0)
        r[f'reference_acceleration_{a}'] = (-10.0, 10.0)
    for a in 'xy':
        r[f'reference_position_{a}'] = (-10.0, 10.0)
    for a in 'wxyz':
        r[f'solver_quat_weight_{a}'] = (0.0, 20.0)
        r[f'solver_terminal_weight_q{a}'] = (0.0, 50.0)
    for i in '1234':
        r[f'solver_control_weight_f{i}'] = (0.0, 5.0)
    for a in ('xx', 'yy', 'zz'):
        r[f'vehicle_inertia_{a}'] = (0.0, 0.1)
    for a in ('xy', 'xz', 'yx', 'yz', 'zx', 'zy'):
        r[f'vehicle_inertia_{a}'] = (-0.05, 0.05)
    return r


DYNAMIC_RANGES = _ranges()


def clamp_dynamic_configuration(config: Any) -> List[str]:
    """Clamp the numeric fields of a dynamic_reconfigure-style ``config`` object to the slider
    ranges of ``cfg/rotors_mpc.cfg:8-92`` in place -- what the dynamic_reconfigure server does
    before ``_reconfigure_cb`` (``nodes/mpc_controller_node:166-198``) sees the values.  Returns
    the names that were changed.  The reference then rebuilds and recompiles the acados solver
    on every event (:196-198); here ``BatchedPositionNMPC.reconfigure`` swaps constants in the
    existing handle (``nmpc_reconfigure``) unless the horizon length changed."""
    changed = []
    for name, (lo, hi) in DYNAMIC_RANGES.items():
        if not hasattr(config, name):
            continue
        v = getattr(config, name)
        c = min(max(v, type(v)(lo)), type(v)(hi))
        if c != v:
            setattr(config, name, c)
            changed.append(name)
    return changed
