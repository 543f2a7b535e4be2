"""Host-side logic (no GPU): params loading / coercion, dynamic reconfigure mapping, reference
generator, controller constants, C-ABI library symbols and struct layout."""
import copy
import ctypes as C
import os
import re
import subprocess
from pathlib import Path
from types import SimpleNamespace

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent


# ---------------------------------------------------------------- params.py
def test_load_params_default_and_env(tmp_path, monkeypatch):
    from rotors_mpc_controller_b200 import load_params
    monkeypatch.delenv('ROTORS_MPC_PARAMS', raising=False)
    p = load_params()
    assert p['solver']['horizon_steps'] == 20 and p['solver']['dt'] == 0.05
    assert p['vehicle']['motor_max_speed'] == 838.0 and p['controller']['thrust_limits'] == [4.0, 20.0]
    assert p['vehicle']['drag_coefficients'] == [0.0, 0.0, 0.0]      # default filled in
    assert p['params_yaml'].endswith('config/params.yaml')
    import yaml
    q = copy.deepcopy(p)
    q.pop('params_yaml')
    q['solver']['horizon_steps'] = '35'          # strings are coerced
    q['node']['yaw_rate_gain'] = 1.0             # stale key dropped (reference params.py:147-149)
    f = tmp_path / 'alt.yaml'
    f.write_text(yaml.safe_dump(q))
    monkeypatch.setenv('ROTORS_MPC_PARAMS', str(f))
    p2 = load_params()
    assert p2['solver']['horizon_steps'] == 35 and 'yaw_rate_gain' not in p2['node']
    assert p2['params_yaml'] == str(f)


def test_params_errors(params):
    from rotors_mpc_controller_b200.params import coerce_params
    bad = copy.deepcopy(params)
    del bad['topics']
    with pytest.raises(ValueError, match='Missing required top-level sections'):
        coerce_params(bad)
    bad = copy.deepcopy(params)
    bad['vehicle']['inertia'] = [1.0, 2.0]
    with pytest.raises(ValueError, match='inertia'):
        coerce_params(bad)
    bad = copy.deepcopy(params)
    bad['controller']['thrust_limits'] = [1.0]
    with pytest.raises(ValueError, match='thrust_limits'):
        coerce_params(bad)
    bad = copy.deepcopy(params)
    del bad['topics']['motor']
    with pytest.raises(ValueError, match="'motor'"):
        coerce_params(bad)


def _reconfigure_object(params):
    """A dynamic_reconfigure-like config carrying every field of cfg/rotors_mpc.cfg."""
    s, v = params['solver'], params['vehicle']
    d = dict(solver_horizon_steps=40, solver_dt=0.02, solver_regularization=1e-3,
             solver_iter_max=100, solver_codegen_directory='/tmp/x',
             vehicle_mass=0.9, vehicle_arm_length=0.2, vehicle_rotor_force_constant=9e-6,
             vehicle_rotor_moment_constant=0.02, vehicle_motor_min_speed=10.0,
             vehicle_motor_max_speed=900.0, world_gravity=9.8, reference_frame='map',
             reference_yaw=0.3, node_rate=100.0, node_log_interval=1.0,
             controller_thrust_min=1.0, controller_thrust_max=10.0,
             topic_state='/a', topic_motor='/b', topic_reference='/c')
    for i, n in enumerate('xyz'):
        d[f'solver_position_weight_{n}'] = 1.0 + i
        d[f'solver_velocity_weight_{n}'] = 2.0 + i
        d[f'solver_rate_weight_{n}'] = 3.0 + i
        d[f'vehicle_drag_{n}'] = 0.1 * i
        d[f'reference_position_{n}'] = float(i)
        d[f'reference_velocity_{n}'] = -float(i)
    for i, n in enumerate('wxyz'):
        d[f'solver_quat_weight_{n}'] = 4.0 + i
    for i in range(4):
        d[f'solver_control_weight_f{i + 1}'] = 5.0 + i
    for i, n in enumerate(['px', 'py', 'pz', 'vx', 'vy', 'vz', 'qw', 'qx', 'qy', 'qz', 'wx', 'wy', 'wz']):
        d[f'solver_terminal_weight_{n}'] = 10.0 + i
    for i, n in enumerate(['xx', 'xy', 'xz', 'yx', 'yy', 'yz', 'zx', 'zy', 'zz']):
        d[f'vehicle_inertia_{n}'] = 0.01 * (i + 1)
    return d


def test_apply_dynamic_configuration(params):
    from rotors_mpc_controller_b200 import apply_dynamic_configuration
    p = copy.deepcopy(params)
    d = _reconfigure_object(p)
    topics, ref_defaults, node_meta = apply_dynamic_configuration(p, SimpleNamespace(**d))
    assert p['solver']['horizon_steps'] == 40 and p['solver']['dt'] == 0.02
    assert p['solver']['position_weight'] == [1.0, 2.0, 3.0]
    assert p['solver']['quaternion_weight'] == [4.0, 5.0, 6.0, 7.0]
    assert p['solver']['terminal_weight'] == [10.0 + i for i in range(13)]
    assert p['vehicle']['inertia'][4] == 0.05 and p['vehicle']['mass'] == 0.9
    assert p['controller']['thrust_limits'] == [1.0, 10.0]
    assert topics == {'state': '/a', 'motor': '/b', 'reference': '/c'} and p['topics'] == topics
    assert ref_defaults == {'position': [0.0, 1.0, 2.0], 'velocity': [-0.0, -1.0, -2.0],
                            'yaw': 0.3, 'frame': 'map'}
    assert node_meta == {'rate': 100.0, 'log_interval': 1.0}
    # optional fields fall back to the current values (reference uses getattr defaults)
    p = copy.deepcopy(params)
    for k in [k for k in d if k.startswith('solver_quat_weight') or k == 'solver_iter_max']:
        d.pop(k)
    apply_dynamic_configuration(p, SimpleNamespace(**d))
    assert p['solver']['quaternion_weight'] == params['solver']['quaternion_weight']
    assert p['solver']['iter_max'] == params['solver']['iter_max']
    # required fields raise when absent
    d.pop('solver_dt')
    with pytest.raises(AttributeError):
        apply_dynamic_configuration(copy.deepcopy(params), SimpleNamespace(**d))


# ---------------------------------------------------------------- reference.py
def test_reference_generator(params):
    from rotors_mpc_controller_b200 import ReferenceGenerator, stack_reference
    ref = ReferenceGenerator(params['reference'])
    h = ref.build_horizon(20, 0.05)
    assert h['positions'].shape == (21, 3) and h['thrusts'].shape == (20, 4)
    assert np.all(h['thrusts'] == 0.0)               # bare generator: zero thrust (reference.py:28)
    assert np.allclose(h['quaternions'], [1, 0, 0, 0]) and h['yaws'].shape == (21,)
    ref.update_hover_thrust(1.6677)
    ref.set_target(position=[2, 0, 1.5], yaw=np.pi / 2)
    h = ref.build_horizon(20, 0.05)
    assert np.allclose(h['quaternions'][0], [np.cos(np.pi / 4), 0, 0, np.sin(np.pi / 4)])
    yref, yref_e = stack_reference(h, 20)
    assert yref.shape == (20, 17) and yref_e.shape == (13,)
    assert np.allclose(yref[3], np.r_[[2, 0, 1.5], [0, 0, 0], h['quaternions'][0], [0, 0, 0],
                                      [1.6677] * 4])
    assert np.allclose(yref_e, yref[0, :13])
    ref.set_target(position=[0, 0, 1], quaternion=[2.0, 0, 0, 0])
    assert np.allclose(ref.build_horizon(2, 0.05)['quaternions'][0], [1, 0, 0, 0])
    with pytest.raises(ValueError):
        ref.set_target(position=[0, 0, 1], thrust=[1.0, 2.0])
    ref.update_defaults([1, 1, 1], [0, 0, 0], 0.0, frame='odom')
    assert ref.frame == 'odom'


def test_batch_reference():
    from rotors_mpc_controller_b200 import BatchReference
    br = BatchReference(20, 0.05, 1.6677)
    yref, yref_e = br.setpoints([[1, 2, 3], [4, 5, 6]], yaws=[0.0, np.pi])
    assert yref.shape == (2, 20, 17) and np.allclose(yref[1, 7, :3], [4, 5, 6])
    assert np.allclose(yref[1, 0, 6:10], [np.cos(np.pi / 2), 0, 0, 1.0], atol=1e-12)
    assert np.allclose(yref[:, :, 13:], 1.6677) and np.allclose(yref_e[0, :3], [1, 2, 3])
    yref, yref_e = br.figure8(t0=[0.0, 1.0], amplitude=2.0, period=4.0)
    w = 2 * np.pi / 4.0
    assert np.allclose(yref[0, 5, 0], 2.0 * np.sin(w * 0.25))
    assert np.allclose(yref[0, 5, 3], 2.0 * w * np.cos(w * 0.25))
    assert np.allclose(yref_e[1, 1], 1.0 * np.sin(2 * w * 2.0))


# ---------------------------------------------------------------- controller constants
def test_controller_params_follow_reference(params):
    from rotors_mpc_controller_b200.controller import (abi_config, config_signature,
                                                       controller_params_from_dict)
    cp = controller_params_from_dict(params)
    assert cp.horizon_steps == 20 and cp.solver_iter_max == 600
    assert np.allclose(cp.input_lower_bounds, 8.54858e-6 * 50 ** 2)      # NOT 4.0
    assert np.allclose(cp.input_upper_bounds, 8.54858e-6 * 838 ** 2)     # NOT 20.0
    assert np.allclose(cp.rotor_z_torque, [-0.016, 0.016, -0.016, 0.016])
    assert np.allclose(cp.rotor_x_offsets, [0.17, 0, -0.17, 0])
    assert np.allclose(cp.inertia, [0.007, 0.007, 0.012])
    assert abs(cp.hover_thrust_per_motor - 0.68 * 9.81 / 4) < 1e-15
    c = abi_config(cp, params['solver'])
    assert (c.N, c.integ_ns, c.integ_steps, c.max_sqp_iter, c.cost_scale_dt) == (20, 2, 2, 1, 1)
    assert list(c.W)[:3] == [10.0, 10.0, 8.0] and list(c.W)[13:] == [1.75] * 4
    assert c.lm == 7e-3 and c.qp_iter_max == 600 and c.precision == 0
    p = copy.deepcopy(params)
    p['solver']['use_thrust_limits'] = True
    assert np.allclose(controller_params_from_dict(p).input_lower_bounds, 4.0)
    p = copy.deepcopy(params)
    p['vehicle']['rotor_configuration'] = 'x'
    with pytest.raises(ValueError, match='"\\+" rotor lay-out'):
        controller_params_from_dict(p)
    p = copy.deepcopy(params)
    del p['vehicle']['mass']
    with pytest.raises(KeyError):
        controller_params_from_dict(p)
    p = copy.deepcopy(params)
    p['solver']['integrator'] = 'bogus'
    with pytest.raises(ValueError):
        abi_config(cp, p['solver'])
    s1 = config_signature(cp)
    p = copy.deepcopy(params)
    p['solver']['dt'] = 0.04
    assert s1 != config_signature(controller_params_from_dict(p)) and len(s1) == 10


# ---------------------------------------------------------------- C ABI
def test_abi_exports_every_declared_symbol(built):
    from rotors_mpc_controller_b200 import _abi
    header = (ROOT / 'include' / 'nmpc_b200.h').read_text()
    declared = set(re.findall(r'\b(nmpc_[a-z0-9_]+)\s*\(', header))
    assert declared == set(_abi.SYMBOLS), declared ^ set(_abi.SYMBOLS)
    lib = C.CDLL(str(_abi.LIB_PATH))
    for name in declared:
        assert hasattr(lib, name), name
    assert b'sm_100a' in _abi.lib().nmpc_version()
    assert _abi.lib().nmpc_strerror(-4).decode().startswith('no CUDA device')


def test_abi_struct_layout_matches_header(tmp_path):
    from rotors_mpc_controller_b200 import _abi
    src = tmp_path / 'sz.c'
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "nmpc_b200.h"\n'
                   'int main(){printf("%zu %zu %zu %zu %zu", sizeof(nmpc_config_t),'
                   'offsetof(nmpc_config_t, dt), offsetof(nmpc_config_t, We),'
                   'offsetof(nmpc_config_t, lbu), offsetof(nmpc_config_t, ipm_tol));return 0;}')
    exe = tmp_path / 'sz'
    subprocess.run(['gcc', '-I', str(ROOT / 'include'), str(src), '-o', str(exe)], check=True)
    got = [int(v) for v in subprocess.run([str(exe)], capture_output=True, text=True).stdout.split()]
    K = _abi.NmpcConfig
    assert got == [C.sizeof(K), K.dt.offset, K.We.offset, K.lbu.offset, K.ipm_tol.offset]


def test_no_cpu_fallback(built, params):
    """Without a CUDA device the product path must fail loudly (never route to the oracle)."""
    import torch
    if torch.cuda.is_available():
        pytest.skip('GPU present')
    from rotors_mpc_controller_b200 import PositionNMPC, _abi
    with pytest.raises(_abi.NmpcError) as ei:
        PositionNMPC(params)
    assert ei.value.code == _abi.E_NODEVICE
    src = ''.join(p.read_text() for p in (ROOT / 'rotors_mpc_controller_b200').glob('*.py'))
    assert 'oracle' not in src.replace('the oracle', '')   # product never imports oracle/


def test_slider_ranges_cover_every_numeric_cfg_field(params):
    """DYNAMIC_RANGES mirrors cfg/rotors_mpc.cfg:8-92 (67 numeric sliders) and clamping acts
    like the dynamic_reconfigure server: out-of-range values are pulled to the edge."""
    import copy
    from types import SimpleNamespace
    from rotors_mpc_controller_b200 import (DYNAMIC_RANGES, apply_dynamic_configuration,
                                            clamp_dynamic_configuration)
    assert len(DYNAMIC_RANGES) == 67
    assert DYNAMIC_RANGES['solver_horizon_steps'] == (1, 600)
    assert DYNAMIC_RANGES['solver_terminal_weight_qz'] == (0.0, 50.0)
    assert DYNAMIC_RANGES['vehicle_motor_max_speed'] == (100.0, 2000.0)
    cfg = SimpleNamespace(solver_horizon_steps=900, solver_dt=0.001, vehicle_mass=0.68,
                          solver_position_weight_x=75.0, node_rate=60.0, topic_state='/odom')
    changed = clamp_dynamic_configuration(cfg)
    assert sorted(changed) == ['solver_dt', 'solver_horizon_steps', 'solver_position_weight_x']
    assert cfg.solver_horizon_steps == 600 and isinstance(cfg.solver_horizon_steps, int)
    assert cfg.solver_dt == 0.01 and cfg.solver_position_weight_x == 50.0
    assert cfg.vehicle_mass == 0.68 and cfg.topic_state == '/odom'
