"""Generates tests/golden/nmpc_golden.npz with the CPU oracle.

The reference ships no golden vectors and cannot run here (acados absent), so these fixtures
pin the ORACLE's output (flag set below) -- they guard against regressions of oracle and CUDA
path alike, they do not pin parity with acados ("parity unpinned").

    python tests/golden/make_golden.py
"""
import json
import sys
from pathlib import Path

import numpy as np
import yaml

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
from oracle import oracle as O  # noqa: E402
from tests.problems import aggressive, random_states, setpoint_refs  # noqa: E402

FLAGS = dict(integ_ns=2, integ_steps=2, cost_scale_dt=1, lm_scaled_by_dt=1, max_sqp_iter=1,
             ipm_tol=1e-14, nan_status=4, bounds='kf*w^2 -> [0.02137145, 6.00318901] N')


def main():
    params = yaml.safe_load(open(ROOT / 'rotors_mpc_controller_b200' / 'config' / 'params.yaml'))
    cfg = O.config_from_params(params)
    hover = cfg.mass * cfg.gravity / 4.0
    out = {}
    # discretisation (midpoint x2 = reference setting, and RK4 x1)
    rng = np.random.default_rng(100)
    x = rng.normal(size=(6, 13))
    u = rng.uniform(0, 6, (6, 4))
    out['disc_x'], out['disc_u'] = x, u
    for tag, kw in (('mid', {}), ('rk4', dict(integ_ns=4, integ_steps=1))):
        c = O.config_from_params(params, **kw)
        r = [O.discretize(c, x[i], u[i]) for i in range(6)]
        out[f'disc_{tag}_xn'] = np.stack([a[0] for a in r])
        out[f'disc_{tag}_A'] = np.stack([a[1] for a in r])
        out[f'disc_{tag}_B'] = np.stack([a[2] for a in r])
    # cfg1: hover (0,0,1) -> set-point (2,0,1.5), closed loop with the nominal plant, 30 steps
    N = 20
    ob = O.OracleBatch(cfg, 1)
    xs = np.array([[0, 0, 1, 0, 0, 0, 1, 0, 0, 0, 0, 0, 0.0]])
    yref = np.zeros((1, N, 17))
    yref[:, :, :3] = [2.0, 0.0, 1.5]
    yref[:, :, 6] = 1.0
    yref[:, :, 13:] = hover
    yref_e = yref[:, 0, :13].copy()
    us, xs_log = [], []
    for k in range(30):
        xs_log.append(xs[0].copy())
        u0, st, th, sp = ob.solve(xs, yref, yref_e)
        assert st[0] == 0
        us.append(u0[0].copy())
        xs = O.plant_step(cfg, xs[0], th[0], 0.05, 5)[None]
    out['cfg1_x'], out['cfg1_u0'] = np.array(xs_log), np.array(us)
    out['cfg1_yref'], out['cfg1_yref_e'] = yref[0], yref_e[0]
    # cfg2-like batch: cold RTI + one warm RTI
    B = 16
    rng = np.random.default_rng(101)
    x0 = random_states(rng, B)
    yr, ye = setpoint_refs(rng, x0, N, hover)
    ob = O.OracleBatch(cfg, B)
    u_c, s_c, _, _ = ob.solve(x0, yr, ye)
    xt_c = ob.xtraj.copy()
    u_w, s_w, th_w, sp_w = ob.solve(x0, yr, ye)
    out.update(b_x0=x0, b_yref=yr, b_yref_e=ye, b_u0_cold=u_c, b_status_cold=s_c, b_xtraj_cold=xt_c,
               b_u0_warm=u_w, b_status_warm=s_w, b_thrust_warm=th_w, b_speed_warm=sp_w)
    # cfg4-like aggressive batch (bounds active)
    rng = np.random.default_rng(102)
    x0, yr, ye = aggressive(rng, B, N, hover)
    ob = O.OracleBatch(cfg, B)
    u_a, s_a, _, _ = ob.solve(x0, yr, ye)
    out.update(a_x0=x0, a_yref=yr, a_yref_e=ye, a_u0=u_a, a_status=s_a, a_utraj=ob.utraj.copy(),
               a_iters=ob.qp_iters.copy())
    # cfg2 SQP to convergence
    cs = O.config_from_params(params, max_sqp_iter=50, sqp_tol=1e-8)
    rng = np.random.default_rng(103)
    x0 = random_states(rng, 8)
    yr, ye = setpoint_refs(rng, x0, N, hover)
    ob = O.OracleBatch(cs, 8)
    u_s, s_s, _, _ = ob.solve(x0, yr, ye)
    out.update(s_x0=x0, s_yref=yr, s_yref_e=ye, s_u0=u_s, s_status=s_s)
    # cfg5: N = 600
    p6 = yaml.safe_load(open(ROOT / 'rotors_mpc_controller_b200' / 'config' / 'params.yaml'))
    p6['solver']['horizon_steps'] = 600
    c6 = O.config_from_params(p6)
    rng = np.random.default_rng(104)
    x0 = random_states(rng, 2)
    yr, ye = setpoint_refs(rng, x0, 600, hover)
    ob = O.OracleBatch(c6, 2)
    u_l, s_l, _, _ = ob.solve(x0, yr, ye)
    out.update(l_x0=x0, l_sp=yr[:, 0, :].copy(), l_u0=u_l, l_status=s_l)
    np.savez_compressed(Path(__file__).with_name('nmpc_golden.npz'), **out)
    Path(__file__).with_name('nmpc_golden.json').write_text(json.dumps(
        {'generator': 'tests/golden/make_golden.py', 'source': 'oracle/ (C port, FP64)',
         'flags': FLAGS, 'arrays': sorted(out)}, indent=1))
    print('wrote', len(out), 'arrays')


if __name__ == '__main__':
    main()
