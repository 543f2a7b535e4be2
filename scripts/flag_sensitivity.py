"""How far does u0 move when ONE of the UPSTREAM (acados-version dependent, unverifiable here)
assumptions flips?  The table a later cross-check against a real acados needs to tell a wrong flag
from solver noise (controller.py:179-190, :257-261).

CPU only (oracle).  Writes profiles/r02_flag_sensitivity.json:
    python scripts/flag_sensitivity.py [--batch 512]
For each workload (configs[1] random set-point steps, configs[3] aggressive / bounds active) and
each solve mode (one cold RTI step = the reference's SQP_RTI call; SQP to convergence) the baseline
is the default flag set; every row flips one flag and reports max / median |u0 - u0_default| over
the batch and how many statuses changed.
"""
import argparse
import json
import sys
import time
from pathlib import Path

import numpy as np
import yaml

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
from oracle import oracle as O  # noqa: E402
from tests.problems import aggressive, random_states, setpoint_refs  # noqa: E402

FLIPS = [
    ('lm_scaled_by_dt=0 (LM term unscaled on stages, older acados)', dict(lm_scaled_by_dt=0)),
    ('cost_scale_dt=0 (stage cost not multiplied by dt)', dict(cost_scale_dt=0)),
    ('integrator rk4 x1 (ns=4, steps=1) instead of midpoint x2', dict(integ_ns=4, integ_steps=1)),
    ('integrator rk4 x2 (ns=4, steps=2)', dict(integ_ns=4, integ_steps=2)),
    ('integrator midpoint x1 (steps=1)', dict(integ_steps=1)),
    ('ipm_tol=1e-6 (HPIPM-like stop instead of 1e-14)', dict(ipm_tol=1e-6)),
    ('ipm_tol=1e-10', dict(ipm_tol=1e-10)),
    ('keep_state_box=1 (+-1e6 rows on all states, stages 1..N-1, controller.py:257-261)',
     dict(keep_state_box=1)),
    ('ipm_warm_start=0 (cold interior point for every QP of an SQP run)', dict(ipm_warm=0)),
]


def solve(params, x0, yr, ye, **flags):
    cfg = O.config_from_params(params, **flags)
    ob = O.OracleBatch(cfg, x0.shape[0])
    u0, st, _, _ = ob.solve(x0, yr, ye)
    return u0, st, ob.qp_iters.copy()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--batch', type=int, default=512)
    ap.add_argument('--out', default=str(ROOT / 'profiles' / 'r02_flag_sensitivity.json'))
    a = ap.parse_args()
    params = yaml.safe_load(open(ROOT / 'rotors_mpc_controller_b200' / 'config' / 'params.yaml'))
    base_cfg = O.config_from_params(params)
    hover = base_cfg.mass * base_cfg.gravity / 4.0
    N, B = base_cfg.N, a.batch
    rng = np.random.default_rng(0)
    x2 = random_states(rng, B)
    y2, e2 = setpoint_refs(rng, x2, N, hover)
    x4, y4, e4 = aggressive(np.random.default_rng(1), B, N, hover)
    out = {'batch': B, 'horizon': N, 'unit': 'N (thrust per rotor; hover = %.4f N)' % hover,
           'default_flags': dict(integ_ns=2, integ_steps=2, cost_scale_dt=1, lm_scaled_by_dt=1,
                                 ipm_tol=1e-14, keep_state_box=0, nan_status=4),
           'rows': []}
    t0 = time.time()
    for wname, (x0, yr, ye) in (('configs[1] random set-point steps', (x2, y2, e2)),
                                ('configs[3] aggressive (bounds active)', (x4, y4, e4))):
        for mname, mode in (('rti_cold', dict(max_sqp_iter=1)),
                            ('sqp_converged', dict(max_sqp_iter=50, sqp_tol=1e-8))):
            ub, sb, ib = solve(params, x0, yr, ye, **mode)
            ok = sb == 0
            for fname, flip in FLIPS:
                if mname == 'rti_cold' and 'ipm_warm' in flip:
                    continue
                u, s, it = solve(params, x0, yr, ye, **{**mode, **flip})
                both = ok & (s == 0)
                d = np.abs(u - ub).max(axis=1)[both]
                out['rows'].append({
                    'workload': wname, 'mode': mname, 'flip': fname,
                    'max_abs_du0': float(d.max()), 'median_abs_du0': float(np.median(d)),
                    'p99_abs_du0': float(np.percentile(d, 99)),
                    'status_changed': int((s != sb).sum()),
                    'mean_factorisations': float(it.mean()),
                    'mean_factorisations_default': float(ib.mean())})
                print(f'{wname[:10]} {mname:14s} {fname[:48]:48s} max {d.max():.3e}  '
                      f'median {np.median(d):.3e}  status changed {int((s != sb).sum())}', flush=True)
    out['wall_s'] = time.time() - t0
    Path(a.out).write_text(json.dumps(out, indent=1))
    print('wrote', a.out)


if __name__ == '__main__':
    main()
