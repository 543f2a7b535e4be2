#!/usr/bin/env python
"""BASELINE configs[3]: aggressive figure-8 and step references that drive the thrust bounds
active, batch 65 536 -- status and active-set agreement between the CUDA path and the oracle.

  python scripts/cfg4_agreement.py [--batch 65536] [--box reference|4-20]

Half of the batch tracks figure-8 references (A ~ U(1,3) m, period ~ U(2,4) s) from
a hover-like state, the other half takes 3-6 m set-point steps from large tilts / rates.  One RTI
solve on the GPU; the oracle solves the same problems on all host cores (one process per core).
Compared per problem: status code, u0, the whole input trajectory and the set of (stage, rotor)
pairs sitting on a bound.  Prints one JSON line.
"""
from __future__ import annotations

import argparse
import json
import multiprocessing as mp
import os
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

N = 20


def make_problems(B, hover):
    from rotors_mpc_controller_b200.reference import BatchReference
    from tests.problems import aggressive, random_states
    rng = np.random.default_rng(4)
    h = B // 2
    x0a, yra, yea = aggressive(rng, B - h, N, hover)
    x0b = random_states(rng, h, tilt_deg=30.0, v_sigma=1.0, w_sigma=1.0)
    ref = BatchReference(N, 0.05, hover)
    yrb, yeb = ref.figure8(rng.uniform(0, 4, h), rng.uniform(1, 3, h), rng.uniform(2, 4, h))
    return (np.concatenate([x0a, x0b]), np.concatenate([yra, yrb]), np.concatenate([yea, yeb]))


def _worker(args):
    lo, hi, box, x0, yref, yref_e = args
    import yaml
    from oracle import oracle as O
    from rotors_mpc_controller_b200.params import default_params_path
    params = yaml.safe_load(open(default_params_path()))
    cfg = O.config_from_params(params, thrust_box=box)
    ob = O.OracleBatch(cfg, hi - lo)
    u0, st, _, _ = ob.solve(x0, yref, yref_e)
    return lo, u0, st, ob.utraj.copy(), ob.qp_iters.copy()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--batch', type=int, default=65536)
    ap.add_argument('--box', default='reference', choices=['reference', '4-20'])
    args = ap.parse_args()
    import yaml
    from oracle import oracle as O
    from rotors_mpc_controller_b200 import BatchedPositionNMPC
    from rotors_mpc_controller_b200.params import default_params_path
    O.build()
    params = yaml.safe_load(open(default_params_path()))
    box = None
    if args.box == '4-20':                       # config/params.yaml:30, unused by the reference
        params['solver']['use_thrust_limits'] = True
        box = tuple(float(v) for v in params['controller']['thrust_limits'])
    B = args.batch
    ctl = BatchedPositionNMPC(params, batch=B)
    lo_b, hi_b = ctl.input_bounds
    x0, yref, yref_e = make_problems(B, ctl.hover_thrust)
    t0 = time.perf_counter()
    u0, st = ctl.solve(x0, yref, yref_e)
    t_gpu = time.perf_counter() - t0
    u0, st = u0.copy(), st.copy()
    _, ut, _ = ctl.trajectories()
    it = ctl.stats()['qp_iters']

    cores = os.cpu_count() or 1
    per = (B + cores - 1) // cores
    jobs = [(i, min(i + per, B), box, x0[i:i + per], yref[i:i + per], yref_e[i:i + per])
            for i in range(0, B, per)]
    t0 = time.perf_counter()
    with mp.get_context('fork').Pool(len(jobs)) as pool:
        parts = pool.map(_worker, jobs)
    t_cpu = time.perf_counter() - t0
    ur, sr = np.zeros_like(u0), np.zeros_like(st)
    utr, itr = np.zeros_like(ut), np.zeros_like(it)
    for lo, a, b, c, d in parts:
        ur[lo:lo + len(a)], sr[lo:lo + len(a)] = a, b
        utr[lo:lo + len(a)], itr[lo:lo + len(a)] = c, d

    ok = (st == 0) & (sr == 0)
    tol = 1e-7 * (hi_b - lo_b)
    act_g = (ut <= lo_b + tol) | (ut >= hi_b - tol)          # [B,N,4] on a bound
    act_o = (utr <= lo_b + tol) | (utr >= hi_b - tol)
    same_set = np.all(act_g == act_o, axis=(1, 2)) | ~ok
    scale = np.maximum(np.abs(ur), 1e-3 * (hi_b - lo_b))
    out = {
        'workload': f'configs[3]: {B} problems (half 3-6 m steps from large tilts, half figure-8 '
                    f'tracks), one RTI solve each, thrust box {lo_b[0]:.4f}..{hi_b[0]:.4f} N',
        'status_equal': bool(np.array_equal(st, sr)),
        'status_histogram_gpu': {int(k): int(v) for k, v in zip(*np.unique(st, return_counts=True))},
        'status_histogram_oracle': {int(k): int(v) for k, v in zip(*np.unique(sr, return_counts=True))},
        'max_abs_u0_diff': float(np.abs(u0 - ur)[ok].max()),
        'max_rel_u0_diff': float((np.abs(u0 - ur) / scale)[ok].max()),
        'max_abs_utraj_diff': float(np.abs(ut - utr)[ok].max()),
        'active_set_equal_problems': int(same_set.sum()),
        'problems_with_active_bounds': int(act_o.any(axis=(1, 2)).sum()),
        'mean_active_pairs_per_problem': float(act_o.sum(axis=(1, 2)).mean()),
        'stage0_saturated_frac': float(act_o[:, 0, :].any(axis=1).mean()),
        'factorisations_equal': bool(np.array_equal(it, itr)),
        'mean_factorisations': float(it.mean()), 'max_factorisations': int(it.max()),
        'gpu_host_call_s': t_gpu, 'oracle_wall_s': t_cpu, 'oracle_cores': len(jobs),
    }
    print(json.dumps(out), flush=True)
    ctl.close()
    bad = (not out['status_equal'] or out['max_rel_u0_diff'] > 1e-6
           or out['active_set_equal_problems'] != B)
    return 1 if bad else 0


if __name__ == '__main__':
    sys.exit(main())
