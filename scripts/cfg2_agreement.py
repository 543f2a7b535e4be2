#!/usr/bin/env python
"""BASELINE configs[1] (the bench workload): batch of randomised hover states with random set-point
steps, full SQP to convergence -- CUDA path against the oracle on EVERY problem of the batch.

  python scripts/cfg2_agreement.py [--batch 4096] [--seed 1000]

The inputs are bench.py's (make_inputs(seed, batch, hover)); the oracle solves them on all host
cores.  Compared per problem: status, u0, SQP iteration count, Riccati factorisation count.
Prints one JSON line; exit code 1 if a status differs or u0 is off by more than 1e-6 relative.
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
import bench  # noqa: E402  (workload definition: SQP_CAP, SQP_TOL, make_inputs, load_params)


def _worker(args):
    lo, x0, yref, yref_e = args
    from oracle import oracle as O
    cfg = O.config_from_params(bench.load_params(), max_sqp_iter=bench.SQP_CAP,
                               sqp_tol=bench.SQP_TOL)
    ob = O.OracleBatch(cfg, x0.shape[0])
    u0, st, _, _ = ob.solve(x0, yref, yref_e)
    return lo, u0, st, ob.qp_iters.copy()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--batch', type=int, default=4096)
    ap.add_argument('--seed', type=int, default=1000)
    args = ap.parse_args()
    from oracle import oracle as O
    from rotors_mpc_controller_b200 import BatchedPositionNMPC
    O.build()
    params = bench.load_params()
    params['solver']['max_sqp_iter'] = bench.SQP_CAP
    params['solver']['sqp_tol'] = bench.SQP_TOL
    B = args.batch
    ctl = BatchedPositionNMPC(params, batch=B)
    x0, yref, yref_e = bench.make_inputs(args.seed, B, ctl.hover_thrust)
    t0 = time.perf_counter()
    u0, st = ctl.solve(x0, yref, yref_e)
    t_gpu = time.perf_counter() - t0
    u0, st = u0.copy(), st.copy()
    s = ctl.stats()

    cores = os.cpu_count() or 1
    per = (B + 4 * cores - 1) // (4 * cores)
    jobs = [(i, x0[i:i + per], yref[i:i + per], yref_e[i:i + per]) for i in range(0, B, per)]
    t0 = time.perf_counter()
    with mp.get_context('fork').Pool(cores) as pool:
        parts = pool.map(_worker, jobs)
    t_cpu = time.perf_counter() - t0
    ur, sr, qr = np.zeros_like(u0), np.zeros_like(st), np.zeros_like(s['qp_iters'])
    for lo, a, b, c in parts:
        ur[lo:lo + len(a)], sr[lo:lo + len(a)], qr[lo:lo + len(a)] = a, b, c
    ok = (st == 0) & (sr == 0)
    rel = np.abs(u0 - ur) / np.maximum(np.abs(ur), 1e-3)
    out = {
        'workload': f'configs[1]: {B} problems (bench.py inputs, seed {args.seed}), cold start, SQP '
                    f'to convergence (step <= {bench.SQP_TOL:g}, cap {bench.SQP_CAP})',
        'status_equal': bool(np.array_equal(st, sr)),
        'status_histogram_gpu': {int(k): int(v) for k, v in zip(*np.unique(st, return_counts=True))},
        'max_abs_u0_diff': float(np.abs(u0 - ur)[ok].max()),
        'max_rel_u0_diff': float(rel[ok].max()),
        'factorisations_equal_frac': float(np.mean(s['qp_iters'] == qr)),
        'mean_factorisations_gpu': float(s['qp_iters'].mean()),
        'mean_factorisations_oracle': float(qr.mean()),
        'mean_sqp_iters': float(s['sqp_iters'].mean()), 'max_sqp_iters': int(s['sqp_iters'].max()),
        'gpu_host_call_s': t_gpu, 'oracle_wall_s': t_cpu, 'oracle_cores': cores,
    }
    print(json.dumps(out), flush=True)
    ctl.close()
    return 0 if out['status_equal'] and out['max_rel_u0_diff'] <= 1e-6 else 1


if __name__ == '__main__':
    sys.exit(main())
