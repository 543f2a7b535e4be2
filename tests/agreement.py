"""Full-size CUDA-vs-oracle agreement helpers (test infrastructure): the oracle side of a large
batch is split over all host cores, one forked process per chunk, as bench.py's cpu_sample does.
Used by tests/test_gpu_fullsize.py and scripts/cfg2_agreement.py / cfg4_agreement.py."""
from __future__ import annotations

import multiprocessing as mp
import os

import numpy as np


def _worker(args):
    params, flags, lo, x0, yref, yref_e = args
    from oracle import oracle as O
    cfg = O.config_from_params(params, **flags)
    ob = O.OracleBatch(cfg, x0.shape[0])
    u0, st, thrust, speed = ob.solve(x0, yref, yref_e)
    return lo, u0, st, ob.utraj.copy(), ob.qp_iters.copy()


def oracle_parallel(params, flags, x0, yref, yref_e, chunks_per_core=4):
    """Solve the batch with the C oracle on every host core.
    Returns dict(u0, status, utraj, qp_iters, cores)."""
    from oracle import oracle as O
    O.build()
    B = x0.shape[0]
    cores = os.cpu_count() or 1
    per = max(1, (B + chunks_per_core * cores - 1) // (chunks_per_core * cores))
    jobs = [(params, flags, i, x0[i:i + per], yref[i:i + per], yref_e[i:i + per])
            for i in range(0, B, per)]
    with mp.get_context('fork').Pool(min(cores, len(jobs))) as pool:
        parts = pool.map(_worker, jobs)
    N = yref.shape[1]
    out = dict(u0=np.zeros((B, 4)), status=np.zeros(B, np.int32), utraj=np.zeros((B, N, 4)),
               qp_iters=np.zeros(B, np.int32), cores=cores)
    for lo, u0, st, ut, it in parts:
        n = len(st)
        out['u0'][lo:lo + n], out['status'][lo:lo + n] = u0, st
        out['utraj'][lo:lo + n], out['qp_iters'][lo:lo + n] = ut, it
    return out


def cfg4_problems(B, N, hover, dt=0.05, seed=4):
    """configs[3] inputs: half 3-6 m set-point steps from large tilts / rates, half
    figure-8 tracks (A ~ U(1,3) m, period ~ U(2,4) s) from hover-like states."""
    from rotors_mpc_controller_b200.reference import BatchReference
    from tests.problems import aggressive, random_states
    rng = np.random.default_rng(seed)
    h = B // 2
    x0a, yra, yea = aggressive(rng, B - h, N, hover)
    x0b = random_states(rng, h, tilt_deg=30.0, v_sigma=1.0, w_sigma=1.0)
    ref = BatchReference(N, dt, hover)
    yrb, yeb = ref.figure8(rng.uniform(0, 4, h), rng.uniform(1, 3, h), rng.uniform(2, 4, h))
    return (np.concatenate([x0a, x0b]), np.concatenate([yra, yrb]), np.concatenate([yea, yeb]))


def active_sets(utraj, lo, hi):
    tol = 1e-7 * (hi - lo)
    return (utraj <= lo + tol) | (utraj >= hi - tol)
