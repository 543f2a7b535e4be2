"""Full-size agreement of the CUDA path with the oracle on EVERY problem of BASELINE.json's
configs at their stated sizes: configs[1] at the bench batch and seed,
configs[3] at 65 536 problems with both thrust boxes, configs[4] (N = 600) at batch 1 024.
The oracle side runs on all host cores (tests/agreement.py).  FP64 tolerance: statuses identical,
|u0 - oracle| <= 1e-9 N absolute (north_star asks 1e-6 relative), identical active sets and
Riccati-factorisation counts."""
import copy

import numpy as np
import pytest

from tests import agreement as A

pytestmark = pytest.mark.gpu


def _mk(params, B, **over):
    from rotors_mpc_controller_b200 import BatchedPositionNMPC
    p = copy.deepcopy(params)
    p['solver'].update(over)
    return BatchedPositionNMPC(p, batch=B), p


def test_cfg2_bench_batch_every_problem(params, built):
    """configs[1] exactly as bench.py times it: rank-0 inputs of the first input set (seed 0),
    4 096 problems, cold start, SQP to convergence."""
    import bench
    B = bench.BATCH
    ctl, p = _mk(bench.load_params(), B, max_sqp_iter=bench.SQP_CAP, sqp_tol=bench.SQP_TOL)
    x0, yref, yref_e = bench.make_inputs(0, B, ctl.hover_thrust)
    u0, st = ctl.solve(x0, yref, yref_e)
    u0, st = u0.copy(), st.copy()
    s = ctl.stats()
    o = A.oracle_parallel(p, dict(max_sqp_iter=bench.SQP_CAP, sqp_tol=bench.SQP_TOL),
                          x0, yref, yref_e)
    assert np.array_equal(st, o['status'])
    assert (st == 0).all()
    assert np.array_equal(s['qp_iters'], o['qp_iters'])       # same factorisation count everywhere
    assert np.max(np.abs(u0 - o['u0'])) <= 1e-9
    ctl.close()


@pytest.mark.parametrize('box', ['reference', '4-20'])
def test_cfg4_65536_status_and_active_sets(params, built, box):
    """configs[3]: 65 536 aggressive problems, one RTI solve each; status codes, u0, the whole input
    trajectory and the set of (stage, rotor) pairs on a bound agree with the oracle on every
    problem -- with the reference's box [0.0214, 6.003] N and the explicit [4, 20] N override."""
    B, N = 65536, 20
    over, flags = {}, {}
    if box == '4-20':      # config/params.yaml:30, parsed and never used by the reference
        over['use_thrust_limits'] = True
        flags['thrust_box'] = tuple(float(v) for v in params['controller']['thrust_limits'])
    ctl, p = _mk(params, B, **over)
    lo, hi = ctl.input_bounds
    x0, yref, yref_e = A.cfg4_problems(B, N, ctl.hover_thrust)
    u0, st = ctl.solve(x0, yref, yref_e)
    u0, st = u0.copy(), st.copy()
    _, ut, _ = ctl.trajectories()
    it = ctl.stats()['qp_iters']
    o = A.oracle_parallel(p, flags, x0, yref, yref_e, chunks_per_core=1)
    assert np.array_equal(st, o['status'])
    ok = st == 0
    assert ok.mean() > 0.99
    assert np.max(np.abs(u0 - o['u0'])[ok]) <= 1e-9
    assert np.max(np.abs(ut - o['utraj'])[ok]) <= 1e-9
    ag, ao = A.active_sets(ut, lo, hi), A.active_sets(o['utraj'], lo, hi)
    assert np.array_equal(ag[ok], ao[ok])
    assert ao[ok].any(axis=(1, 2)).sum() > (B // 4 if box == 'reference' else 0.99 * B)
    assert np.array_equal(it, o['qp_iters'])
    ctl.close()


def test_cfg5_long_horizon_1024_every_problem(params, built):
    """configs[4]: N = 600 (the slider maximum, cfg/rotors_mpc.cfg:9), iter_max = 600, batch 1 024,
    one cold RTI solve each, all 1 024 problems against the oracle."""
    B, N = 1024, 600
    ctl, p = _mk(params, B, horizon_steps=N)
    from tests.problems import random_states, setpoint_refs
    rng = np.random.default_rng(0)
    x0 = random_states(rng, B)
    yref, yref_e = setpoint_refs(rng, x0, N, ctl.hover_thrust)
    u0, st = ctl.solve(x0, yref, yref_e)
    u0, st = u0.copy(), st.copy()
    it = ctl.stats()['qp_iters']
    o = A.oracle_parallel(p, {}, x0, yref, yref_e)
    assert np.array_equal(st, o['status'])
    assert (st == 0).all()
    assert np.array_equal(it, o['qp_iters'])
    # 600 stages of Riccati recursion: round-off grows with the horizon (|u0| ~ 1.7 N)
    assert np.max(np.abs(u0 - o['u0'])) <= 1e-8
    ctl.close()
