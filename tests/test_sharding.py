"""Batch sharding + result gather (host logic), world_size 2 on gloo / CPU.  The rank-local
"solver" here is the CPU oracle (tests only); on GPUs it is BatchedPositionNMPC over NCCL."""
import os

import numpy as np
import pytest

from tests.problems import random_states, setpoint_refs


def test_shard_bounds_partition():
    from rotors_mpc_controller_b200.sharding import shard_bounds, shard_sizes
    for B in (0, 1, 7, 8, 4096, 1000003):
        for W in (1, 2, 3, 8):
            spans = [shard_bounds(r, W, B) for r in range(W)]
            assert spans[0][0] == 0 and spans[-1][1] == B
            assert all(spans[i][1] == spans[i + 1][0] for i in range(W - 1))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1 and sizes == shard_sizes(W, B)
    with pytest.raises(ValueError):
        shard_bounds(2, 2, 10)


def _worker(rank, world, port, B, tmp):
    import torch.distributed as dist
    import yaml
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    from oracle import oracle as O
    from rotors_mpc_controller_b200.params import default_params_path
    from rotors_mpc_controller_b200.sharding import ShardedBatchNMPC
    params = yaml.safe_load(open(default_params_path()))
    cfg = O.config_from_params(params)

    class OracleSolver:                       # same solve() surface as BatchedPositionNMPC
        def __init__(self, n):
            self.b = O.OracleBatch(cfg, n)

        def solve(self, x0, yref, yref_e):
            u0, st, _, _ = self.b.solve(x0, yref, yref_e)
            return u0, st

    rng = np.random.default_rng(0)            # every rank builds the same global problem set
    x0 = random_states(rng, B)
    yref, yref_e = setpoint_refs(rng, x0, cfg.N, cfg.mass * cfg.gravity / 4)
    x0[3, 0] = np.nan                         # one failing problem (status 4) in rank 0's shard
    sh = ShardedBatchNMPC(OracleSolver, B)
    u0, st = sh.solve_local(sh.local_slice(x0), sh.local_slice(yref), sh.local_slice(yref_e))
    U, S = sh.gather(u0, st)
    stats = sh.reduce_stats([float(np.sum(st != 0)), float(sh.local_batch)])
    if rank == 0:
        np.savez(tmp, U=U.numpy(), S=S.numpy(), stats=stats, x0=x0, yref=yref, yref_e=yref_e)
    dist.destroy_process_group()


def test_two_rank_gather_matches_single_process(tmp_path, params):
    import torch.multiprocessing as mp
    from oracle import oracle as O
    O.build()
    B = 11                                    # ragged: 6 + 5
    out = str(tmp_path / 'gather.npz')
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_worker, args=(2, port, B, out), nprocs=2, join=True)
    g = np.load(out)
    ob = O.OracleBatch(O.config_from_params(params), B)
    u0, st, _, _ = ob.solve(g['x0'], g['yref'], g['yref_e'])
    assert np.array_equal(g['S'], st) and st[3] == 4 and np.sum(st != 0) == 1
    assert np.max(np.abs(g['U'] - u0)) == 0.0
    assert list(g['stats']) == [1.0, float(B)]
