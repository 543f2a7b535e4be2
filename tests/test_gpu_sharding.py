"""World-size-2 NCCL test of the batch sharding with the REAL rank-local solver
(BatchedPositionNMPC on cuda:rank): the (u0, status) all-gather north_star names, over
NVLink / NVSwitch.  Skipped on boxes with fewer than two GPUs (tests/test_sharding.py covers the
host logic on gloo with world_size 2)."""
import os

import numpy as np
import pytest

from tests.problems import random_states, setpoint_refs

pytestmark = pytest.mark.gpu


def _worker(rank, world, port, B, tmp):
    import torch
    import torch.distributed as dist
    import yaml
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = torch.device('cuda', rank)
    dist.init_process_group('nccl', rank=rank, world_size=world, device_id=dev)
    from rotors_mpc_controller_b200 import BatchedPositionNMPC
    from rotors_mpc_controller_b200.params import default_params_path
    from rotors_mpc_controller_b200.sharding import ShardedBatchNMPC
    params = yaml.safe_load(open(default_params_path()))
    rng = np.random.default_rng(0)            # every rank builds the same global problem set
    x0 = random_states(rng, B)
    hover = params['vehicle']['mass'] * 9.81 / 4
    yref, yref_e = setpoint_refs(rng, x0, 20, hover)
    x0[3, 0] = np.nan                         # one failing problem (status 4) in rank 0's shard
    sh = ShardedBatchNMPC(lambda n: BatchedPositionNMPC(params, batch=n, device=rank), B)
    tens = [torch.as_tensor(np.ascontiguousarray(sh.local_slice(a)), device=dev)
            for a in (x0, yref, yref_e)]
    u0, st = sh.solve_local(*tens)            # device tensors -> NCCL gathers device memory
    U, S = sh.gather(u0, st)
    assert U.is_cuda and U.shape == (B, 4)
    stats = sh.reduce_stats([float((st != 0).sum().item()), float(sh.local_batch)], device=dev)
    if rank == 0:
        np.savez(tmp, U=U.cpu().numpy(), S=S.cpu().numpy(), stats=stats, x0=x0, yref=yref,
                 yref_e=yref_e)
    sh.solver.close()
    dist.destroy_process_group()


def test_two_gpu_nccl_gather_matches_oracle(tmp_path, params, built):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip('needs two GPUs (gpurun --gpus 2)')
    import torch.multiprocessing as mp
    from oracle import oracle as O
    B = 1001                                  # ragged: 501 + 500
    out = str(tmp_path / 'gather.npz')
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_worker, args=(2, port, B, out), nprocs=2, join=True)
    g = np.load(out)
    ob = O.OracleBatch(O.config_from_params(params), B)
    u0, st, _, _ = ob.solve(g['x0'], g['yref'], g['yref_e'])
    assert np.array_equal(g['S'], st) and st[3] == 4 and np.sum(st != 0) == 1
    assert np.max(np.abs(g['U'] - u0)) < 1e-9
    assert list(g['stats']) == [1.0, float(B)]
