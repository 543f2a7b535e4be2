"""Latency of small batches with / without helper teams (NMPC_B200_NO_TEAMS=1)."""
import copy, os, sys
import numpy as np, torch, yaml
sys.path.insert(0, '.')
from rotors_mpc_controller_b200 import BatchedPositionNMPC
from rotors_mpc_controller_b200.params import default_params_path
from tests.problems import random_states, setpoint_refs, aggressive
params = yaml.safe_load(open(default_params_path()))
dev = torch.device('cuda', 0)
tag = 'no-teams' if os.environ.get('NMPC_B200_NO_TEAMS') else 'teams'
def run(name, B, N=20, sqp=False, hard=False):
    p = copy.deepcopy(params)
    p['solver']['horizon_steps'] = N
    if sqp:
        p['solver']['max_sqp_iter'] = 50; p['solver']['sqp_tol'] = 1e-8
    ctl = BatchedPositionNMPC(p, batch=B)
    rng = np.random.default_rng(0)
    if hard:
        x0, yref, yref_e = aggressive(rng, B, N, ctl.hover_thrust)
    else:
        x0 = random_states(rng, B); yref, yref_e = setpoint_refs(rng, x0, N, ctl.hover_thrust)
    X, Y, Z = (torch.as_tensor(a, device=dev) for a in (x0, yref, yref_e))
    ms = []
    for rep in range(7):
        ctl.invalidate(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); u0, st = ctl.solve(X, Y, Z); e1.record(); torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    it = ctl.stats()
    m = float(np.median(ms[2:]))
    print(f'{tag:9s} {name:10s} B={B:5d} N={N:3d} {m*1e3:10.1f} us  fact/solve {it["qp_iters"].mean():.2f} sqp {it["sqp_iters"].mean():.1f} u0sum {u0.sum().item():.12e}', flush=True)
    ctl.close()
run('aggr rti', 1, hard=True)
run('aggr rti', 148, hard=True)
run('aggr rti', 444, hard=True)
run('aggr rti', 888, hard=True)
run('cfg2 sqp', 444, sqp=True)
run('cfg2 sqp', 1776, sqp=True)
run('cfg2 sqp', 4096, sqp=True)
run('n600 rti', 64, N=600)
run('n600 rti', 1024, N=600)
