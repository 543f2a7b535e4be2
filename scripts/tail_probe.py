"""How much of the configs[1] launch (4096 problems, SQP to convergence) is the tail of the longest
problems?  Times the same workload at several batch sizes and prints the iteration histogram."""
import sys, copy, numpy as np, torch, yaml
sys.path.insert(0, '.')
from rotors_mpc_controller_b200 import BatchedPositionNMPC
from rotors_mpc_controller_b200.params import default_params_path
from tests.problems import random_states, setpoint_refs
params = yaml.safe_load(open(default_params_path()))
params['solver']['max_sqp_iter'] = 50
params['solver']['sqp_tol'] = 1e-8
dev = torch.device('cuda', 0)
for B in (444, 1776, 4096, 4096 * 4, 4096 * 16):
    ctl = BatchedPositionNMPC(copy.deepcopy(params), batch=B)
    rng = np.random.default_rng(0)
    x0 = random_states(rng, B)
    yref, yref_e = setpoint_refs(rng, x0, 20, ctl.hover_thrust)
    X, Y, Z = (torch.as_tensor(a, device=dev) for a in (x0, yref, yref_e))
    ms = []
    for rep in range(4):
        ctl.invalidate(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); ctl.solve(X, Y, Z); e1.record(); torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    s = ctl.stats()
    q, it = s['qp_iters'], s['sqp_iters']
    m = float(np.median(ms[1:]))
    print(f'B={B:6d} {m:8.3f} ms {B / m * 1e3:9.0f} solves/s  fact/s {q.sum() / m * 1e3:.3e}  sqp mean {it.mean():.1f} '
          f'p50 {np.percentile(it, 50):.0f} p90 {np.percentile(it, 90):.0f} p99 {np.percentile(it, 99):.0f} max {it.max()}  '
          f'fact mean {q.mean():.1f} max {q.max()}', flush=True)
    ctl.close()
