"""Fixed workload for ncu: configs[1] problems (SQP to convergence) at a chosen batch size.
B=444 puts one warp on each SM sub-partition: the stall profile of the serial chain."""
import copy
import sys
import numpy as np
import torch
import yaml
sys.path.insert(0, '.')
from rotors_mpc_controller_b200 import BatchedPositionNMPC
from rotors_mpc_controller_b200.params import default_params_path
from tests.problems import random_states, setpoint_refs

B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
params = yaml.safe_load(open(default_params_path()))
params['solver']['max_sqp_iter'] = 50
params['solver']['sqp_tol'] = 1e-8
dev = torch.device('cuda', 0)
ctl = BatchedPositionNMPC(copy.deepcopy(params), batch=B)
rng = np.random.default_rng(0)
x0 = random_states(rng, B)
yref, yref_e = setpoint_refs(rng, x0, 20, ctl.hover_thrust)
X, Y, Z = (torch.as_tensor(a, device=dev) for a in (x0, yref, yref_e))
for rep in range(3):
    ctl.invalidate()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    ctl.solve(X, Y, Z)
    e1.record()
    torch.cuda.synchronize()
    it = ctl.stats()['qp_iters']
    print(f'rep {rep}: {e0.elapsed_time(e1):.3f} ms, {B / e0.elapsed_time(e1) * 1e3:.0f} solves/s, '
          f'mean factorisations {it.mean():.2f}')
