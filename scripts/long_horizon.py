"""BASELINE configs[4]: N=600 (slider maximum), iter_max=600, batch 1024: timing + oracle spot check."""
import sys, copy, numpy as np, torch, yaml
sys.path.insert(0, '.')
from rotors_mpc_controller_b200 import BatchedPositionNMPC
from rotors_mpc_controller_b200.params import default_params_path
from tests.problems import random_states, setpoint_refs
from oracle import oracle as O
params = yaml.safe_load(open(default_params_path()))
params['solver']['horizon_steps'] = 600
B, N = 1024, 600
ctl = BatchedPositionNMPC(params, batch=B)
rng = np.random.default_rng(0)
x0 = random_states(rng, B); yref, yref_e = setpoint_refs(rng, x0, N, ctl.hover_thrust)
dev = torch.device('cuda', 0)
X, Y, Z = (torch.as_tensor(a, device=dev) for a in (x0, yref, yref_e))
for rep in range(3):
    ctl.invalidate(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); u0, st = ctl.solve(X, Y, Z); e1.record(); torch.cuda.synchronize()
    it = ctl.stats()['qp_iters']
    print(f'N=600 B={B} cold RTI rep {rep}: {e0.elapsed_time(e1):.2f} ms, {B/e0.elapsed_time(e1)*1e3:.0f} solves/s, '
          f'factorisations mean {it.mean():.2f} max {it.max()}, status {np.bincount(st.cpu().numpy())}, '
          f'workspace {ctl.workspace_bytes/2**20:.0f} MiB')
from tests import agreement as A
o = A.oracle_parallel(params, {}, x0, yref, yref_e)
stn, u0n = st.cpu().numpy(), u0.cpu().numpy()
print('oracle check (all', B, 'problems): status equal', np.array_equal(stn, o['status']),
      'factorisations equal', np.array_equal(ctl.stats()['qp_iters'], o['qp_iters']),
      'max |u0 - oracle|', np.abs(u0n - o['u0']).max())
