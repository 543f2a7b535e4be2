import sys, time, numpy as np, torch, yaml
sys.path.insert(0, '.'); sys.path.insert(0, '..')
from rotors_mpc_controller_b200 import BatchedPositionNMPC
from rotors_mpc_controller_b200.params import default_params_path
from tests.problems import random_states, setpoint_refs, aggressive
params = yaml.safe_load(open(default_params_path()))
dev = torch.device('cuda', 0)
for name, B in (('cfg2', 4096), ('cfg2', 65536), ('aggr', 65536)):
    ctl = BatchedPositionNMPC(params, batch=B)
    rng = np.random.default_rng(0)
    if name == 'cfg2':
        x0 = random_states(rng, B); yref, yref_e = setpoint_refs(rng, x0, 20, ctl.hover_thrust)
    else:
        x0, yref, yref_e = aggressive(rng, B, 20, ctl.hover_thrust)
    X, Y, Z = (torch.as_tensor(a, device=dev) for a in (x0, yref, yref_e))
    for rep in range(4):
        ctl.invalidate()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); u0, st = ctl.solve(X, Y, Z); e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        it = ctl.stats()['qp_iters']
        print(name, B, 'cold rep', rep, f'{ms:.3f} ms', f'{B/ms*1e3:.0f} solves/s', 'iters mean', it.mean(), 'max', it.max(), 'status', np.bincount(st.cpu().numpy()))
    for rep in range(3):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); u0, st = ctl.solve(X, Y, Z); e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        it = ctl.stats()['qp_iters']
        print(name, B, 'warm rep', rep, f'{ms:.3f} ms', f'{B/ms*1e3:.0f} solves/s', 'iters mean', it.mean(), 'max', it.max())
    ctl.close()
