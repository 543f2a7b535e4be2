"""Time the three reference workloads with whatever library NMPC_B200_LIB points at."""
import sys, os, numpy as np, torch, yaml
sys.path.insert(0, '.')
from rotors_mpc_controller_b200 import BatchedPositionNMPC
from rotors_mpc_controller_b200.params import default_params_path
from tests.problems import random_states, setpoint_refs, aggressive
import copy
params = yaml.safe_load(open(default_params_path()))
dev = torch.device('cuda', 0)
tag = os.environ.get('NMPC_B200_LIB', 'default').split('/')[-1]
def run(name, B, sqp):
    p = copy.deepcopy(params)
    if sqp:
        p['solver']['max_sqp_iter'] = 50; p['solver']['sqp_tol'] = 1e-8
    ctl = BatchedPositionNMPC(p, batch=B)
    rng = np.random.default_rng(0)
    if name == 'aggr':
        x0, yref, yref_e = aggressive(rng, B, 20, ctl.hover_thrust)
    else:
        x0 = random_states(rng, B); yref, yref_e = setpoint_refs(rng, x0, 20, ctl.hover_thrust)
    X, Y, Z = (torch.as_tensor(a, device=dev) for a in (x0, yref, yref_e))
    ms = []
    for rep in range(5):
        ctl.invalidate(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); u0, st = ctl.solve(X, Y, Z); e1.record(); torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    it = ctl.stats()['qp_iters']
    m = float(np.median(ms[1:]))
    print(f'{tag:16s} {name:5s} B={B:6d} sqp={int(sqp)} {m:8.3f} ms {B/m*1e3:10.0f} solves/s  fact/solve {it.mean():.2f} ok {(st==0).float().mean().item():.3f} u0sum {u0.sum().item():.12e}', flush=True)
    ctl.close()
run('cfg2', 4096, True)
run('aggr', 65536, False)
run('cfg2', 65536, False)


def closed_loop(B=65536, steps=50):
    """Warm-started RTI in closed loop with the device plant (configs[2] shape)."""
    from rotors_mpc_controller_b200 import ClosedLoop, Plant
    ctl = BatchedPositionNMPC(copy.deepcopy(params), batch=B)
    cl = ClosedLoop(ctl, Plant(dt=ctl.dt, substeps=5))
    rng = np.random.default_rng(1)
    x = random_states(rng, B)
    sp = x[:, 0:3] + np.concatenate([rng.uniform(-2, 2, (B, 2)), rng.uniform(-0.5, 0.5, (B, 1))], 1)
    ms = []
    for rep in range(3):
        cl.reset(x)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); cl.run_setpoint(steps, sp); e1.record(); torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    m = float(np.median(ms[1:]))
    s = cl.summary(reduce=False)
    print(f'{tag:16s} loop  B={B:6d} steps={steps} {m:8.3f} ms {B*steps/m*1e3:10.0f} solves/s  failed {s["failed_solves"]}', flush=True)
    ctl.close()


closed_loop()
