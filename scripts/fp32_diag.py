"""FP32 variant vs FP64 (both on the GPU): where does the u0 error come from?"""
import sys, copy, numpy as np, yaml
sys.path.insert(0, '.')
from rotors_mpc_controller_b200 import BatchedPositionNMPC
from rotors_mpc_controller_b200.params import default_params_path
from tests.problems import random_states, setpoint_refs, aggressive
params = yaml.safe_load(open(default_params_path()))
B, N = 2048, 20
def mk(**kw):
    p = copy.deepcopy(params); p['solver'].update(kw); return BatchedPositionNMPC(p, batch=B)
import itertools
for kind, tol in itertools.product(('cfg2', 'aggr'), (1e-7, 1e-8, 1e-9)):
    rng = np.random.default_rng(21)
    c64, c32 = mk(), mk(precision='f32', ipm_tol=tol)
    print('== ipm_tol', tol)
    if kind == 'cfg2':
        x0 = random_states(rng, B); yref, yref_e = setpoint_refs(rng, x0, N, c64.hover_thrust)
    else:
        x0, yref, yref_e = aggressive(rng, B, N, c64.hover_thrust)
    u64, s64 = c64.solve(x0, yref, yref_e); u64 = u64.copy(); it64 = c64.stats()['qp_iters']
    u32, s32 = c32.solve(x0, yref, yref_e); u32 = u32.copy(); it32 = c32.stats()['qp_iters']
    okm = (s32 == 0)
    rel = np.abs(u32 - u64).max(axis=1) * okm / np.maximum(np.abs(u64).max(axis=1), c64.hover_thrust)
    print(kind, 'status32 hist', np.bincount(s32, minlength=5), 'status eq', np.array_equal(s64, s32), 'iters64 mean', it64.mean(), 'iters32 mean', it32.mean())
    for lo, hi in ((1, 1), (2, 5), (6, 10), (11, 100)):
        m = (it64 >= lo) & (it64 <= hi)
        if m.any():
            print(f'  it64 in [{lo},{hi}] n={m.sum():5d} rel err max {rel[m].max():.2e} median {np.median(rel[m]):.2e} p99 {np.quantile(rel[m],0.99):.2e}')
    x64, uu64, _ = c64.trajectories(); x32, uu32, _ = c32.trajectories()
    print('  traj err max x', np.abs(x64 - x32).max(), 'u', np.abs(uu64 - uu32).max())
    w = np.argmax(rel); print('  worst', w, rel[w], it64[w], it32[w], u64[w], u32[w])
    # linearisation only
    xs = rng.normal(size=(64, 13)); us = rng.uniform(0, 6, (64, 4))
    a = c64.discretize(xs, us); b = c32.discretize(xs, us)
    print('  discretize err', [float(np.abs(p - q).max()) for p, q in zip(a, b)])
