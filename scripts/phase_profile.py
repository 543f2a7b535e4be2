"""Per-phase cycle breakdown of the solve kernel (instrumented build: scripts/build_variant.sh prof
-DNMPC_PROFILE, then NMPC_B200_LIB=$PWD/variants/prof.so python scripts/phase_profile.py)."""
import copy, ctypes as C, os, sys
import numpy as np, torch, yaml
sys.path.insert(0, '.')
from rotors_mpc_controller_b200 import BatchedPositionNMPC, _abi
from rotors_mpc_controller_b200.params import default_params_path
from tests.problems import random_states, setpoint_refs, aggressive
params = yaml.safe_load(open(default_params_path()))
dev = torch.device('cuda', 0)
NAMES = ['load', 'linearize', 'gradients', 'factor', 'forward', 'delta', 'elementwise', 'fullstep', 'epilogue', 'total']
lib = C.CDLL(str(_abi.LIB_PATH))

def cycles():
    out = (C.c_double * 10)()
    assert lib.nmpc_debug_phase_cycles(out, 10) == 0
    return np.array(list(out))

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
    for rep in range(3):
        ctl.invalidate(); torch.cuda.synchronize(); cycles()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); ctl.solve(X, Y, Z); e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
    c = cycles()
    st = ctl.stats()
    tot = c[:9].sum()
    print(f'{name}: B={B} N={N} {ms:.3f} ms {B/ms*1e3:.0f} solves/s, fact/solve {st["qp_iters"].mean():.2f}, sqp {st["sqp_iters"].mean():.2f}, '
          f'warp-cycles/solve {c[9]/B:.0f}')
    print('   ' + '  '.join(f'{n} {100*v/tot:.1f}%' for n, v in zip(NAMES[:9], c[:9])), flush=True)
    ctl.close()

which = sys.argv[1:] or ['all']
if 'all' in which or 'n20' in which:
    run('cfg2 sqp', 4096, sqp=True)
    run('cfg2 sqp', 65536, sqp=True)
    run('aggr rti', 65536, hard=True)
    run('cfg2 rti', 65536)
    run('cfg2 rti solo', 148, hard=True)
if 'all' in which or 'n600' in which:
    run('n600 rti', 1024, N=600)
    run('n600 rti', 148, N=600)
if 'small' in which:
    run('aggr rti', 1, hard=True)
    run('aggr rti', 148, hard=True)
