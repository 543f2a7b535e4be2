"""Small invocations of every kernel variant, meant to run under compute-sanitizer
(scripts/gpu_sanitize.sh: memcheck, racecheck, synccheck, initcheck).  SURVEY.md section 5.

Each case goes through the public Python front-end -> C ABI -> kernels and prints one line; a
sanitizer error makes the tool exit non-zero (--error-exitcode)."""
import copy
import sys
from pathlib import Path

import numpy as np
import torch
import yaml

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
from rotors_mpc_controller_b200 import BatchedPositionNMPC, ClosedLoop, Plant  # noqa: E402
from rotors_mpc_controller_b200.params import default_params_path  # noqa: E402
from tests.problems import aggressive, random_states, setpoint_refs  # noqa: E402


def mk(params, B, **over):
    p = copy.deepcopy(params)
    p['solver'].update(over)
    return BatchedPositionNMPC(p, batch=B)


def main():
    which = set(sys.argv[1:]) or {'rti', 'sqp', 'n15', 'misaligned', 'n600', 'f32', 'rollout',
                                  'cta'}
    params = yaml.safe_load(open(default_params_path()))
    dev = torch.device('cuda', 0)
    rng = np.random.default_rng(7)

    def inputs(ctl, B, N, hard=False):
        if hard:
            return aggressive(rng, B, N, ctl.hover_thrust)
        x0 = random_states(rng, B)
        return (x0,) + setpoint_refs(rng, x0, N, ctl.hover_thrust)

    if 'rti' in which:      # N = 20 fixed-horizon kernel, TMA-staged inputs, cold then warm
        ctl = mk(params, 64)
        x0, yr, ye = inputs(ctl, 64, 20, hard=True)
        for _ in range(2):
            u0, st = ctl.solve(x0, yr, ye)
        print('rti   B=64 N=20 staged  status', np.bincount(st, minlength=5).tolist(), flush=True)
        ctl.close()
    if 'sqp' in which:      # SQP to convergence with the interior-point warm start
        ctl = mk(params, 32, max_sqp_iter=50, sqp_tol=1e-8)
        x0, yr, ye = inputs(ctl, 32, 20)
        u0, st = ctl.solve(x0, yr, ye)
        print('sqp   B=32 N=20         status', np.bincount(st, minlength=5).tolist(),
              'factorisations', int(ctl.stats()['qp_iters'].sum()), flush=True)
        ctl.close()
    if 'n15' in which:      # odd horizon: generic kernel, direct loads
        ctl = mk(params, 16, horizon_steps=15)
        x0, yr, ye = inputs(ctl, 16, 15, hard=True)
        for _ in range(2):
            u0, st = ctl.solve(x0, yr, ye)
        print('n15   B=16 N=15 direct  status', np.bincount(st, minlength=5).tolist(), flush=True)
        ctl.close()
    if 'misaligned' in which:   # device yref at an 8 (mod 16) byte address -> direct-load path
        ctl = mk(params, 16)
        x0, yr, ye = inputs(ctl, 16, 20)
        buf = torch.empty(yr.size + 1, dtype=torch.float64, device=dev)
        yt = buf[1:].view(16, 20, 17)
        yt.copy_(torch.as_tensor(yr))
        assert yt.data_ptr() % 16 == 8
        u0, st = ctl.solve(torch.as_tensor(x0, device=dev), yt, torch.as_tensor(ye, device=dev))
        torch.cuda.synchronize()
        print('misal B=16 N=20 direct  status', np.bincount(st.cpu().numpy(), minlength=5).tolist(),
              flush=True)
        ctl.close()
    if 'n600' in which:     # long horizon, workspace streamed through L2/HBM
        ctl = mk(params, 4, horizon_steps=600)
        x0, yr, ye = inputs(ctl, 4, 600)
        u0, st = ctl.solve(x0, yr, ye)
        print('n600  B=4  N=600        status', np.bincount(st, minlength=5).tolist(), flush=True)
        ctl.close()
    if 'f32' in which:
        ctl = mk(params, 32, precision='f32')
        x0, yr, ye = inputs(ctl, 32, 20, hard=True)
        u0, st = ctl.solve(x0, yr, ye)
        print('f32   B=32 N=20         status', np.bincount(st, minlength=5).tolist(), flush=True)
        ctl.close()
    if 'rollout' in which:  # reference kernel + solve + plant kernel, 5 control steps, motor lag
        ctl = mk(params, 32)
        cl = ClosedLoop(ctl, Plant(dt=ctl.dt, substeps=5, motor_model=1, rotor_drag=8.06428e-5))
        x = random_states(rng, 32)
        cl.reset(x)
        cl.run_figure8(5, rng.uniform(0, 4, 32), rng.uniform(1, 3, 32), rng.uniform(2, 4, 32))
        torch.cuda.synchronize()
        print('roll  B=32 5 steps      failed', cl.summary(reduce=False)['failed_solves'], flush=True)
        ctl.close()
    print('sanitize_cases: done', flush=True)


if __name__ == '__main__':
    main()
