#!/usr/bin/env python
"""BASELINE configs[2]: closed-loop Monte Carlo, batch-sharded over the GPUs of one node.

  python scripts/monte_carlo.py --trajectories 131072 --steps 200                 # 1 GPU
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
      scripts/monte_carlo.py --trajectories 1000000 --steps 200                   # 8 GPUs

Every trajectory: random initial state around hover (SURVEY 8d cfg2 distribution), set-point step
at t=0 (or a figure-8 track with --figure8), `steps` control steps of SQP-RTI with warm start and
last-valid fallback, nominal RK4 plant, re-solved at the node's rate (60 Hz, params.yaml node.rate;
--control-rate 20 steps the plant by the solver dt instead).  Rank g owns the contiguous slice [g*B/G, (g+1)*B/G) for
the whole run; the only communication is the final reduction of the statistics vector.
Prints one JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--trajectories', type=int, default=131072)
    ap.add_argument('--steps', type=int, default=200)
    ap.add_argument('--substeps', type=int, default=5)
    ap.add_argument('--control-rate', type=float, default=None,
                    help='re-solve rate in Hz (default: node.rate of params.yaml = 60 Hz, the '
                         'reference node; the horizon keeps the solver dt = 0.05 s)')
    ap.add_argument('--figure8', action='store_true')
    ap.add_argument('--motor-lag', action='store_true')
    ap.add_argument('--rotors-plant', action='store_true',
                    help="RotorS-faithful model-mismatch plant: motor lag, rotor drag, rolling "
                         "moment 1e-6, mass 0.716 kg and RotorS' yaw-torque sign (SURVEY Q3, Q5)")
    ap.add_argument('--chunk', type=int, default=131072, help='vehicles per handle (HBM bound)')
    ap.add_argument('--precision', default='f64')
    args = ap.parse_args()

    import torch
    import torch.distributed as dist
    import yaml
    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        dist.init_process_group('nccl', device_id=dev)
    from rotors_mpc_controller_b200 import BatchedPositionNMPC, ClosedLoop, Plant
    from rotors_mpc_controller_b200.params import default_params_path
    from rotors_mpc_controller_b200.sharding import shard_bounds
    from tests.problems import random_states

    params = yaml.safe_load(open(default_params_path()))
    params['solver']['precision'] = args.precision
    rate = args.control_rate or float(params.get('node', {}).get('rate', 60.0))   # mpc_controller_node:55
    lo, hi = shard_bounds(rank, world, args.trajectories)
    n_local = hi - lo
    plant_kw = dict(substeps=args.substeps)
    if args.motor_lag:
        plant_kw.update(motor_model=1, rotor_drag=8.06428e-5)
    if args.rotors_plant:   # hummingbird.xacro:29-46, gazebo_motor_model.cpp:415-483
        plant_kw.update(motor_model=1, rotor_drag=8.06428e-5, rolling_moment=1e-6, mass=0.716,
                        rotors_yaw_sign=True)

    agg = np.zeros(9)
    mx = np.zeros(2)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    gpu_ms, launches = 0.0, 0
    done = 0
    ctl = cl = None
    while done < n_local:
        b = min(args.chunk, n_local - done)
        if ctl is None or ctl.batch != b:
            if ctl is not None:
                ctl.close()
            ctl = BatchedPositionNMPC(params, batch=b, device=local)
            cl = ClosedLoop(ctl, Plant(dt=1.0 / rate, **plant_kw))
        rng = np.random.default_rng([rank, done])
        x = random_states(rng, b)
        cl.reset(x)
        l0 = ctl.kernel_launches
        torch.cuda.synchronize(dev)
        ev0.record()
        if args.figure8:
            cl.run_figure8(args.steps, rng.uniform(0, 4, b), rng.uniform(1, 3, b),
                           rng.uniform(2, 4, b))
        else:
            sp = x[:, 0:3] + np.concatenate([rng.uniform(-2, 2, (b, 2)),
                                             rng.uniform(-0.5, 0.5, (b, 1))], axis=1)
            cl.run_setpoint(args.steps, sp, rng.uniform(-np.pi, np.pi, b))
        ev1.record()
        torch.cuda.synchronize(dev)
        gpu_ms += ev0.elapsed_time(ev1)
        launches += ctl.kernel_launches - l0
        s = cl.summary(reduce=False)                      # local part only; reduced once below
        agg += np.array([s['problems'], s['mean_sq_pos_err_sum'] * s['problems'],
                         s['mean_final_pos_err'] * s['problems'], s['failed_solves'],
                         s['qp_iters'], s['saturated_steps'], s['thrust_sum'], s['nonfinite'],
                         s['problems_with_failures']])
        mx = np.maximum(mx, [s['max_pos_err'], s['max_final_pos_err']])
        done += b
    t = torch.tensor(np.concatenate([agg, [launches]]), dtype=torch.float64, device=dev)
    m = torch.tensor(np.concatenate([mx, [gpu_ms]]), dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM)          # the <=64-entry statistics vector
        dist.all_reduce(m, op=dist.ReduceOp.MAX)
    t, m = t.cpu().numpy(), m.cpu().numpy()
    if rank == 0:
        n, solves = t[0], t[0] * args.steps
        print(json.dumps({
            'workload': 'configs[2] closed-loop Monte Carlo' + (' (figure-8)' if args.figure8 else ''),
            'n_gpus': world, 'trajectories': int(n), 'steps': args.steps,
            'rti_solves': int(solves), 'gpu_ms_max_over_ranks': m[2],
            'solves_per_s': solves / (m[2] * 1e-3), 'kernel_launches': int(t[9]),
            'mean_factorisations_per_solve': t[4] / solves,
            'failed_solves': int(t[3]), 'trajectories_with_failures': int(t[8]),
            'nonfinite': int(t[7]), 'saturated_step_frac': t[5] / solves,
            'rms_pos_err': float(np.sqrt(t[1] / solves)), 'mean_final_pos_err': t[2] / n,
            'max_final_pos_err': m[1], 'precision': args.precision,
            'plant': plant_kw, 'control_rate_hz': rate, 'horizon_dt': params['solver']['dt']}), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
