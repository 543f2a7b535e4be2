#!/usr/bin/env python
"""bench.py -- NMPC solves/s of the hot path on B200 (see DESIGN.md "Measurement").

  python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
  python bench.py --impl reference --gpus N --steps K ...  # CPU arm (oracle port, all host cores)

A *step* is one pass of the hot path over one batch: `batch` independent PositionNMPC.solve()
calls (BASELINE.json configs[1]: 4096 hummingbirds per GPU, randomised states around hover,
random set-point steps, N=20, cold start, SQP to convergence).  One JSON line is printed by
rank 0.  The acados reference itself is unavailable on this box (not vendored, not
installable), so the reference arm times the CPU oracle port of the same algorithm.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

BATCH = 4096          # per GPU (configs[1])
HORIZON = 20
SQP_CAP = 50          # "full SQP to convergence": step inf-norm <= 1e-8, cap 50
SQP_TOL = 1e-8
N_INPUT_SETS = 5      # seeds 0..4, cycled over the steps
# algorithmic work per solve (SURVEY.md 8d / DESIGN.md): flops = S*F_LIN + I*F_IPM
F_LIN = 0.507e6       # per SQP iteration: 20 stages x 2 steps x 2 RK stages, dense-block count
F_IPM = 0.244e6       # per Riccati factorisation (+ predictor / corrector solves)
BYTES_PER_SOLVE = 5752 + 2924   # compulsory HBM I/O per solve, FP64


def load_params():
    import yaml
    from rotors_mpc_controller_b200.params import default_params_path
    p = yaml.safe_load(open(default_params_path()))
    p['solver']['horizon_steps'] = HORIZON
    return p


def make_inputs(seed, batch, hover):
    from tests.problems import random_states, setpoint_refs
    rng = np.random.default_rng(seed)
    x0 = random_states(rng, batch)
    yref, yref_e = setpoint_refs(rng, x0, HORIZON, hover)
    return x0, yref, yref_e


# ------------------------------------------------------------------ CPU arm (oracle port)
def _cpu_worker(args):
    seed, lo, hi, sqp = args
    from oracle import oracle as O
    params = load_params()
    cfg = O.config_from_params(params, max_sqp_iter=sqp, sqp_tol=SQP_TOL)
    hover = cfg.mass * cfg.gravity / 4.0
    x0, yref, yref_e = make_inputs(seed, hi, hover)
    ob = O.OracleBatch(cfg, hi - lo)
    t0 = time.perf_counter()
    ob.solve(x0[lo:hi], yref[lo:hi], yref_e[lo:hi])
    return time.perf_counter() - t0, int(ob.qp_iters.sum())


def cpu_sample(n_problems, cores, seed=0, sqp=SQP_CAP):
    """Solve `n_problems` of the workload with the oracle, one process per core.
    Returns (solves/s over the wall time of the parallel section, wall seconds)."""
    import multiprocessing as mp
    from oracle import oracle as O
    O.build()
    per = (n_problems + cores - 1) // cores
    jobs = [(seed, i * per, min((i + 1) * per, n_problems), sqp) for i in range(cores)
            if i * per < n_problems]
    ctx = mp.get_context('fork')
    with ctx.Pool(len(jobs)) as pool:
        pool.map(_cpu_worker, [(seed, 0, 8, sqp)] * len(jobs))  # warm the workers
        t0 = time.perf_counter()
        pool.map(_cpu_worker, jobs)
        wall = time.perf_counter() - t0
    return n_problems / wall, wall


def run_reference(args):
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return 0
    cores = os.cpu_count() or 1
    n = 256 * cores if cores <= 32 else 128 * cores
    times = []
    for i in range(args.warmup + args.steps):
        v, wall = cpu_sample(n, cores, seed=i % N_INPUT_SETS)
        if i >= args.warmup:
            times.append(wall)
    value = n * len(times) / sum(times)
    sample = (f'{n} problems per step of the configs[1] workload (cold start, SQP to '
              f'convergence), split over {cores} processes')
    line = {
        'impl': 'reference', 'metric': 'NMPC solves/sec (hummingbird, N=20)', 'value': value,
        'unit': 'solves/s', 'n_gpus': args.gpus, 'steps': args.steps, 'warmup': args.warmup,
        'ms_per_step': 1e3 * sum(times) / len(times), 'higher_is_better': True,
        'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f64', 'data': 'synthetic',
        'config': {**workload_config(args.gpus), 'batch_per_gpu': None, 'global_batch': n,
                   'parallelism': f'one process per host core x{cores} (rank 0 only), no GPU',
                   'l2': 'n/a (CPU arm)',
                   'note': f'each step is a bounded sample of the workload: {n} problems'},
        'cpu_baseline': {'value': value, 'unit': 'solves/s', 'cores': cores, 'kind': 'port',
                         'sample': sample},
        'e2e': {'value': value, 'unit': 'solves/s', 'h2d_bytes_per_step': 0,
                'd2h_bytes_per_step': 0},
        'note': 'acados/HPIPM (the true reference implementation) is unavailable on this box; '
                'this arm times oracle/ (C port of the same algorithm), one process per core',
    }
    print(json.dumps(line), flush=True)
    return 0


def workload_config(gpus, dtype='f64'):
    par = (f'batch-sharded x{gpus}; per step an NCCL all-gather of (u0[4], status) per problem '
           f'(40 B) on a side stream, overlapped with the next step' if gpus > 1
           else 'single GPU (batch-sharded x1, nothing to gather)')
    return {'workload': 'configs[1]: batch 4096 hummingbirds per GPU, randomised states around '
                        'hover, random set-point steps, N=20, dt=0.05, cold start, SQP to '
                        f'convergence (step<={"1e-4" if dtype == "f32" else "1e-8"}, cap 50), ERK midpoint x2, '
                        'u-box [0.0214,6.003] N',
            'batch_per_gpu': BATCH, 'horizon': HORIZON, 'global_batch': BATCH * gpus,
            'parallelism': par,
            'l2': 'flushed between timed steps (256 MiB write, outside the event pair)'}


# ------------------------------------------------------------------ clocks sampler
class ClockSampler:
    Q = ('clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,'
         'clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,'
         'clocks_event_reasons.sw_power_cap')

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ['nvidia-smi', f'--query-gpu={self.Q}', '--format=csv,noheader,nounits',
                 '-i', str(self.index), '-lms', '100'], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(',')])

    def stop(self):
        if self.proc is None:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
                for n, v in zip(names, r[2:6]):
                    if v.lower().startswith('active'):
                        reasons.add(n)
            except Exception:
                pass
        return {'sm_mhz': float(np.median(sm)) if sm else None,
                'sm_max_mhz': max(mx) if mx else None, 'reasons': sorted(reasons),
                'samples': len(sm)}


# ------------------------------------------------------------------ result gather (N > 1)
class Gather:
    """Per-step NCCL all-gather of (u0[B,4], status[B]) -- 40 B per problem -- on a side stream.
    launch() is asynchronous: it orders itself behind the solve (event / stream), packs the two
    outputs into one [B,5] tensor and all-gathers it; the next step's solve overlaps it."""

    def __init__(self, ctl, dev, world):
        import torch
        import torch.distributed as dist
        self.torch, self.dist, self.dev, self.world = torch, dist, dev, world
        self.side = torch.cuda.Stream(dev)
        B = ctl.batch
        self.pack = torch.zeros((B, 5), dtype=torch.float64, device=dev)
        self.out = torch.empty((world * B, 5), dtype=torch.float64, device=dev)
        self.events, self.launches, self.ctl = [], 0, ctl
        self.last_local = None

    def handle_outputs(self):
        """Device-resident (u0, status) of the last nmpc_solve_host call (zero-copy views)."""
        import ctypes as C
        from rotors_mpc_controller_b200 import _abi
        from rotors_mpc_controller_b200.closed_loop import _DeviceArray
        B = self.ctl.batch
        ptrs = [C.c_void_p() for _ in range(4)]
        _abi.check(_abi.lib().nmpc_last_outputs(self.ctl._h.ptr, *[C.byref(p) for p in ptrs]),
                   'nmpc_last_outputs')
        u0 = self.torch.as_tensor(_DeviceArray(ptrs[0].value, (B, 4), '<f8'), device=self.dev)
        st = self.torch.as_tensor(_DeviceArray(ptrs[1].value, (B,), '<i4'), device=self.dev)
        return u0, st

    def launch(self, u0, status, after=None):
        torch = self.torch
        g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        if after is not None:
            self.side.wait_event(after)
        with torch.cuda.stream(self.side):
            g0.record()
            self.pack[:, :4].copy_(u0)
            self.pack[:, 4].copy_(status)
            self.dist.all_gather_into_tensor(self.out, self.pack)
            g1.record()
        self.events.append((g0, g1))
        self.launches += 3
        self.last_local = (u0, status)

    def drain(self):
        """Wait for the outstanding gathers; returns the wait in ms (host clock)."""
        t0 = time.perf_counter()
        self.side.synchronize()
        return (time.perf_counter() - t0) * 1e3

    def mean_ms(self):
        ms = [a.elapsed_time(b) for a, b in self.events[3:]]
        return float(np.mean(ms)) if ms else None

    def check(self, ctl):
        """The slice of this rank inside the gathered tensor equals its local outputs."""
        rank = self.dist.get_rank()
        B = ctl.batch
        mine = self.out[rank * B:(rank + 1) * B]
        u0, st = self.last_local
        return bool(self.torch.equal(mine[:, :4], u0) and
                    self.torch.equal(mine[:, 4].to(self.torch.int32), st.to(self.torch.int32)))


# ------------------------------------------------------------------ GPU arm
def run_gpu(args):
    import torch
    import torch.distributed as dist
    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    if not torch.cuda.is_available():
        raise SystemExit('bench.py: no CUDA device (the product path has no CPU fallback)')
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        dist.init_process_group('nccl', device_id=dev)

    import __graft_entry__ as entry
    if rank == 0:
        entry.build()
    if world > 1:
        dist.barrier()
    from rotors_mpc_controller_b200 import BatchedPositionNMPC, _abi
    params = load_params()
    params['solver']['max_sqp_iter'] = SQP_CAP
    params['solver']['sqp_tol'] = SQP_TOL
    params['solver']['precision'] = args.dtype
    if args.dtype == 'f32':
        params['solver']['sqp_tol'] = 1e-4     # FP32 resolves steps to ~1e-5 (parity bar: 1e-3)
    ctl = BatchedPositionNMPC(params, batch=BATCH, device=local)
    hover = ctl.hover_thrust

    # Multi-GPU: the (u0, status) all-gather north_star names, issued after every step's solve on a
    # side stream so that it overlaps the next step (NCCL over NVLink 5 / NVSwitch).
    gather = Gather(ctl, dev, world) if world > 1 else None

    # inputs: N_INPUT_SETS seeded batches per rank, resident in HBM and in pinned host memory
    dev_in, host_in = [], []
    for s in range(N_INPUT_SETS):
        arrs = make_inputs(1000 * rank + s, BATCH, hover)
        dev_in.append(tuple(torch.as_tensor(a, device=dev) for a in arrs))
        pinned = []
        for a in arrs:
            t = torch.empty(a.shape, dtype=torch.float64, pin_memory=True)
            t.numpy()[...] = a
            pinned.append(t.numpy())
        host_in.append(tuple(pinned))
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    main = torch.cuda.current_stream(dev)

    def one_step(i, timed):
        x0, yref, yref_e = dev_in[i % N_INPUT_SETS]
        ctl.invalidate()          # cold start every step: each step is a fresh batch
        flush.zero_()             # evict the previous step's data from L2
        main.synchronize()        # (not the side stream: the previous step's gather may still run)
        e0 = torch.cuda.Event(enable_timing=True)
        e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        u0, st = ctl.solve(x0, yref, yref_e)
        e1.record()
        if gather is not None:
            gather.launch(u0, st, after=e1)
        main.synchronize()
        return (e0.elapsed_time(e1) if timed else 0.0), st

    W, K = max(args.warmup, 3), args.steps
    for i in range(W):
        one_step(i, False)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    barrier()
    launches0 = ctl.kernel_launches
    step_ms, qp_sum, sqp_sum, ok_count = [], 0, 0, 0
    wall0 = time.perf_counter()
    for i in range(K):
        ms, st = one_step(W + i, True)
        step_ms.append(ms)
        stats = ctl.stats()
        qp_sum += int(stats['qp_iters'].sum())
        sqp_sum += int(stats['sqp_iters'].sum())
        ok_count += int((st == 0).sum().item())
    drain_ms = gather.drain() if gather is not None else 0.0   # last gather(s) not yet complete
    barrier()
    wall = time.perf_counter() - wall0
    launches = ctl.kernel_launches - launches0 + (gather.launches if gather is not None else 0)
    total_ms = float(sum(step_ms)) + drain_ms
    nccl_ms = gather.mean_ms() if gather is not None else None
    gathered_ok = gather.check(ctl) if gather is not None else None

    # end-to-end through the public host API: pinned host buffers, H2D + solve + D2H per step
    e2e_t = []
    for i in range(W + K):
        x0, yref, yref_e = host_in[i % N_INPUT_SETS]
        ctl.invalidate()
        flush.zero_()
        main.synchronize()
        t0 = time.perf_counter()
        u0h, sth = ctl.solve(x0, yref, yref_e)     # returns after the D2H copies completed
        if gather is not None:                     # device copies of the same outputs -> all ranks
            gather.launch(*gather.handle_outputs(), after=None)
        t1 = time.perf_counter()
        if i >= W:
            e2e_t.append(t1 - t0)
    if gather is not None:
        t0 = time.perf_counter()
        gather.drain()
        e2e_t.append(time.perf_counter() - t0)
    e2e_total = float(sum(e2e_t))
    clocks = sampler.stop() if rank == 0 else None

    if world > 1:
        t = torch.tensor([total_ms, e2e_total], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms, e2e_total = float(t[0]), float(t[1])
        c = torch.tensor([qp_sum, sqp_sum, ok_count, launches], dtype=torch.int64, device=dev)
        dist.all_reduce(c, op=dist.ReduceOp.SUM)
        qp_sum, sqp_sum, ok_count, launches = (int(v) for v in c)
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    n_solves = world * BATCH * K
    is32 = args.dtype == 'f32'
    value = n_solves / (total_ms * 1e-3)
    e2e_value = n_solves / e2e_total
    mean_qp = qp_sum / n_solves
    mean_sqp = sqp_sum / n_solves
    # roofline of the (single, fused) kernel; launch duration = CUDA-event time of the step
    launch_s = (total_ms * 1e-3) / K
    flops_per_launch = BATCH * (mean_sqp * F_LIN + mean_qp * F_IPM)
    bytes_per_launch = BATCH * BYTES_PER_SOLVE
    peaks = {}
    try:
        peaks = json.load(open(ROOT / 'MEASURED_PEAKS.json'))
    except Exception:
        pass
    hbm_peak = float(peaks.get('hbm_gbs', 6650.0))
    import ctypes as C
    tf = C.c_double(0.0)
    peak_fn = _abi.lib().nmpc_fp32_peak if is32 else _abi.lib().nmpc_fp64_peak
    _abi.check(peak_fn(local, C.byref(tf)), 'nmpc_fp32_peak' if is32 else 'nmpc_fp64_peak')
    traffic, traffic_src, executed = None, None, None
    try:   # dram__bytes_read.sum + dram__bytes_write.sum of ONE launch of this workload's kernel,
           # from the committed `ncu --set full` capture of this very command (written by
           # scripts/ncu_traffic.py; its "source" and "head" fields say which capture and kernel)
        tj = json.load(open(ROOT / 'profiles' / 'traffic.json'))
        if tj.get('dtype', 'f64') == args.dtype:
            traffic, traffic_src = tj['dram_bytes_per_launch'], tj.get('source')
            executed = tj.get('flops_executed_per_launch')   # thread-level FMA x2 + ADD + MUL of that capture
    except Exception:
        pass
    achieved_gbs = bytes_per_launch / launch_s * 1e-9
    achieved_tf = flops_per_launch / launch_s * 1e-12
    line = {
        'metric': 'NMPC solves/sec (hummingbird, N=20)', 'value': value, 'unit': 'solves/s',
        'n_gpus': world, 'steps': K, 'warmup': W, 'ms_per_step': total_ms / K,
        'p50_batch_latency_ms': float(np.median(step_ms)),
        'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': args.dtype,
        'data': 'synthetic', 'config': workload_config(world, args.dtype),
        'clocks': clocks,
        'e2e': {'value': e2e_value, 'unit': 'solves/s',
                'h2d_bytes_per_step': BATCH * (13 + HORIZON * 17 + 13) * 8,
                'd2h_bytes_per_step': BATCH * (3 * 4 * 8 + 4)},
        'gpu_launches': launches,
        # the bound that limits this kernel is the FP64 pipe (north_star: no tensor cores, FP64
        # FLOP/s and HBM GB/s against peak); the HBM roofline of the same launch is beside it
        'roofline': {'bound': 'fp32' if is32 else 'fp64', 'achieved': achieved_tf, 'peak': tf.value,
                     'unit': 'TFLOP/s', 'frac': achieved_tf / tf.value if tf.value else None,
                     'traffic': traffic, 'traffic_source': traffic_src,
                     'peak_source': ('nmpc_fp32_peak: FFMA' if is32 else 'nmpc_fp64_peak: DFMA')
                                    + ' micro-benchmark on this GPU, same run '
                                    '(MEASURED_PEAKS.json holds no vector-FMA figure)',
                     'flops_per_solve': flops_per_launch / BATCH,
                     'executed_flops_per_solve': executed / BATCH if executed else None,
                     'note': 'achieved = ALGORITHMIC flops (dense-block count of SURVEY 8d for the '
                             'reference algorithm: S linearisations + I factorisations with two vector '
                             'sweeps each, DESIGN.md section 4) / CUDA-event launch time; '
                             'executed_flops_per_solve = what the kernel really issued in the committed '
                             'ncu capture (constant Jacobian columns never multiplied, nominal '
                             'trajectory integrated once per stage, multipliers from the roll-out '
                             'instead of a costate sweep, checkpoint restarts)'},
        'roofline_hbm': {'bound': 'hbm', 'achieved': achieved_gbs, 'peak': hbm_peak, 'unit': 'GB/s',
                         'frac': achieved_gbs / hbm_peak, 'traffic': traffic,
                         'peak_source': 'MEASURED_PEAKS.json' if peaks else 'fallback',
                         'note': 'compulsory I/O only (8.7 KB/solve): the path is not HBM bound'},
        'solver_stats': {'mean_sqp_iters': mean_sqp, 'mean_riccati_factorisations': mean_qp,
                         'status_ok_frac': ok_count / n_solves},
        'wall_s_timed_region': wall,
    }
    if is32:
        line['parity'] = 'FP32 variant: u0 within 1e-3 relative of the FP64 oracle (tests/test_gpu_parity.py::test_fp32_variant), SQP step tolerance 1e-4'
    if gather is not None:
        line['nccl_ms_per_step'] = nccl_ms
        line['nccl'] = {'collective': 'all_gather_into_tensor', 'bytes_per_rank_per_step': BATCH * 40,
                        'stream': 'side stream, overlapped with the next step; drained inside the '
                                  'timed region', 'gathered_equals_local': gathered_ok}
    if world == 1 and not args.no_cpu:
        cores = os.cpu_count() or 1
        n = 256 * cores if cores <= 32 else 128 * cores
        v, w = cpu_sample(n, cores, seed=0)
        line['cpu_baseline'] = {
            'value': v, 'unit': 'solves/s', 'cores': cores, 'kind': 'port',
            'sample': f'{n} problems of the same workload (seed 0), oracle/ C port, one process '
                      f'per core, {w:.1f} s wall'}
    if world == 1 and not args.no_extras and not is32:
        line['extras'] = extras(params, dev)
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


def extras(params, dev):
    """Context numbers outside the headline workload (N=1 only): single RTI solves at large
    batch (north_star's 10^6 solves/s target is an RTI figure) and the closed-loop Monte Carlo
    of configs[2] at reduced length, all through the same kernel."""
    import copy
    import torch
    from rotors_mpc_controller_b200 import BatchedPositionNMPC, ClosedLoop, Plant
    from tests.problems import aggressive, random_states
    out = {}
    p = copy.deepcopy(params)
    p['solver']['max_sqp_iter'] = 1
    B = 65536
    ctl = BatchedPositionNMPC(p, batch=B, device=dev.index)
    for name in ('rti_cfg2_cold', 'rti_aggressive_cold'):
        rng = np.random.default_rng(0)
        if name == 'rti_cfg2_cold':
            arrs = make_inputs(0, B, ctl.hover_thrust)
        else:
            arrs = aggressive(rng, B, HORIZON, ctl.hover_thrust)
        X, Y, Z = (torch.as_tensor(a, device=dev) for a in arrs)
        ms = []
        for rep in range(6):
            ctl.invalidate()
            torch.cuda.synchronize(dev)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            ctl.solve(X, Y, Z)
            e1.record()
            torch.cuda.synchronize(dev)
            ms.append(e0.elapsed_time(e1))
        it = ctl.stats()['qp_iters']
        med = float(np.median(ms[1:]))
        out[name] = {'batch': B, 'solves_per_s': B / (med * 1e-3), 'ms': med,
                     'mean_factorisations': float(it.mean()),
                     'max_factorisations': int(it.max()),
                     'fp64_tflops_algorithmic': B * (F_LIN + float(it.mean()) * F_IPM) / (med * 1e-3) * 1e-12}
    # the headline workload at 16x the batch: the share of the batch-4096 launch that is the
    # tail of its longest problems (no work left to hide them behind) disappears
    ps = copy.deepcopy(params)
    cs = BatchedPositionNMPC(ps, batch=B, device=dev.index)
    X, Y, Z = (torch.as_tensor(a, device=dev) for a in make_inputs(0, B, cs.hover_thrust))
    ms = []
    for rep in range(3):
        cs.invalidate()
        torch.cuda.synchronize(dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        cs.solve(X, Y, Z)
        e1.record()
        torch.cuda.synchronize(dev)
        ms.append(e0.elapsed_time(e1))
    st = cs.stats()
    med = float(np.median(ms[1:]))
    out['sqp_cfg2_batch65536'] = {
        'batch': B, 'solves_per_s': B / (med * 1e-3), 'ms': med,
        'mean_factorisations': float(st['qp_iters'].mean()),
        'max_factorisations': int(st['qp_iters'].max()),
        'fp64_tflops_algorithmic': B * (float(st['sqp_iters'].mean()) * F_LIN
                                        + float(st['qp_iters'].mean()) * F_IPM) / (med * 1e-3) * 1e-12}
    cs.close()
    # closed loop: 65536 vehicles x 50 control steps, warm-started RTI, RK4 plant on the device
    rate = float(params.get('node', {}).get('rate', 60.0))     # the node re-solves at 60 Hz
    cl = ClosedLoop(ctl, Plant(dt=1.0 / rate, substeps=5))
    rng = np.random.default_rng(1)
    x = random_states(rng, B)
    sp = x[:, 0:3] + np.concatenate([rng.uniform(-2, 2, (B, 2)), rng.uniform(-0.5, 0.5, (B, 1))], 1)
    steps = 50
    cl.reset(x)
    cl.run_setpoint(5, sp)                      # warm-up
    cl.reset(x)
    torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    cl.run_setpoint(steps, sp)
    e1.record()
    torch.cuda.synchronize(dev)
    ms = e0.elapsed_time(e1)
    s = cl.summary(reduce=False)
    out['closed_loop_rti'] = {'vehicles': B, 'steps': steps, 'control_rate_hz': rate,
                              'solves_per_s': B * steps / (ms * 1e-3),
                              'ms_per_control_step': ms / steps,
                              'mean_factorisations': s['qp_iters'] / (B * steps),
                              'failed_solves': s['failed_solves']}
    ctl.close()
    # configs[0]: ONE vehicle through the reference's own call surface (PositionNMPC.solve with
    # state / reference dicts, host buffers), closed loop hover (0,0,1) -> set-point (2,0,1.5)
    from rotors_mpc_controller_b200 import PositionNMPC, ReferenceGenerator
    p1 = copy.deepcopy(params)
    p1['solver']['max_sqp_iter'] = 1
    one = PositionNMPC(p1, device=dev.index)
    ref = ReferenceGenerator(p1['reference'])
    ref.update_hover_thrust(one.hover_thrust)
    ref.set_target(position=[2.0, 0.0, 1.5], yaw=0.0)
    horizon = ref.build_horizon(one.horizon, one.dt)
    plant = ClosedLoop(one, Plant(dt=one.dt, substeps=5))     # nominal RK4 plant (device kernel)
    x = np.array([0, 0, 1, 0, 0, 0, 1, 0, 0, 0, 0, 0, 0], float)
    plant.reset(x[None])
    lat = []
    for step in range(120):
        state = {'position': x[0:3], 'velocity': x[3:6], 'quaternion': x[6:10], 'body_rates': x[10:13]}
        t0 = time.perf_counter()
        u0, status = one.solve(state, horizon)
        lat.append(time.perf_counter() - t0)
        thrust, _ = one.command()
        x = plant.plant_step(thrust[None]).cpu().numpy()[0]
    out['single_vehicle_rti'] = {
        'call': 'PositionNMPC.solve(state, reference) + command(), batch 1, host buffers',
        'p50_latency_us': float(np.median(lat[20:]) * 1e6), 'p99_latency_us': float(np.percentile(lat[20:], 99) * 1e6),
        'final_position': [float(v) for v in x[0:3]], 'last_status': int(status)}
    one.close()
    # the same closed loop through the CPU oracle's PositionNMPC-shaped wrapper on ONE host core:
    # configs[0] is the reference's own (latency) case, so this is the baseline beside it
    from oracle import oracle as O
    orc = O.OraclePositionNMPC(p1)
    x = np.array([0, 0, 1, 0, 0, 0, 1, 0, 0, 0, 0, 0, 0], float)
    lat_o = []
    for step in range(120):
        state = {'position': x[0:3], 'velocity': x[3:6], 'quaternion': x[6:10], 'body_rates': x[10:13]}
        t0 = time.perf_counter()
        u0, status = orc.solve(state, horizon)
        lat_o.append(time.perf_counter() - t0)
        thrust = np.clip(u0, *[np.asarray(b) for b in (orc.cfg.lbu[:], orc.cfg.ubu[:])])
        x = O.plant_step(orc.cfg, x, thrust, orc.cfg.dt, 5)
    out['single_vehicle_rti']['cpu_oracle_p50_latency_us'] = float(np.median(lat_o[20:]) * 1e6)
    out['single_vehicle_rti']['cpu_oracle_p99_latency_us'] = float(np.percentile(lat_o[20:], 99) * 1e6)
    out['single_vehicle_rti']['cpu_oracle_final_position'] = [float(v) for v in x[0:3]]
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=40)
    ap.add_argument('--warmup', type=int, default=5)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--dtype', default='f64', choices=['f64', 'f32'],
                    help='compute precision of the solve kernel (I/O stays FP64)')
    ap.add_argument('--no-cpu', action='store_true', help='skip the cpu_baseline leg')
    ap.add_argument('--no-extras', action='store_true',
                    help='skip the large-batch RTI / closed-loop context numbers')
    args = ap.parse_args()
    if args.impl == 'reference':
        return run_reference(args)
    return run_gpu(args)


if __name__ == '__main__':
    sys.exit(main())
