"""GPU tests of the handle's API surface beyond a plain solve: ragged batch sizes, fast
``reconfigure`` (constants swap, no code generation), run-time bounds, masked
invalidation, warm-start store round trip and the C ABI's error codes.  Checked against the CPU
oracle with the tolerances of tests/test_gpu_parity.py (FP64: |u0 - oracle| <= 1e-8, statuses
identical)."""
import copy
import ctypes as C

import numpy as np
import pytest

from tests.problems import aggressive, random_states, setpoint_refs

pytestmark = pytest.mark.gpu

TOL_U = 1e-8


@pytest.fixture(scope='module')
def O(built):
    from oracle import oracle
    return oracle


def _ctl(params, B):
    from rotors_mpc_controller_b200 import BatchedPositionNMPC
    return BatchedPositionNMPC(copy.deepcopy(params), batch=B)


@pytest.mark.parametrize('B', [1, 3, 5, 127, 130, 1777])
def test_ragged_batches(params, O, B):
    """Batches smaller than an SM's share, around one wave (1 776 resident warps) and odd sizes."""
    ctl = _ctl(params, B)
    rng = np.random.default_rng(B)
    x0, yref, yref_e = aggressive(rng, B, 20, ctl.hover_thrust)
    u0, st = ctl.solve(x0, yref, yref_e)
    ob = O.OracleBatch(O.config_from_params(params), B)
    ur, sr, thr, _ = ob.solve(x0, yref, yref_e)
    assert np.array_equal(st, sr)
    assert np.max(np.abs(u0 - ur)) < TOL_U
    assert np.max(np.abs(ctl.command()[0] - thr)) < TOL_U
    assert np.array_equal(ctl.stats()['qp_iters'], ob.qp_iters)
    ctl.close()


def test_reconfigure_swaps_constants(params, O):
    """controller.py:63-172 rebuilds and recompiles the solver on every slider move; here the same
    call swaps the constant block of the handle.  Results must follow the NEW parameters and the
    warm start must be dropped (controller.py:167)."""
    B = 96
    ctl = _ctl(params, B)
    rng = np.random.default_rng(5)
    x0 = random_states(rng, B)
    yref, yref_e = setpoint_refs(rng, x0, 20, ctl.hover_thrust)
    ctl.solve(x0, yref, yref_e)
    assert ctl.trajectories()[2].all()
    launches = ctl.kernel_launches

    p2 = copy.deepcopy(params)
    p2['solver']['position_weight'] = [25.0, 20.0, 4.0]
    p2['solver']['control_weight'] = [0.5, 0.6, 0.7, 0.8]
    p2['solver']['terminal_weight'] = [float(v) * 1.5 for v in np.arange(1, 14)]
    p2['solver']['regularization'] = 2.0e-2
    p2['solver']['iter_max'] = 80
    p2['vehicle']['mass'] = 0.9
    p2['vehicle']['motor_max_speed'] = 700.0
    ctl.reconfigure(p2)
    assert not ctl.trajectories()[2].any()              # warm start invalidated
    assert abs(ctl.hover_thrust - 0.9 * ctl.gravity / 4.0) < 1e-12
    yref[:, :, 13:17] = ctl.hover_thrust
    u0, st = ctl.solve(x0, yref, yref_e)
    assert ctl.kernel_launches == launches + 2          # same handle, no rebuild: invalidate + solve
    ob = O.OracleBatch(O.config_from_params(p2), B)
    ur, sr, thr, spr = ob.solve(x0, yref, yref_e)
    assert np.array_equal(st, sr)
    assert np.max(np.abs(u0 - ur)) < TOL_U
    assert np.max(np.abs(ctl.command()[0] - thr)) < TOL_U
    lo, hi = ctl.input_bounds
    assert np.allclose(hi, ctl.rotor_force_constant * 700.0 ** 2)

    # a different horizon needs new buffers: the handle is replaced behind the same object
    p3 = copy.deepcopy(p2)
    p3['solver']['horizon_steps'] = 12
    ctl.reconfigure(p3)
    assert ctl.horizon == 12
    yr, yre = setpoint_refs(rng, x0, 12, ctl.hover_thrust)
    u0, st = ctl.solve(x0, yr, yre)
    ob = O.OracleBatch(O.config_from_params(p3), B)
    ur, sr, _, _ = ob.solve(x0, yr, yre)
    assert np.array_equal(st, sr)
    assert np.max(np.abs(u0 - ur)) < TOL_U
    ctl.close()


def test_set_bounds_at_run_time(params, O):
    B = 64
    ctl = _ctl(params, B)
    rng = np.random.default_rng(6)
    x0, yref, yref_e = aggressive(rng, B, 20, ctl.hover_thrust)
    ctl.set_bounds(0.5, 3.0)
    u0, st = ctl.solve(x0, yref, yref_e)
    ob = O.OracleBatch(O.config_from_params(params, thrust_box=(0.5, 3.0)), B)
    ur, sr, thr, _ = ob.solve(x0, yref, yref_e)
    assert np.array_equal(st, sr)
    assert np.max(np.abs(u0 - ur)) < TOL_U
    ok = st == 0
    assert ok.any()
    assert u0[ok].min() >= 0.5 - 1e-9 and u0[ok].max() <= 3.0 + 1e-9
    assert (np.abs(u0[ok] - 0.5) < 1e-6).any() or (np.abs(u0[ok] - 3.0) < 1e-6).any()
    from rotors_mpc_controller_b200 import _abi
    with pytest.raises(_abi.NmpcError):
        ctl.set_bounds(2.0, 1.0)                          # empty box
    ctl.close()


def test_masked_invalidate_and_store_round_trip(params, O):
    """nmpc_invalidate(mask) is the batched form of `_prev_solution_valid = False`
    (controller.py:449, :167); get/set_trajectories move the warm-start store between handles."""
    B = 80
    ctl = _ctl(params, B)
    cfg = O.config_from_params(params)
    ob = O.OracleBatch(cfg, B)
    rng = np.random.default_rng(7)
    x0 = random_states(rng, B)
    yref, yref_e = setpoint_refs(rng, x0, 20, ctl.hover_thrust)
    ctl.solve(x0, yref, yref_e)
    ob.solve(x0, yref, yref_e)
    mask = (np.arange(B) % 3 == 0).astype(np.uint8)
    ctl.invalidate(mask)
    ob.valid[mask.astype(bool)] = 0
    assert np.array_equal(ctl.trajectories()[2], ob.valid.astype(bool))
    xs, us, vs = ctl.trajectories()
    twin = _ctl(params, B)
    twin.set_trajectories(xs, us, vs)
    x1 = x0 + 0.01 * rng.normal(size=x0.shape)
    x1[:, 6:10] /= np.linalg.norm(x1[:, 6:10], axis=1, keepdims=True)
    u0, st = ctl.solve(x1, yref, yref_e)
    u0 = u0.copy()
    ur, sr, _, _ = ob.solve(x1, yref, yref_e)
    assert np.array_equal(st, sr)
    assert np.max(np.abs(u0 - ur)) < TOL_U
    ut, stt = twin.solve(x1, yref, yref_e)
    assert np.array_equal(ut, u0) and np.array_equal(stt, st)   # same store -> same bits
    ctl.close()
    twin.close()


def test_device_mask_invalidate_and_stream_scope(params, O):
    """nmpc_invalidate_device: a device mask on the caller's stream, no host synchronisation.
    nmpc_invalidate / nmpc_get_* are stream-scoped: they wait for THEIR handle's last kernel
    only, so a long solve in flight on another handle's stream is not waited for."""
    import time

    import torch
    dev = torch.device('cuda', 0)
    B = 64
    ctl = _ctl(params, B)
    rng = np.random.default_rng(11)
    x0 = random_states(rng, B)
    yref, yref_e = setpoint_refs(rng, x0, 20, ctl.hover_thrust)
    ctl.solve(x0, yref, yref_e)
    mask = torch.as_tensor((np.arange(B) % 4 == 1).astype(np.uint8), device=dev)
    ctl.invalidate(mask)                                   # device mask, async
    valid = ctl.trajectories()[2]                          # ordered behind the kernel
    assert np.array_equal(valid, np.arange(B) % 4 != 1)
    with pytest.raises(ValueError):
        ctl.invalidate(torch.zeros(B + 1, dtype=torch.uint8, device=dev))
    # a long solve on another handle / another stream (N = 600, one CTA per SM: the persistent
    # kernel leaves room on every SM for the small handle's kernels) ...
    p = copy.deepcopy(params)
    p['solver'].update(horizon_steps=600)
    from rotors_mpc_controller_b200 import BatchedPositionNMPC
    big = BatchedPositionNMPC(p, batch=148)
    xb = random_states(rng, 148)
    yb, yeb = setpoint_refs(rng, xb, 600, big.hover_thrust)
    xb, yb, yeb = (torch.as_tensor(a, device=dev) for a in (xb, yb, yeb))
    side = torch.cuda.Stream(dev)
    big.solve(xb, yb, yeb)      # first launch of this instantiation (lazy module load) stays outside
    big.invalidate()
    torch.cuda.synchronize(dev)
    t0 = time.perf_counter()
    with torch.cuda.stream(side):
        big.solve(xb, yb, yeb)
    # ... does not serialise the small handle's host-side calls
    ctl.invalidate(np.ones(B, np.uint8))
    ctl.invalidate()
    ctl.stats()
    t_small = time.perf_counter() - t0
    side.synchronize()
    t_big = time.perf_counter() - t0
    assert not ctl.trajectories()[2].any()
    assert t_small < 0.5 * t_big, (t_small, t_big)
    big.close()
    ctl.close()


@pytest.mark.parametrize('case', ['rti_warm', 'sqp', 'n15', 'n600', 'f32', 'aggressive'])
def test_poisoned_memory_changes_nothing(params, case):
    """Stand-in for compute-sanitizer initcheck (closed on this pool): with the workspace NaN-filled
    before every launch and the shared tile NaN-filled before every problem, every kernel variant
    returns bit-identical results -- no problem reads memory it has not written itself (stale data
    of the previous problem on the same warp, uninitialised padding, another warp's slot)."""
    over = {'rti_warm': {}, 'sqp': dict(max_sqp_iter=50, sqp_tol=1e-8),
            'n15': dict(horizon_steps=15), 'n600': dict(horizon_steps=600),
            'f32': dict(precision='f32'), 'aggressive': {}}[case]
    B = 8 if case == 'n600' else 300
    p = copy.deepcopy(params)
    p['solver'].update(over)
    from rotors_mpc_controller_b200 import BatchedPositionNMPC
    N = p['solver']['horizon_steps']
    rng = np.random.default_rng(21)
    res = []
    for poison in (False, True):
        ctl = BatchedPositionNMPC(p, batch=B)
        ctl.debug_poison(poison)
        r = np.random.default_rng(21)
        if case in ('aggressive', 'f32'):
            x0, yref, yref_e = aggressive(r, B, N, ctl.hover_thrust)
        else:
            x0 = random_states(r, B)
            yref, yref_e = setpoint_refs(r, x0, N, ctl.hover_thrust)
        outs = []
        for rep in range(2 if case != 'sqp' else 1):        # second solve is warm-started
            u0, st = ctl.solve(x0, yref, yref_e)
            outs += [u0.copy(), st.copy()]
        xs, us, vs = ctl.trajectories()
        outs += [xs, us, vs, ctl.stats()['qp_iters']]
        res.append(outs)
        ctl.close()
    for a, b in zip(*res):
        assert np.array_equal(a, b, equal_nan=True)
    assert (res[0][1] == 0).mean() > 0.9


def test_abi_error_codes(params, built):
    """Every entry point answers a bad argument with NMPC_E_INVALID instead of crashing; a
    per-problem solver failure is data (status), never an error code."""
    from rotors_mpc_controller_b200 import _abi
    from rotors_mpc_controller_b200.controller import abi_config, controller_params_from_dict
    lib = _abi.lib()
    cfg = abi_config(controller_params_from_dict(params), params['solver'])
    h = C.c_void_p()
    assert lib.nmpc_create(C.byref(cfg), 0, 0, C.byref(h)) == -1            # empty batch
    assert lib.nmpc_create(C.byref(cfg), 4, 99, C.byref(h)) == -1           # no such device
    assert lib.nmpc_create(None, 4, 0, C.byref(h)) == -1
    bad = copy.copy(cfg)
    bad.N = 601                                                               # slider maximum is 600
    assert lib.nmpc_create(C.byref(bad), 4, 0, C.byref(h)) == -1
    bad = copy.copy(cfg)
    bad.lbu[0] = bad.ubu[0] + 1.0
    assert lib.nmpc_create(C.byref(bad), 4, 0, C.byref(h)) == -1
    assert lib.nmpc_create(C.byref(cfg), 4, 0, C.byref(h)) == 0
    out = np.zeros(64)
    st = np.zeros(4, np.int32)
    assert lib.nmpc_solve_host(h, None, _abi.ptr(out), _abi.ptr(out), _abi.ptr(out),
                               _abi.ptr(st), None, None) == -1
    assert lib.nmpc_solve_host(None, _abi.ptr(out), _abi.ptr(out), _abi.ptr(out), _abi.ptr(out),
                               _abi.ptr(st), None, None) == -1
    other = copy.copy(cfg)
    other.N = cfg.N + 1
    assert lib.nmpc_reconfigure(h, C.byref(other)) == -1                     # shape change
    assert lib.nmpc_set_bounds(h, None, None) == -1
    assert lib.nmpc_batch(h) == 4 and lib.nmpc_horizon(h) == cfg.N
    assert lib.nmpc_batch(None) == -1
    assert b'invalid' in lib.nmpc_strerror(-1)
    assert lib.nmpc_destroy(h) == 0


def test_sqp_ipm_warm_start_is_a_pure_speed_up(params, O):
    """solver.ipm_warm_start (default on) only changes where the interior point of the 2nd, 3rd,
    ... QP of an SQP run starts: every QP is still solved to the same tolerance, so the converged
    iterate agrees with the cold-start run to the SQP tolerance (1e-8 step) and the statuses are
    equal, with a third fewer Riccati factorisations.  Both variants match the oracle run with
    the same flag (iteration counts included); RTI solves are untouched."""
    from rotors_mpc_controller_b200 import BatchedPositionNMPC
    B = 384
    rng = np.random.default_rng(21)
    x0 = random_states(rng, B)
    out = {}
    for warm in (False, True):
        p = copy.deepcopy(params)
        p['solver'].update(max_sqp_iter=50, sqp_tol=1e-8, ipm_warm_start=warm,
                           active_set_shortcut=False)     # isolate the interior-point warm start
        ctl = BatchedPositionNMPC(p, batch=B)
        if not warm:
            yref, yref_e = setpoint_refs(np.random.default_rng(22), x0, 20, ctl.hover_thrust)
        u0, st = ctl.solve(x0, yref, yref_e)
        s = ctl.stats()
        ob = O.OracleBatch(O.config_from_params(p, max_sqp_iter=50, sqp_tol=1e-8), B)
        ur, sr, _, _ = ob.solve(x0, yref, yref_e)
        assert np.array_equal(st, sr)
        assert np.max(np.abs(u0 - ur)) < 1e-7             # two SQP runs stopping at step <= 1e-8
        same = s['sqp_iters'] > 0
        assert np.mean(s['qp_iters'] == ob.qp_iters) > 0.98   # same path up to rounding ties
        out[warm] = (u0.copy(), st.copy(), s['qp_iters'].copy(), s['sqp_iters'].copy())
        ctl.close()
        assert same.all()
    assert np.array_equal(out[True][1], out[False][1]) and np.all(out[True][1] == 0)
    assert np.max(np.abs(out[True][0] - out[False][0])) < 1e-7
    assert out[True][2].sum() < 0.75 * out[False][2].sum()
    assert out[True][2].max() < 0.75 * out[False][2].max()
    assert np.mean(out[True][3] == out[False][3]) > 0.95      # same number of SQP iterations


@pytest.mark.parametrize('kind', ['cfg2', 'aggressive'])
def test_active_set_shortcut_is_a_pure_speed_up(params, O, kind):
    """solver.active_set_shortcut (default on): inside one SQP run a QP first tries the previous
    QP's active set -- an exact equality-constrained solve in ONE factorisation, accepted only if it
    is primal and dual feasible (then it is the unique QP solution).  Same statuses, same SQP
    iteration counts, converged u0 equal to the interior-point-only run within the SQP tolerance,
    far fewer factorisations; the oracle with the same flag takes the same path."""
    from rotors_mpc_controller_b200 import BatchedPositionNMPC
    B = 384
    out = {}
    for on in (False, True):
        p = copy.deepcopy(params)
        p['solver'].update(max_sqp_iter=50, sqp_tol=1e-8, active_set_shortcut=on)
        ctl = BatchedPositionNMPC(p, batch=B)
        rng = np.random.default_rng(31)
        if kind == 'cfg2':
            x0 = random_states(rng, B)
            yref, yref_e = setpoint_refs(rng, x0, 20, ctl.hover_thrust)
        else:
            x0, yref, yref_e = aggressive(rng, B, 20, ctl.hover_thrust)
        u0, st = ctl.solve(x0, yref, yref_e)
        s = ctl.stats()
        ob = O.OracleBatch(O.config_from_params(p, max_sqp_iter=50, sqp_tol=1e-8), B)
        ur, sr, _, _ = ob.solve(x0, yref, yref_e)
        assert np.array_equal(st, sr)
        ok = st == 0
        assert np.max(np.abs(u0 - ur)[ok]) < 1e-7
        assert np.mean(s['qp_iters'] == ob.qp_iters) > 0.97    # same path up to rounding ties
        out[on] = (u0.copy(), st.copy(), s['qp_iters'].copy(), s['sqp_iters'].copy())
        ctl.close()
    assert np.array_equal(out[True][1], out[False][1])
    ok = out[True][1] == 0
    assert ok.mean() > (0.99 if kind == 'cfg2' else 0.3)
    assert np.max(np.abs(out[True][0] - out[False][0])[ok]) < 1e-7
    assert np.mean(out[True][3][ok] == out[False][3][ok]) > 0.95    # same number of SQP iterations
    assert out[True][2][ok].sum() < (0.7 if kind == 'cfg2' else 0.9) * out[False][2][ok].sum()


def test_fp32_sqp_with_warm_start(params):
    """FP32 variant in SQP mode (tolerance an FP32 iterate can reach): statuses equal to the FP64
    run, u0 within north_star's 1e-3 relative, warm start does not cost accuracy."""
    from rotors_mpc_controller_b200 import BatchedPositionNMPC
    B = 2048
    x0 = random_states(np.random.default_rng(0), B)
    res = {}
    for prec in ('f64', 'f32'):
        p = copy.deepcopy(params)
        p['solver'].update(max_sqp_iter=30, sqp_tol=1e-4, precision=prec)
        ctl = BatchedPositionNMPC(p, batch=B)
        if not res:
            yref, yref_e = setpoint_refs(np.random.default_rng(1), x0, 20, ctl.hover_thrust)
        u0, st = ctl.solve(x0, yref, yref_e)
        res[prec] = (u0.copy(), st.copy(), ctl.stats()['qp_iters'].mean())
        ctl.close()
    assert np.array_equal(res['f32'][1], res['f64'][1]) and np.all(res['f64'][1] == 0)
    rel = np.abs(res['f32'][0] - res['f64'][0]).max() / np.abs(res['f64'][0]).max()
    assert rel < 1e-3
    assert res['f32'][2] < 1.2 * res['f64'][2]        # no extra interior-point iterations


def test_tma_staging_matches_direct_loads(params, O, monkeypatch):
    """The solve kernel stages yref / the warm-start rows with TMA bulk copies when the horizon is
    even and the caller's yref is 16-byte aligned, and reads them directly otherwise.  Same
    problems through (a) the staged path, (b) a device yref that is only 8-byte aligned (falls
    back to direct loads), (c) the NMPC_B200_NO_TMA handle: bit-identical results, cold and warm
    (odd problem indices exercise the 8-byte offset of the staged warm-start rows)."""
    import torch
    B = 257
    rng = np.random.default_rng(11)
    ctl = _ctl(params, B)
    x0, yref, yref_e = aggressive(rng, B, 20, ctl.hover_thrust)
    dev = torch.device('cuda', 0)
    X, Z = torch.as_tensor(x0, device=dev), torch.as_tensor(yref_e, device=dev)
    Ya = torch.as_tensor(yref, device=dev)
    assert Ya.data_ptr() % 16 == 0
    raw = torch.empty(yref.size + 1, dtype=torch.float64, device=dev)
    Ym = raw[1:].view(B, 20, 17)
    Ym.copy_(Ya)
    assert Ym.data_ptr() % 16 == 8 and Ym.is_contiguous()

    def run(c, Y):
        out = []
        for _ in range(3):  # cold, then two warm-started solves
            u0, st = c.solve(X, Y, Z)
            out.append((u0.cpu().numpy().copy(), st.cpu().numpy().copy(),
                        c.stats()['qp_iters'].copy(), c.trajectories()[0].copy()))
        return out

    staged = run(ctl, Ya)
    ctl.invalidate()
    direct = run(ctl, Ym)
    ctl.close()
    monkeypatch.setenv('NMPC_B200_NO_TMA', '1')
    ctl2 = _ctl(params, B)
    monkeypatch.delenv('NMPC_B200_NO_TMA')
    knob = run(ctl2, Ya)
    ctl2.close()
    for a, b, c in zip(staged, direct, knob):
        for i in range(4):
            assert np.array_equal(a[i], b[i]) and np.array_equal(a[i], c[i])
    ob = O.OracleBatch(O.config_from_params(params), B)
    ur, sr, _, _ = ob.solve(x0, yref, yref_e)
    assert np.array_equal(staged[0][1], sr)
    assert np.max(np.abs(staged[0][0] - ur)) < TOL_U


def test_odd_horizon_uses_direct_loads(params, O):
    """Odd N: yref rows are not 16-byte multiples, staging is off; results still match the oracle."""
    p = copy.deepcopy(params)
    p['solver']['horizon_steps'] = 15
    B = 64
    ctl = _ctl(p, B)
    rng = np.random.default_rng(12)
    x0 = random_states(rng, B)
    yref, yref_e = setpoint_refs(rng, x0, 15, ctl.hover_thrust)
    u0, st = (a.copy() for a in ctl.solve(x0, yref, yref_e))   # the handle reuses its output arrays
    u1, st1 = ctl.solve(x0, yref, yref_e)                      # warm-started second solve
    ob = O.OracleBatch(O.config_from_params(p), B)
    ur, sr, _, _ = ob.solve(x0, yref, yref_e)
    assert np.array_equal(st, sr)
    assert np.max(np.abs(u0 - ur)) < TOL_U
    ur1, sr1, _, _ = ob.solve(x0, yref, yref_e)
    assert np.array_equal(st1, sr1)
    assert np.max(np.abs(u1 - ur1)) < TOL_U
    ctl.close()


def test_fixed_horizon_kernel_matches_generic(params, monkeypatch):
    """N = 20 with the midpoint integrator runs the kernel instantiation with the horizon folded
    in at compile time; NMPC_B200_GENERIC_N=1 (read at nmpc_create) forces the run-time-N
    instantiation that every other horizon uses.  Same arithmetic in the same order: bit-identical
    results, cold and warm, RTI and SQP."""
    B = 300
    rng = np.random.default_rng(21)
    p = copy.deepcopy(params)
    p['solver']['max_sqp_iter'] = 50
    p['solver']['sqp_tol'] = 1e-8
    outs = []
    for generic in (False, True):
        if generic:
            monkeypatch.setenv('NMPC_B200_GENERIC_N', '1')
        res = []
        for prm in (params, p):
            ctl = _ctl(prm, B)
            r = np.random.default_rng(21)
            x0, yref, yref_e = aggressive(r, B, 20, ctl.hover_thrust)
            for _ in range(2):
                u0, st = ctl.solve(x0, yref, yref_e)
                res.append((u0.copy(), st.copy(), ctl.stats()['qp_iters'].copy(),
                            ctl.trajectories()[0].copy()))
            ctl.close()
        outs.append(res)
        if generic:
            monkeypatch.delenv('NMPC_B200_GENERIC_N')
    for a, b in zip(*outs):
        for i in range(4):
            assert np.array_equal(a[i], b[i])
