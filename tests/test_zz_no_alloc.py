"""GPU test of ``nmpc_solve_static`` -- the allocation-free entry point whose scratch memory is a
static ``__device__`` array inside libnmpc_b200.so (include/nmpc_b200.h; no counterpart in the
reference).  It is called through the C ABI (ctypes) and compared with the existing
``nmpc_create`` + ``nmpc_solve_device`` path on the same inputs: same kernel, same arithmetic per
problem, so the results are expected to be identical; the assertion uses the tolerance of the
other GPU tests (|u0 - reference| <= 1e-8, statuses and iteration counts identical).

Collected last on purpose (file name).  Every test here
makes TWO calls in a row (cold start, then warm start from the caller-owned store) -- the second
call finds the scratch array as the first one left it.
"""
import copy
import ctypes as C

import numpy as np
import pytest

from tests.problems import aggressive, random_states, setpoint_refs

pytestmark = pytest.mark.gpu

TOL_U = 1e-8


class _Store:
    """Caller-owned warm-start store + outputs of the allocation-free path (torch tensors are only
    the device memory the CALLER brings; the entry point itself allocates nothing)."""

    def __init__(self, torch, B, N, hover):
        dev = torch.device('cuda', 0)
        f64 = dict(dtype=torch.float64, device=dev)
        self.xtraj = torch.zeros((B, N + 1, 13), **f64)
        self.utraj = torch.zeros((B, N, 4), **f64)
        self.valid = torch.zeros((B,), dtype=torch.uint8, device=dev)
        self.last_cmd = torch.full((B, 4), hover, **f64)
        self.u0 = torch.empty((B, 4), **f64)
        self.status = torch.empty((B,), dtype=torch.int32, device=dev)
        self.thrust = torch.empty((B, 4), **f64)
        self.speed = torch.empty((B, 4), **f64)
        self.qp = torch.empty((B,), dtype=torch.int32, device=dev)
        self.sqp = torch.empty((B,), dtype=torch.int32, device=dev)


def _solve_static(torch, cfg, B, S, x0, yref, yref_e):
    from rotors_mpc_controller_b200 import _abi
    stream = torch.cuda.current_stream().cuda_stream
    rc = _abi.lib().nmpc_solve_static(
        C.byref(cfg), B, 0, x0.data_ptr(), yref.data_ptr(), yref_e.data_ptr(), S.xtraj.data_ptr(),
        S.utraj.data_ptr(), S.valid.data_ptr(), S.last_cmd.data_ptr(), S.u0.data_ptr(),
        S.status.data_ptr(), S.thrust.data_ptr(), S.speed.data_ptr(), S.qp.data_ptr(),
        S.sqp.data_ptr(), stream)
    _abi.check(rc, 'nmpc_solve_static')
    torch.cuda.synchronize()


def _compare(ctl, S, u0, st):
    """handle path (ctl, u0, st) vs. static path (S): outputs, statistics and the warm-start store."""
    assert np.array_equal(S.status.cpu().numpy(), st.cpu().numpy())
    assert np.max(np.abs(S.u0.cpu().numpy() - u0.cpu().numpy())) <= TOL_U
    thr, spd = ctl.command()
    assert np.max(np.abs(S.thrust.cpu().numpy() - thr.cpu().numpy())) <= TOL_U
    assert np.max(np.abs(S.speed.cpu().numpy() - spd.cpu().numpy())) <= 1e-6   # rad/s, ~1e3 scale
    stats = ctl.stats()
    assert np.array_equal(S.qp.cpu().numpy(), stats['qp_iters'])
    assert np.array_equal(S.sqp.cpu().numpy(), stats['sqp_iters'])
    x, u, v = ctl.trajectories()
    assert np.array_equal(S.valid.cpu().numpy().astype(bool), v)
    ok = v
    assert np.max(np.abs(S.xtraj.cpu().numpy()[ok] - x[ok])) <= TOL_U
    assert np.max(np.abs(S.utraj.cpu().numpy()[ok] - u[ok])) <= TOL_U


def _two_calls(params, B, N, seed, scenario):
    import torch
    from rotors_mpc_controller_b200 import BatchedPositionNMPC
    ctl = BatchedPositionNMPC(copy.deepcopy(params), batch=B)
    S = _Store(torch, B, N, ctl.hover_thrust)
    rng = np.random.default_rng(seed)
    dev = torch.device('cuda', 0)
    for call in range(2):   # cold start, then warm start from the stores the first call wrote
        if scenario == 'aggressive':
            x0, yref, yref_e = aggressive(rng, B, N, ctl.hover_thrust)
        else:
            x0 = random_states(rng, B)
            yref, yref_e = setpoint_refs(rng, x0, N, ctl.hover_thrust)
        x0, yref, yref_e = (torch.as_tensor(a, device=dev).contiguous() for a in (x0, yref, yref_e))
        u0, st = ctl.solve(x0, yref, yref_e)
        torch.cuda.synchronize()
        _solve_static(torch, ctl._cfg, B, S, x0, yref, yref_e)
        _compare(ctl, S, u0, st)
        assert int((st == 0).sum()) > B // 2   # the comparison is about real solves
    ctl.close()


def test_static_rti_reference_configuration(params, built):
    """configs[0] (the reference's RTI step, N = 20, FP64); batch above the 624 warps the scratch
    array holds at N = 20, so the work queue hands out several problems per warp."""
    _two_calls(params, 2500, 20, 11, 'aggressive')


def test_static_sqp_to_convergence(params, built):
    """configs[1] (SQP to convergence): many QPs per problem through the same scratch slots."""
    p = copy.deepcopy(params)
    p['solver']['max_sqp_iter'] = 50
    _two_calls(p, 700, 20, 12, 'setpoint')


def test_static_other_horizon_and_fp32(params, built):
    """Generic-horizon instantiation (N = 40) and the FP32 kernel."""
    p = copy.deepcopy(params)
    p['solver']['horizon_steps'] = 40
    _two_calls(p, 300, 40, 13, 'setpoint')
    p = copy.deepcopy(params)
    p['solver']['precision'] = 'f32'
    _two_calls(p, 300, 20, 14, 'setpoint')


def test_static_rejects_sizes_beyond_the_maximum(params, built):
    import torch
    from rotors_mpc_controller_b200 import BatchedPositionNMPC, _abi
    ctl = BatchedPositionNMPC(copy.deepcopy(params), batch=4)
    S = _Store(torch, 4, 20, ctl.hover_thrust)
    dev = torch.device('cuda', 0)
    x0 = torch.zeros((4, 13), dtype=torch.float64, device=dev)
    yref = torch.zeros((4, 20, 17), dtype=torch.float64, device=dev)
    lib = _abi.lib()

    def call(cfg, B, xt=S.xtraj.data_ptr()):
        return lib.nmpc_solve_static(C.byref(cfg), B, 0, x0.data_ptr(), yref.data_ptr(),
                                     x0.data_ptr(), xt, S.utraj.data_ptr(), S.valid.data_ptr(),
                                     S.last_cmd.data_ptr(), None, None, None, None, None, None, None)

    assert call(ctl._cfg, _abi.STATIC_MAX_BATCH + 1) == _abi.E_INVALID
    assert b'NMPC_STATIC_MAX_BATCH' in lib.nmpc_last_error()
    assert call(ctl._cfg, 0) == _abi.E_INVALID
    assert call(ctl._cfg, 4, None) == _abi.E_INVALID          # the warm-start store is required
    bad = type(ctl._cfg).from_buffer_copy(ctl._cfg)
    bad.N = 601                                               # beyond the supported horizon
    assert call(bad, 4) == _abi.E_INVALID
    torch.cuda.synchronize()
    ctl.close()
