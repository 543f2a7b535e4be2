// nmpc_abi.cu -- kernels + C ABI of libnmpc_b200.so (see include/nmpc_b200.h).
// Built for sm_100a only; there is no CPU fallback: without a CUDA device nmpc_create fails.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>
#include <new>

#include "../../include/nmpc_b200.h"
#include "nmpc_solver.cuh"
#include "nmpc_closedloop.cuh"

using namespace nmpc;

namespace {

// CTAs of four independent warps (warps never synchronise with each other; 4 CTAs of 4 x 10.6 KB
// tiles fit the 196 KB shared-memory configuration).  The register file is partitioned per
// scheduler (16 K registers each): 4 warps per scheduler = 16 problems in flight per SM need <= 128
// registers, 3 per scheduler fit up to 170.  History of the FP64 cap: 160 (12 warps per SM) through
// most of round 2 -- at 128 the sweeps spilled and lost more than the extra warp brought (7.91 vs
// 7.60 ms on the bench step); once the linearisation and the sweeps had shed their instructions the
// 128-register build spills 8 bytes and wins (bench 4.73 vs 4.82 ms, large RTI batches +2-4 %).
#ifndef NMPC_MAXNREG
#define NMPC_MAXNREG 128
#endif
// (the run-time-horizon instantiations spill 104 bytes at 128 registers: they keep 160 / 12 warps per SM)
#ifndef NMPC_MAXNREG_GENERIC
#define NMPC_MAXNREG_GENERIC 160
#endif
#ifndef NMPC_MAXNREG_F32
#define NMPC_MAXNREG_F32 128
#endif
constexpr int WARPS = CTA_WARPS;  // warps (= problems in flight) per CTA
constexpr size_t SMALL_HOST_BYTES = 256 * 1024;  // nmpc_solve_host packs batches up to this input size
constexpr size_t ZERO_COPY_BYTES = 16 * 1024;    // ... and up to this size the kernel reads / writes the
                                                 // pinned mirrors directly (no copy nodes on the stream)

constexpr int smem_elems_per_warp() { return smem_tile_elems(); }

// ---------------------------------------------------------------- kernels
// NF > 0: instantiation for a horizon known at compile time (see fixed_layout); NF = 0: any N.
template <typename T, int NS, int NF>
__global__ void __maxnreg__(sizeof(T) == 4 ? NMPC_MAXNREG_F32 : (NF > 0 ? NMPC_MAXNREG : NMPC_MAXNREG_GENERIC))
    nmpc_solve_kernel(const __grid_constant__ SolveArgs<T> A, const __grid_constant__ WsLayout Lr,
                      int* __restrict__ work_counter) {
  const WsLayout Lfix_ = NF > 0 ? make_layout(NF > 0 ? NF : 1) : WsLayout();
  const WsLayout& L = NF > 0 ? Lfix_ : Lr;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  T* sm = reinterpret_cast<T*>(smem_raw) + warp * smem_elems_per_warp();
  if (A.stage && lane == 0) {  // mbarrier of the TMA input staging + its phase parity
    reinterpret_cast<unsigned*>(sm + SmemLayout::bar)[2] = 0u;
    mbar_init(sm + SmemLayout::bar);
  }
  {  // constants shared by the CTA (see VBLOCK_ELEMS)
    T* vblock = reinterpret_cast<T*>(smem_raw) + WARPS * smem_elems_per_warp();
    for (int e = threadIdx.x; e < VBLOCK_ELEMS; e += blockDim.x) {
      const int i = e / M_LD, a = e - i * M_LD;
      vblock[e] = (a < 3) ? ((i == a) ? A.cfg.dt : ((i == 3 + a) ? T(1) : T(0))) : T(0);
    }
    __syncthreads();  // the only CTA-wide barrier: afterwards the warps run independently
  }
  const int slot = blockIdx.x * WARPS + warp;
  T* ws = A.ws + (size_t)slot * L.total;
  // Work distribution.  First round: warp w of CTA b takes queue position w * gridDim.x + b, so a
  // batch smaller than the number of resident warps spreads evenly over the SMs (and schedulers)
  // instead of filling the first CTAs.  Afterwards a dynamic queue: iteration counts differ per
  // problem, so problems are handed out one at a time instead of in a fixed stride.
  // (Measured and dropped: positions dealt out so that every scheduler gets one problem of the
  // longest, the middle and the shortest third of the first round -- no change, the queue keeps
  // every scheduler full until it is empty either way; and the block scheduler does not place CTA b
  // on SM b mod 148, see scripts/micro/cta_placement.cu.)
  int pos = warp * gridDim.x + blockIdx.x;
  while (pos < A.batch) {
    int prob = pos;
    if (A.order) {
      if (lane == 0) prob = A.order[pos];
      prob = __shfl_sync(FULL, prob, 0);
    }
    solve_one<T, NS, NF>(A, L, ws, sm, prob, lane);
    if (A.batch <= (int)gridDim.x * WARPS) break;  // one round: the queue is not used (nor reset by the host)
    if (lane == 0) pos = atomicAdd(work_counter, 1) + gridDim.x * WARPS;
    pos = __shfl_sync(FULL, pos, 0);
  }
}

// Longest-expected-first dispatch.  The work of a problem (SQP iterations x factorisations per
// QP) varies by an order of magnitude inside a batch, and a launch with only a few problems per
// resident warp ends when its longest problem does: handing the long ones out FIRST shortens the
// tail.  Nobody knows the iteration count in advance; the value of the problem's own objective at
// the start (controller.py:223-245, state part:  N (x0 - yref_0)' W (x0 - yref_0) + (x0 - yref_e)'
// W_e (x0 - yref_e), the cost of staying put over the horizon) ranks it well enough (Spearman 0.6 with the
// factorisation count on configs[1]).  Keys -> 1024 logarithmic bins (whole grid) -> counting sort in
// one CTA, largest first; the order inside a bin is arbitrary (results do not depend on the order).
constexpr int ORDER_BINS = 1024;
template <typename T>
__device__ __forceinline__ int order_bin(const DevCfg<T>& c, const double* __restrict__ x0,
                                         const double* __restrict__ y0, const double* __restrict__ ye,
                                         double nstages) {
  double key = 0.0;
#pragma unroll
  for (int i = 0; i < NX; i++) {
    const double xi = x0[i], e0 = xi - y0[i], e1 = xi - ye[i];
    key += nstages * (double)c.Wq[i] * e0 * e0 + (double)c.We[i] * e1 * e1;
  }
  if (!(key == key) || key > 1e30) return ORDER_BINS - 1;   // NaN / inf inputs: anywhere
  const int b = (int)(48.0 * log2(1.0 + 16.0 * key));       // ~2 % resolution over 6 decades
  return b < 0 ? 0 : (b > ORDER_BINS - 1 ? ORDER_BINS - 1 : b);
}

// pass 1 (whole grid): bin of every problem + global histogram
template <typename T>
__global__ void __launch_bounds__(256)
    order_bins_kernel(const __grid_constant__ DevCfg<T> c, int batch, int N, const double* __restrict__ x0,
                      const double* __restrict__ yref, const double* __restrict__ yref_e,
                      int* __restrict__ bins, int* __restrict__ hist) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= batch) return;
  const int b = order_bin(c, x0 + (size_t)p * NX, yref + (size_t)p * N * NZ, yref_e + (size_t)p * NX,
                          (double)N);
  bins[p] = b;
  atomicAdd(&hist[b], 1);
}

// pass 2 (one CTA): exclusive prefix sum over the bins in DESCENDING key order, then the scatter
__global__ void __launch_bounds__(1024)
    order_scatter_kernel(int batch, const int* __restrict__ bins, int* __restrict__ hist,
                         int* __restrict__ order) {
  __shared__ int start[ORDER_BINS];
  const int t = threadIdx.x;
  const int mine = hist[ORDER_BINS - 1 - t];   // thread t owns bin 1023 - t
  hist[ORDER_BINS - 1 - t] = 0;                // ready for the next launch
  start[t] = mine;
  __syncthreads();
  for (int off = 1; off < ORDER_BINS; off <<= 1) {
    const int v = (t >= off) ? start[t - off] : 0;
    __syncthreads();
    start[t] += v;
    __syncthreads();
  }
  const int excl = start[t] - mine;
  __syncthreads();
  start[t] = excl;
  __syncthreads();
  for (int p = t; p < batch; p += 1024) order[atomicAdd(&start[ORDER_BINS - 1 - bins[p]], 1)] = p;
}

// Test hook: discretise n independent (x,u) pairs, one warp each.
template <typename T, int NS>
__global__ void __launch_bounds__(WARPS * 32)
    nmpc_discretize_kernel(const __grid_constant__ DevCfg<T> c, int n, const double* __restrict__ x,
                           const double* __restrict__ u, double* __restrict__ xn,
                           double* __restrict__ Aout, double* __restrict__ Bout) {
  const int lane = threadIdx.x & 31;
  const int idx = blockIdx.x * WARPS + (threadIdx.x >> 5);
  if (idx >= n) return;
  T xs[NX], us[NU], s[NX], su[NU];
#pragma unroll
  for (int i = 0; i < NX; i++) {
    xs[i] = (T)x[(size_t)idx * NX + i];
    s[i] = (lane - NU == i) ? T(1) : T(0);
  }
#pragma unroll
  for (int i = 0; i < NU; i++) {
    us[i] = (T)u[(size_t)idx * NU + i];
    su[i] = (lane == i) ? T(1) : T(0);
  }
  erk_column<T, NS>(c, xs, us, s, su);
#pragma unroll
  for (int i = 0; i < NX; i++) {
    if (lane < NU) Bout[((size_t)idx * NX + i) * NU + lane] = (double)s[i];
    else if (lane < NZ) Aout[((size_t)idx * NX + i) * NX + (lane - NU)] = (double)s[i];
    if (lane == i) xn[(size_t)idx * NX + i] = (double)xs[i];
  }
}

__global__ void fill_last_cmd_kernel(double* last, int n, double hover) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) last[i] = hover;
}

// _prev_solution_valid = False (controller.py:167, :449) for the problems with mask[i] != 0
// (all problems when mask is null).
__global__ void invalidate_kernel(uint8_t* valid, const uint8_t* __restrict__ mask, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && (!mask || mask[i])) valid[i] = 0;
}

// DFMA throughput micro-benchmark: 8 independent FMA chains per thread (the FP64 roofline
// denominator; MEASURED_PEAKS.json carries no FP64 figure).
__global__ void __launch_bounds__(256) fp64_peak_kernel(double* out, int iters, double a, double b) {
  double v0 = threadIdx.x, v1 = v0 + 1, v2 = v0 + 2, v3 = v0 + 3, v4 = v0 + 4, v5 = v0 + 5,
         v6 = v0 + 6, v7 = v0 + 7;
  for (int i = 0; i < iters; i++) {
#pragma unroll
    for (int r = 0; r < 8; r++) {
      v0 = fma(v0, a, b);
      v1 = fma(v1, a, b);
      v2 = fma(v2, a, b);
      v3 = fma(v3, a, b);
      v4 = fma(v4, a, b);
      v5 = fma(v5, a, b);
      v6 = fma(v6, a, b);
      v7 = fma(v7, a, b);
    }
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = v0 + v1 + v2 + v3 + v4 + v5 + v6 + v7;
}

// FFMA twin of fp64_peak_kernel (FP32 roofline denominator of the FP32 variant).
__global__ void __launch_bounds__(256) fp32_peak_kernel(float* out, int iters, float a, float b) {
  float v0 = threadIdx.x, v1 = v0 + 1, v2 = v0 + 2, v3 = v0 + 3, v4 = v0 + 4, v5 = v0 + 5,
        v6 = v0 + 6, v7 = v0 + 7;
  for (int i = 0; i < iters; i++) {
#pragma unroll
    for (int r = 0; r < 8; r++) {
      v0 = fmaf(v0, a, b);
      v1 = fmaf(v1, a, b);
      v2 = fmaf(v2, a, b);
      v3 = fmaf(v3, a, b);
      v4 = fmaf(v4, a, b);
      v5 = fmaf(v5, a, b);
      v6 = fmaf(v6, a, b);
      v7 = fmaf(v7, a, b);
    }
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = v0 + v1 + v2 + v3 + v4 + v5 + v6 + v7;
}

thread_local char g_err[512] = "";

template <typename T>
DevCfg<T> make_devcfg(const nmpc_config_t& c) {
  DevCfg<T> d;
  memset(&d, 0, sizeof(d));
  d.N = c.N;
  d.qp_iter_max = c.qp_iter_max;
  d.integ_steps = c.integ_steps;
  d.max_sqp_iter = c.max_sqp_iter;
  d.nan_status = c.nan_status;
  d.ipm_warm = c.ipm_warm_start != 0;
  d.as_shortcut = c.active_set_shortcut != 0;
  d.dt = (T)c.dt;
  d.h = (T)(c.dt / (double)c.integ_steps);
  const double sc = c.cost_scale_dt ? c.dt : 1.0;
  const double lm_s = c.lm_scaled_by_dt ? c.lm * c.dt : c.lm;
  for (int i = 0; i < NX; i++) {
    d.Wq[i] = (T)(sc * c.W[i]);
    d.Qd[i] = (T)(sc * c.W[i] + lm_s);
    d.We[i] = (T)c.We[i];
    d.QN[i] = (T)(c.We[i] + c.lm);
  }
  for (int i = 0; i < NU; i++) {
    d.Wr[i] = (T)(sc * c.W[NX + i]);
    d.Rd[i] = (T)(sc * c.W[NX + i] + lm_s);
    d.rx[i] = (T)c.rotor_x[i];
    d.ry[i] = (T)c.rotor_y[i];
    d.rz[i] = (T)c.rotor_z[i];
    d.lbu[i] = (T)c.lbu[i];
    d.ubu[i] = (T)c.ubu[i];
  }
  d.inv_mass = (T)(1.0 / c.mass);
  d.gravity = (T)c.gravity;
  for (int i = 0; i < 3; i++) {
    d.J[i] = (T)c.J[i];
    d.invJ[i] = (T)(1.0 / c.J[i]);
  }
  d.kf = (T)c.kf;
  d.w_min = (T)c.w_min;
  d.w_max = (T)c.w_max;
  d.hover = (T)(c.mass * c.gravity / 4.0);
  d.sqp_tol = (T)c.sqp_tol;
  double tol = c.ipm_tol;
  if (sizeof(T) == 4 && tol < 1e-9) tol = 1e-9;  // FP32: slack*multiplier products resolve to ~1e-9
  d.ipm_tol = (T)tol;
  return d;
}

}  // namespace

struct nmpc_handle {
  nmpc_config_t cfg;
  int batch, device, sm_count, grid, nslots, launches;
  int fixed_n;      // 1: the N = 20 instantiation of the solve kernel runs
  int stage_bytes;  // bytes of inputs the solve kernel stages by TMA per problem (0: direct loads)
  int poison;       // debug: NaN-fill workspace (per launch) and shared tile (per problem)
  WsLayout L;
  size_t ws_bytes;
  void* ws;
  double *xtraj, *utraj, *last_cmd;
  uint8_t* valid;
  int *qp_iters, *sqp_iters, *counter;
  int *order, *bins, *hist;   // dispatch order of the batch, work-key bins, their histogram (order_*_kernel)
  // staging for nmpc_solve_host: ONE device allocation each for the inputs [x0 | yref | yref_e] and the
  // outputs [u0 | thrust | speed | status] (sections padded to 16 bytes), so that a small batch moves
  // with one H2D and one D2H copy through the pinned host mirrors pin_in / pin_out
  unsigned char *s_in, *s_out, *pin_in, *pin_out;
  unsigned char *zc_in, *zc_out;  // device aliases of the pinned mirrors (tiny batches: zero-copy), else null
  unsigned char* last_out;        // where the last handle-owned outputs were written (s_out or zc_out)
  size_t in_bytes, out_bytes, off_yref, off_yref_e, off_thrust, off_speed, off_status;
  double *s_x0, *s_yref, *s_yref_e, *s_u0, *s_thrust, *s_speed;
  int* s_status;
  uint8_t* s_mask;       // staging of a host mask (nmpc_invalidate)
  cudaStream_t stream;   // the handle's own stream (host-buffer calls, copies of the stores)
  cudaEvent_t ev_last;   // recorded behind the last kernel this handle enqueued on a caller stream
  // device-side constants, rebuilt only when the configuration changes
  DevCfg<double> dc64;
  DevCfg<float> dc32;
};

// The handle's stores (valid, warm start, statistics, last command) are written by kernels the
// caller enqueues on streams of his choice.  Calls that touch the stores from the host side order
// themselves behind the last such kernel through ev_last and run on the handle's own stream, so
// they never synchronise the device (other handles / streams of the process keep running).
static void mark_enqueued(nmpc_handle* h, cudaStream_t st) { cudaEventRecord(h->ev_last, st); }
static cudaError_t order_behind_last(nmpc_handle* h) {
  return cudaStreamWaitEvent(h->stream, h->ev_last, 0);
}
static void refresh_devcfg(nmpc_handle* h);

#define CK(call)                                                                         \
  do {                                                                                   \
    cudaError_t e_ = (call);                                                             \
    if (e_ != cudaSuccess) {                                                             \
      snprintf(g_err, sizeof(g_err), "%s:%d %s: %s", __FILE__, __LINE__, #call,          \
               cudaGetErrorString(e_));                                                  \
      return (e_ == cudaErrorMemoryAllocation) ? NMPC_E_NOMEM : NMPC_E_CUDA;             \
    }                                                                                    \
  } while (0)

static int validate(const nmpc_config_t* c) {
  if (!c) return NMPC_E_INVALID;
  if (c->N < 1 || c->N > 600) return NMPC_E_INVALID;
  if (!(c->dt > 0.0)) return NMPC_E_INVALID;
  if (c->integ_ns != 1 && c->integ_ns != 2 && c->integ_ns != 4) return NMPC_E_INVALID;
  if (c->integ_steps < 1 || c->qp_iter_max < 1) return NMPC_E_INVALID;
  if (c->dispatch < NMPC_DISPATCH_AUTO || c->dispatch > NMPC_DISPATCH_LPT) return NMPC_E_INVALID;
  if (c->precision != NMPC_PRECISION_F64 && c->precision != NMPC_PRECISION_F32)
    return NMPC_E_INVALID;
  if (!(c->mass > 0.0) || !(c->J[0] > 0.0) || !(c->J[1] > 0.0) || !(c->J[2] > 0.0))
    return NMPC_E_INVALID;
  for (int i = 0; i < 4; i++)
    if (!(c->lbu[i] < c->ubu[i])) return NMPC_E_INVALID;
  return NMPC_OK;
}

static void refresh_devcfg(nmpc_handle* h) {
  h->dc64 = make_devcfg<double>(h->cfg);
  h->dc32 = make_devcfg<float>(h->cfg);
}
template <typename T>
static const DevCfg<T>& devcfg_of(const nmpc_handle* h);
template <>
const DevCfg<double>& devcfg_of<double>(const nmpc_handle* h) { return h->dc64; }
template <>
const DevCfg<float>& devcfg_of<float>(const nmpc_handle* h) { return h->dc32; }

// The reference's default horizon (config/params.yaml: horizon_steps 20) with its default
// integrator runs the fixed-horizon instantiation.
constexpr int N_FIXED = 20;
static bool use_fixed_horizon(const nmpc_handle* h) {
  return h->cfg.N == N_FIXED && h->cfg.integ_ns == 2 && h->cfg.integ_steps == 2 &&
         !getenv("NMPC_B200_GENERIC_N");
}

template <typename T, int NS, int NF>
static int configure_kernel(nmpc_handle* h, size_t smem) {
  auto kern = nmpc_solve_kernel<T, NS, NF>;
  CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int occ = 0;
  CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, WARPS * 32, smem));
  if (occ < 1) occ = 1;
  if (const char* cap = getenv("NMPC_B200_MAX_CTAS_PER_SM")) {  // tuning knob (L2 footprint)
    const int v = atoi(cap);
    if (v >= 1 && v < occ) occ = v;
  }
  int grid = h->sm_count * occ;
  if (grid > h->batch) grid = h->batch;   // small batches: one problem per CTA before a second warp is used
  h->grid = grid;
  h->nslots = grid * WARPS;
  return NMPC_OK;
}

template <typename T>
static size_t solve_smem_bytes(const nmpc_handle* h) {
  return ((size_t)WARPS * smem_elems_per_warp() + VBLOCK_ELEMS) * sizeof(T);
}

template <typename T>
static int configure_T(nmpc_handle* h) {
  h->fixed_n = use_fixed_horizon(h) ? 1 : 0;
  h->stage_bytes = make_stage_layout(h->cfg.N, (int)sizeof(T)).total;
  if (getenv("NMPC_B200_NO_TMA")) h->stage_bytes = 0;  // A/B knob: direct loads of the inputs
  const size_t smem = solve_smem_bytes<T>(h);
  switch (h->cfg.integ_ns) {
    case 1: return configure_kernel<T, 1, 0>(h, smem);
    case 2:
      return h->fixed_n ? configure_kernel<T, 2, N_FIXED>(h, smem)
                        : configure_kernel<T, 2, 0>(h, smem);
    default: return configure_kernel<T, 4, 0>(h, smem);
  }
}

extern "C" {

static int create_impl(nmpc_handle* h, const nmpc_config_t* cfg, int batch, int device) {
  memset(h, 0, sizeof(*h));
  h->cfg = *cfg;
  h->batch = batch;
  h->device = device;
  refresh_devcfg(h);
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, device));
  h->sm_count = prop.multiProcessorCount;
  h->L = make_layout(cfg->N);
  int rc = (cfg->precision == NMPC_PRECISION_F64) ? configure_T<double>(h) : configure_T<float>(h);
  if (rc != NMPC_OK) return rc;
  const size_t esz = (cfg->precision == NMPC_PRECISION_F64) ? 8 : 4;
  h->ws_bytes = (size_t)h->nslots * h->L.total * esz;
  const size_t N = cfg->N, B = batch;
  CK(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
  CK(cudaEventCreateWithFlags(&h->ev_last, cudaEventDisableTiming));
  CK(cudaMalloc(&h->ws, h->ws_bytes));
  // +16 B: the 16-byte aligned range the TMA copies for the last problem may end 8 B past its row
  CK(cudaMalloc(&h->xtraj, B * (N + 1) * NX * sizeof(double) + 16));
  CK(cudaMalloc(&h->utraj, B * N * NU * sizeof(double)));
  CK(cudaMalloc(&h->last_cmd, B * NU * sizeof(double)));
  CK(cudaMalloc(&h->valid, B));
  CK(cudaMalloc(&h->qp_iters, B * sizeof(int)));
  CK(cudaMalloc(&h->sqp_iters, B * sizeof(int)));
  CK(cudaMalloc(&h->counter, sizeof(int)));
  CK(cudaMalloc(&h->order, B * sizeof(int)));
  CK(cudaMalloc(&h->bins, B * sizeof(int)));
  CK(cudaMalloc(&h->hist, ORDER_BINS * sizeof(int)));
  auto pad16 = [](size_t v) { return (v + 15) & ~(size_t)15; };
  h->off_yref = pad16(B * NX * sizeof(double));
  h->off_yref_e = h->off_yref + pad16(B * N * NZ * sizeof(double));
  h->in_bytes = h->off_yref_e + pad16(B * NX * sizeof(double));
  h->off_thrust = pad16(B * NU * sizeof(double));
  h->off_speed = 2 * h->off_thrust;
  h->off_status = 3 * h->off_thrust;
  h->out_bytes = h->off_status + pad16(B * sizeof(int));
  CK(cudaMalloc(&h->s_in, h->in_bytes));
  CK(cudaMalloc(&h->s_out, h->out_bytes));
  h->s_x0 = reinterpret_cast<double*>(h->s_in);
  h->s_yref = reinterpret_cast<double*>(h->s_in + h->off_yref);
  h->s_yref_e = reinterpret_cast<double*>(h->s_in + h->off_yref_e);
  h->s_u0 = reinterpret_cast<double*>(h->s_out);
  h->s_thrust = reinterpret_cast<double*>(h->s_out + h->off_thrust);
  h->s_speed = reinterpret_cast<double*>(h->s_out + h->off_speed);
  h->s_status = reinterpret_cast<int*>(h->s_out + h->off_status);
  if (h->in_bytes <= SMALL_HOST_BYTES) {  // latency path of nmpc_solve_host (e.g. the single vehicle)
    CK(cudaHostAlloc(&h->pin_in, h->in_bytes, cudaHostAllocMapped));
    CK(cudaHostAlloc(&h->pin_out, h->out_bytes, cudaHostAllocMapped));
    if (h->in_bytes <= ZERO_COPY_BYTES && !getenv("NMPC_B200_NO_ZERO_COPY")) {
      CK(cudaHostGetDevicePointer(reinterpret_cast<void**>(&h->zc_in), h->pin_in, 0));
      CK(cudaHostGetDevicePointer(reinterpret_cast<void**>(&h->zc_out), h->pin_out, 0));
    }
  }
  CK(cudaMalloc(&h->s_mask, B));
  cudaStream_t st = h->stream;
  CK(cudaMemsetAsync(h->ws, 0, h->ws_bytes, st));
  CK(cudaMemsetAsync(h->xtraj, 0, B * (N + 1) * NX * sizeof(double) + 16, st));
  CK(cudaMemsetAsync(h->utraj, 0, B * N * NU * sizeof(double), st));
  CK(cudaMemsetAsync(h->valid, 0, B, st));
  CK(cudaMemsetAsync(h->qp_iters, 0, B * sizeof(int), st));
  CK(cudaMemsetAsync(h->sqp_iters, 0, B * sizeof(int), st));
  CK(cudaMemsetAsync(h->hist, 0, ORDER_BINS * sizeof(int), st));
  CK(cudaEventRecord(h->ev_last, st));
  return nmpc_reset_last_command(h);
}

int nmpc_create(const nmpc_config_t* cfg, int32_t batch, int32_t device, nmpc_handle_t** out) {
  if (!out) return NMPC_E_INVALID;
  *out = nullptr;
  int rc = validate(cfg);
  if (rc != NMPC_OK || batch < 1) return NMPC_E_INVALID;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1) {
    snprintf(g_err, sizeof(g_err), "no CUDA device (libnmpc_b200 has no CPU fallback)");
    return NMPC_E_NODEVICE;
  }
  if (device < 0 || device >= ndev) return NMPC_E_INVALID;
  CK(cudaSetDevice(device));
  nmpc_handle* h = new (std::nothrow) nmpc_handle();
  if (!h) return NMPC_E_NOMEM;
  rc = create_impl(h, cfg, batch, device);
  if (rc != NMPC_OK) {  // nothing of a half-built handle survives (nmpc_destroy tolerates nulls)
    nmpc_destroy(h);
    return rc;
  }
  *out = h;
  return NMPC_OK;
}

int nmpc_destroy(nmpc_handle_t* h) {
  if (!h) return NMPC_OK;
  cudaSetDevice(h->device);
  if (h->stream) {
    if (h->ev_last) cudaStreamWaitEvent(h->stream, h->ev_last, 0);
    cudaStreamSynchronize(h->stream);  // the last kernel that uses the buffers has finished
  }
  cudaFree(h->ws);
  cudaFree(h->xtraj);
  cudaFree(h->utraj);
  cudaFree(h->last_cmd);
  cudaFree(h->valid);
  cudaFree(h->qp_iters);
  cudaFree(h->sqp_iters);
  cudaFree(h->counter);
  cudaFree(h->order);
  cudaFree(h->bins);
  cudaFree(h->hist);
  cudaFree(h->s_in);
  cudaFree(h->s_out);
  if (h->pin_in) cudaFreeHost(h->pin_in);
  if (h->pin_out) cudaFreeHost(h->pin_out);
  cudaFree(h->s_mask);
  if (h->ev_last) cudaEventDestroy(h->ev_last);
  if (h->stream) cudaStreamDestroy(h->stream);
  delete h;
  return NMPC_OK;
}

int nmpc_reset_last_command(nmpc_handle_t* h) {
  if (!h) return NMPC_E_INVALID;
  CK(cudaSetDevice(h->device));
  const int n = h->batch * NU;
  const double hover = h->cfg.mass * h->cfg.gravity / 4.0;  // mpc_controller_node:64
  CK(order_behind_last(h));
  fill_last_cmd_kernel<<<(n + 255) / 256, 256, 0, h->stream>>>(h->last_cmd, n, hover);
  h->launches++;
  CK(cudaGetLastError());
  CK(cudaStreamSynchronize(h->stream));
  return NMPC_OK;
}

int nmpc_invalidate_device(nmpc_handle_t* h, const uint8_t* mask, void* stream) {
  if (!h) return NMPC_E_INVALID;
  CK(cudaSetDevice(h->device));
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  invalidate_kernel<<<(h->batch + 255) / 256, 256, 0, st>>>(h->valid, mask, h->batch);
  h->launches++;
  CK(cudaGetLastError());
  mark_enqueued(h, st);
  return NMPC_OK;
}

int nmpc_reconfigure(nmpc_handle_t* h, const nmpc_config_t* cfg) {
  if (!h) return NMPC_E_INVALID;
  int rc = validate(cfg);
  if (rc != NMPC_OK) return rc;
  if (cfg->N != h->cfg.N || cfg->precision != h->cfg.precision) return NMPC_E_INVALID;
  CK(cudaSetDevice(h->device));
  const int old_ns = h->cfg.integ_ns;
  h->cfg = *cfg;
  refresh_devcfg(h);
  if (cfg->integ_ns != old_ns) {
    const int slots = h->nslots;
    rc = (cfg->precision == NMPC_PRECISION_F64) ? configure_T<double>(h) : configure_T<float>(h);
    if (rc != NMPC_OK) return rc;
    if (h->nslots > slots) {  // never exceed the allocated workspace
      h->nslots = slots;
      h->grid = slots / WARPS;
    }
  }
  return nmpc_invalidate(h, nullptr);  // controller.py:167
}

int nmpc_set_bounds(nmpc_handle_t* h, const double lbu[4], const double ubu[4]) {
  if (!h || !lbu || !ubu) return NMPC_E_INVALID;
  for (int i = 0; i < 4; i++)
    if (!(lbu[i] < ubu[i])) return NMPC_E_INVALID;
  for (int i = 0; i < 4; i++) {
    h->cfg.lbu[i] = lbu[i];
    h->cfg.ubu[i] = ubu[i];
  }
  refresh_devcfg(h);
  return NMPC_OK;
}

int nmpc_invalidate(nmpc_handle_t* h, const uint8_t* mask) {
  if (!h) return NMPC_E_INVALID;
  CK(cudaSetDevice(h->device));
  // behind the last solve of this handle, on the handle's own stream; the host waits for that
  // stream only, so a solve enqueued afterwards on ANY stream sees the cleared flags
  CK(order_behind_last(h));
  const uint8_t* dmask = nullptr;
  if (mask) {
    CK(cudaMemcpyAsync(h->s_mask, mask, h->batch, cudaMemcpyHostToDevice, h->stream));
    dmask = h->s_mask;
  }
  invalidate_kernel<<<(h->batch + 255) / 256, 256, 0, h->stream>>>(h->valid, dmask, h->batch);
  h->launches++;
  CK(cudaGetLastError());
  CK(cudaStreamSynchronize(h->stream));
  return NMPC_OK;
}

extern "C++" {
template <typename T, int NS, int NF>
static int launch_solve(nmpc_handle* h, const double* x0, const double* yref, const double* yref_e,
                        double* u0, int32_t* status, double* thrust, double* speed,
                        cudaStream_t st) {
  SolveArgs<T> A;
  A.cfg = devcfg_of<T>(h);
  A.batch = h->batch;
  A.x0 = x0;
  A.yref = yref;
  A.yref_e = yref_e;
  A.xtraj = h->xtraj;
  A.utraj = h->utraj;
  A.valid = h->valid;
  A.last_cmd = h->last_cmd;
  A.u0 = u0;
  A.status = status;
  A.thrust = thrust;
  A.speed = speed;
  A.qp_iters = h->qp_iters;
  A.sqp_iters = h->sqp_iters;
  A.ws = reinterpret_cast<T*>(h->ws);
  // bulk copies need 16-byte aligned sources: the caller's yref may not be (then direct loads)
  A.stage = (h->stage_bytes > 0 && (reinterpret_cast<uintptr_t>(yref) & 15) == 0) ? 1 : 0;
  A.poison = h->poison;
  const size_t smem = solve_smem_bytes<T>(h);
  // one workspace per handle: a solve on another stream than the previous one queues behind it
  CK(cudaStreamWaitEvent(st, h->ev_last, 0));
  if (h->batch > h->nslots) CK(cudaMemsetAsync(h->counter, 0, sizeof(int), st));  // else: one round, no queue
  // longest-expected-first dispatch when the launch is a few waves of very unequal problems
  const int disp = h->cfg.dispatch;
  const bool lpt = disp == NMPC_DISPATCH_LPT ||
                   (disp == NMPC_DISPATCH_AUTO && h->cfg.max_sqp_iter > 1 && h->batch > h->nslots);
  A.order = nullptr;
  if (lpt) {
    order_bins_kernel<T><<<(h->batch + 255) / 256, 256, 0, st>>>(A.cfg, h->batch, h->cfg.N, x0, yref,
                                                                 yref_e, h->bins, h->hist);
    order_scatter_kernel<<<1, 1024, 0, st>>>(h->batch, h->bins, h->hist, h->order);
    h->launches += 2;
    CK(cudaGetLastError());
    A.order = h->order;
  }
  if (h->poison) CK(cudaMemsetAsync(h->ws, 0xff, h->ws_bytes, st));  // all-ones = NaN in FP64 / FP32
  nmpc_solve_kernel<T, NS, NF>
      <<<h->grid, WARPS * 32, smem, st>>>(A, h->L, h->counter);
  h->launches++;
  CK(cudaGetLastError());
  mark_enqueued(h, st);
  return NMPC_OK;
}

template <typename T>
static int launch_solve_T(nmpc_handle* h, const double* x0, const double* yref,
                          const double* yref_e, double* u0, int32_t* status, double* thrust,
                          double* speed, cudaStream_t st) {
  switch (h->cfg.integ_ns) {
    case 1: return launch_solve<T, 1, 0>(h, x0, yref, yref_e, u0, status, thrust, speed, st);
    case 2:
      return h->fixed_n
                 ? launch_solve<T, 2, N_FIXED>(h, x0, yref, yref_e, u0, status, thrust, speed, st)
                 : launch_solve<T, 2, 0>(h, x0, yref, yref_e, u0, status, thrust, speed, st);
    default: return launch_solve<T, 4, 0>(h, x0, yref, yref_e, u0, status, thrust, speed, st);
  }
}
}  // extern "C++"

int nmpc_solve_device(nmpc_handle_t* h, const double* x0, const double* yref, const double* yref_e,
                      double* u0, int32_t* status, double* thrust, double* speed, void* stream) {
  if (!h || !x0 || !yref || !yref_e) return NMPC_E_INVALID;
  CK(cudaSetDevice(h->device));
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (h->cfg.precision == NMPC_PRECISION_F64)
    return launch_solve_T<double>(h, x0, yref, yref_e, u0, status, thrust, speed, st);
  return launch_solve_T<float>(h, x0, yref, yref_e, u0, status, thrust, speed, st);
}

// ---------------------------------------------------------------- allocation-free solve
// (no counterpart in the reference, see include/nmpc_b200.h).  g_static_scratch in HOST code is a
// stand-in symbol, not a device address: the address a kernel may use comes from
// cudaGetSymbolAddress, once per device with that device current (static_setup).
extern "C++" {
constexpr size_t STATIC_HEAD = 256;  // [work-queue counter | pad] in front of the workspaces
static __device__ __align__(256) unsigned char g_static_scratch[NMPC_STATIC_SCRATCH_BYTES];

constexpr int STATIC_MAX_DEVICES = 64;
struct StaticDevice {
  unsigned char* scratch;  // device address of g_static_scratch on this device
  int sm_count;
  int occ[8];              // resident CTAs per SM of the instantiation [precision * 4 + variant], 0 = not set up
};
static StaticDevice g_static_dev[STATIC_MAX_DEVICES];
static std::mutex g_static_mutex;

template <typename T, int NS, int NF>
static int launch_static(const nmpc_config_t* cfg, int batch, int device, int key, SolveArgs<T>& A,
                         cudaStream_t st) {
  StaticDevice& D = g_static_dev[device];
  const size_t smem = ((size_t)WARPS * smem_elems_per_warp() + VBLOCK_ELEMS) * sizeof(T);
  auto kern = nmpc_solve_kernel<T, NS, NF>;
  {
    std::lock_guard<std::mutex> lock(g_static_mutex);
    if (!D.scratch) {
      void* p = nullptr;
      CK(cudaGetSymbolAddress(&p, g_static_scratch));
      CK(cudaDeviceGetAttribute(&D.sm_count, cudaDevAttrMultiProcessorCount, device));
      D.scratch = static_cast<unsigned char*>(p);
    }
    if (!D.occ[key]) {
      CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      int occ = 0;
      CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, WARPS * 32, smem));
      D.occ[key] = occ < 1 ? 1 : occ;
    }
  }
  const WsLayout L = make_layout(cfg->N);
  const size_t slot_bytes = (size_t)L.total * sizeof(T);
  const size_t fit = (NMPC_STATIC_SCRATCH_BYTES - STATIC_HEAD) / (slot_bytes * WARPS);  // CTAs
  if (fit < 1) {
    snprintf(g_err, sizeof(g_err),
             "nmpc_solve_static: horizon N = %d needs %zu bytes of workspace per CTA, the static "
             "scratch array holds %zu (NMPC_STATIC_SCRATCH_BYTES)",
             cfg->N, slot_bytes * WARPS, (size_t)NMPC_STATIC_SCRATCH_BYTES - STATIC_HEAD);
    return NMPC_E_INVALID;
  }
  int grid = D.sm_count * D.occ[key];
  if ((size_t)grid > fit) grid = (int)fit;
  if (grid > batch) grid = batch;
  int* counter = reinterpret_cast<int*>(D.scratch);
  A.ws = reinterpret_cast<T*>(D.scratch + STATIC_HEAD);
  CK(cudaMemsetAsync(counter, 0, sizeof(int), st));  // the one part of the array that must start at 0
  kern<<<grid, WARPS * 32, smem, st>>>(A, L, counter);
  CK(cudaGetLastError());
  return NMPC_OK;
}

template <typename T>
static int solve_static_T(const nmpc_config_t* cfg, int batch, int device, const double* x0,
                          const double* yref, const double* yref_e, double* xtraj, double* utraj,
                          uint8_t* valid, double* last_cmd, double* u0, int32_t* status,
                          double* thrust, double* speed, int32_t* qp_iters, int32_t* sqp_iters,
                          cudaStream_t st) {
  SolveArgs<T> A;
  A.cfg = make_devcfg<T>(*cfg);
  A.batch = batch;
  A.x0 = x0;
  A.yref = yref;
  A.yref_e = yref_e;
  A.xtraj = xtraj;
  A.utraj = utraj;
  A.valid = valid;
  A.last_cmd = last_cmd;
  A.u0 = u0;
  A.status = status;
  A.thrust = thrust;
  A.speed = speed;
  A.qp_iters = qp_iters;
  A.sqp_iters = sqp_iters;
  A.ws = nullptr;
  A.order = nullptr;  // index order: the longest-expected-first order would need per-batch scratch
  A.stage = 0;        // plain loads: no alignment or padding demands on the caller's buffers
  A.poison = 0;
  const int p4 = (sizeof(T) == 4 ? 4 : 0);
  const bool fixed = cfg->N == N_FIXED && cfg->integ_ns == 2 && cfg->integ_steps == 2 &&
                     !getenv("NMPC_B200_GENERIC_N");
  switch (cfg->integ_ns) {
    case 1: return launch_static<T, 1, 0>(cfg, batch, device, p4 + 0, A, st);
    case 2:
      return fixed ? launch_static<T, 2, N_FIXED>(cfg, batch, device, p4 + 1, A, st)
                   : launch_static<T, 2, 0>(cfg, batch, device, p4 + 2, A, st);
    default: return launch_static<T, 4, 0>(cfg, batch, device, p4 + 3, A, st);
  }
}
}  // extern "C++"

int nmpc_solve_static(const nmpc_config_t* cfg, int32_t batch, int32_t device, const double* x0,
                      const double* yref, const double* yref_e, double* xtraj, double* utraj,
                      uint8_t* valid, double* last_cmd, double* u0, int32_t* status, double* thrust,
                      double* speed, int32_t* qp_iters, int32_t* sqp_iters, void* stream) {
  if (validate(cfg) != NMPC_OK) {
    snprintf(g_err, sizeof(g_err), "nmpc_solve_static: invalid configuration");
    return NMPC_E_INVALID;
  }
  if (batch < 1 || batch > NMPC_STATIC_MAX_BATCH) {
    snprintf(g_err, sizeof(g_err), "nmpc_solve_static: batch %d outside 1..%d (NMPC_STATIC_MAX_BATCH)",
             (int)batch, NMPC_STATIC_MAX_BATCH);
    return NMPC_E_INVALID;
  }
  if (!x0 || !yref || !yref_e || !xtraj || !utraj || !valid || !last_cmd) {
    snprintf(g_err, sizeof(g_err), "nmpc_solve_static: null input or warm-start store pointer");
    return NMPC_E_INVALID;
  }
  if (device < 0 || device >= STATIC_MAX_DEVICES) {
    snprintf(g_err, sizeof(g_err), "nmpc_solve_static: device %d outside 0..%d", (int)device,
             STATIC_MAX_DEVICES - 1);
    return NMPC_E_INVALID;
  }
  CK(cudaSetDevice(device));
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (cfg->precision == NMPC_PRECISION_F64)
    return solve_static_T<double>(cfg, batch, device, x0, yref, yref_e, xtraj, utraj, valid,
                                  last_cmd, u0, status, thrust, speed, qp_iters, sqp_iters, st);
  return solve_static_T<float>(cfg, batch, device, x0, yref, yref_e, xtraj, utraj, valid, last_cmd,
                               u0, status, thrust, speed, qp_iters, sqp_iters, st);
}

int nmpc_solve_host(nmpc_handle_t* h, const double* x0, const double* yref, const double* yref_e,
                    double* u0, int32_t* status, double* thrust, double* speed) {
  if (!h || !x0 || !yref || !yref_e) return NMPC_E_INVALID;
  CK(cudaSetDevice(h->device));
  const size_t B = h->batch, N = h->cfg.N;
  cudaStream_t st = h->stream;
  if (h->pin_in) {
    // small batch: pack into the pinned mirror, ONE H2D copy, solve, ONE D2H copy (7 copies of
    // pageable memory cost more than the solve of a single vehicle)
    memcpy(h->pin_in, x0, B * NX * sizeof(double));
    memcpy(h->pin_in + h->off_yref, yref, B * N * NZ * sizeof(double));
    memcpy(h->pin_in + h->off_yref_e, yref_e, B * NX * sizeof(double));
    // a tiny batch (the single vehicle): the kernel reads the inputs from the pinned mirror and
    // writes its outputs there (one PCIe round trip for the staged rows, posted writes) -- no copy
    // nodes in front of and behind the kernel
    const bool zc = h->zc_in != nullptr;
    unsigned char* in = zc ? h->zc_in : h->s_in;
    unsigned char* out = zc ? h->zc_out : h->s_out;
    h->last_out = out;
    if (!zc) CK(cudaMemcpyAsync(h->s_in, h->pin_in, h->in_bytes, cudaMemcpyHostToDevice, st));
    int rc = nmpc_solve_device(h, reinterpret_cast<double*>(in),
                               reinterpret_cast<double*>(in + h->off_yref),
                               reinterpret_cast<double*>(in + h->off_yref_e),
                               reinterpret_cast<double*>(out),
                               reinterpret_cast<int32_t*>(out + h->off_status),
                               reinterpret_cast<double*>(out + h->off_thrust),
                               reinterpret_cast<double*>(out + h->off_speed), st);
    if (rc != NMPC_OK) return rc;
    if (!zc) CK(cudaMemcpyAsync(h->pin_out, h->s_out, h->out_bytes, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    if (u0) memcpy(u0, h->pin_out, B * NU * sizeof(double));
    if (thrust) memcpy(thrust, h->pin_out + h->off_thrust, B * NU * sizeof(double));
    if (speed) memcpy(speed, h->pin_out + h->off_speed, B * NU * sizeof(double));
    if (status) memcpy(status, h->pin_out + h->off_status, B * sizeof(int));
    return NMPC_OK;
  }
  h->last_out = h->s_out;
  CK(cudaMemcpyAsync(h->s_x0, x0, B * NX * sizeof(double), cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(h->s_yref, yref, B * N * NZ * sizeof(double), cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(h->s_yref_e, yref_e, B * NX * sizeof(double), cudaMemcpyHostToDevice, st));
  int rc = nmpc_solve_device(h, h->s_x0, h->s_yref, h->s_yref_e, h->s_u0, h->s_status, h->s_thrust,
                             h->s_speed, st);
  if (rc != NMPC_OK) return rc;
  if (u0) CK(cudaMemcpyAsync(u0, h->s_u0, B * NU * sizeof(double), cudaMemcpyDeviceToHost, st));
  if (status) CK(cudaMemcpyAsync(status, h->s_status, B * sizeof(int), cudaMemcpyDeviceToHost, st));
  if (thrust)
    CK(cudaMemcpyAsync(thrust, h->s_thrust, B * NU * sizeof(double), cudaMemcpyDeviceToHost, st));
  if (speed)
    CK(cudaMemcpyAsync(speed, h->s_speed, B * NU * sizeof(double), cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  return NMPC_OK;
}

int nmpc_get_trajectories(nmpc_handle_t* h, double* xtraj, double* utraj, uint8_t* valid,
                          int32_t dst_is_device) {
  if (!h) return NMPC_E_INVALID;
  CK(cudaSetDevice(h->device));
  CK(order_behind_last(h));
  const size_t B = h->batch, N = h->cfg.N;
  const cudaMemcpyKind kind = dst_is_device ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost;
  cudaStream_t st = h->stream;
  if (xtraj) CK(cudaMemcpyAsync(xtraj, h->xtraj, B * (N + 1) * NX * sizeof(double), kind, st));
  if (utraj) CK(cudaMemcpyAsync(utraj, h->utraj, B * N * NU * sizeof(double), kind, st));
  if (valid) CK(cudaMemcpyAsync(valid, h->valid, B, kind, st));
  CK(cudaStreamSynchronize(st));
  return NMPC_OK;
}

int nmpc_set_trajectories(nmpc_handle_t* h, const double* xtraj, const double* utraj,
                          const uint8_t* valid) {
  if (!h) return NMPC_E_INVALID;
  CK(cudaSetDevice(h->device));
  CK(order_behind_last(h));
  const size_t B = h->batch, N = h->cfg.N;
  cudaStream_t st = h->stream;
  if (xtraj)
    CK(cudaMemcpyAsync(h->xtraj, xtraj, B * (N + 1) * NX * sizeof(double), cudaMemcpyHostToDevice,
                       st));
  if (utraj)
    CK(cudaMemcpyAsync(h->utraj, utraj, B * N * NU * sizeof(double), cudaMemcpyHostToDevice, st));
  if (valid) CK(cudaMemcpyAsync(h->valid, valid, B, cudaMemcpyHostToDevice, st));
  CK(cudaStreamSynchronize(st));
  return NMPC_OK;
}

int nmpc_get_stats(nmpc_handle_t* h, int32_t* qp_iters, int32_t* sqp_iters,
                   int32_t dst_is_device) {
  if (!h) return NMPC_E_INVALID;
  CK(cudaSetDevice(h->device));
  CK(order_behind_last(h));
  const size_t B = h->batch;
  const cudaMemcpyKind kind = dst_is_device ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost;
  cudaStream_t st = h->stream;
  if (qp_iters) CK(cudaMemcpyAsync(qp_iters, h->qp_iters, B * sizeof(int), kind, st));
  if (sqp_iters) CK(cudaMemcpyAsync(sqp_iters, h->sqp_iters, B * sizeof(int), kind, st));
  CK(cudaStreamSynchronize(st));
  return NMPC_OK;
}

extern "C++" {
template <typename T>
static int launch_disc(nmpc_handle* h, int n, const double* x, const double* u, double* xn,
                       double* A, double* B, cudaStream_t st) {
  const DevCfg<T>& c = devcfg_of<T>(h);
  const int grid = (n + WARPS - 1) / WARPS;
  switch (h->cfg.integ_ns) {
    case 1: nmpc_discretize_kernel<T, 1><<<grid, WARPS * 32, 0, st>>>(c, n, x, u, xn, A, B); break;
    case 2: nmpc_discretize_kernel<T, 2><<<grid, WARPS * 32, 0, st>>>(c, n, x, u, xn, A, B); break;
    default: nmpc_discretize_kernel<T, 4><<<grid, WARPS * 32, 0, st>>>(c, n, x, u, xn, A, B); break;
  }
  h->launches++;
  CK(cudaGetLastError());
  return NMPC_OK;
}
}  // extern "C++"

int nmpc_discretize_device(nmpc_handle_t* h, int32_t n, const double* x, const double* u,
                           double* xn, double* A, double* B, void* stream) {
  if (!h || n < 1 || !x || !u || !xn || !A || !B) return NMPC_E_INVALID;
  CK(cudaSetDevice(h->device));
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (h->cfg.precision == NMPC_PRECISION_F64) return launch_disc<double>(h, n, x, u, xn, A, B, st);
  return launch_disc<float>(h, n, x, u, xn, A, B, st);
}


// ------------------------------------------------------------------ closed loop (f1-f3)
static PlantCfg make_plantcfg(const nmpc_handle* h, const nmpc_plant_t& p) {
  PlantCfg pc;
  pc.substeps = p.substeps;
  pc.motor_model = p.motor_model;
  pc.dt = p.dt;
  pc.tau_up = p.tau_up;
  pc.tau_down = p.tau_down;
  pc.rotor_drag = p.rotor_drag;
  pc.inv_mass = 1.0 / (p.mass > 0.0 ? p.mass : h->cfg.mass);
  pc.rolling_moment = p.rolling_moment;
  pc.rotors_yaw_sign = p.rotors_yaw_sign ? 1 : 0;
  return pc;
}

static int validate_plant(const nmpc_plant_t* p, const double* rotor_speed) {
  if (!p || p->substeps < 0 || p->substeps > 100000) return NMPC_E_INVALID;
  if (p->substeps > 0 && !(p->dt > 0.0)) return NMPC_E_INVALID;
  if (p->motor_model != 0 && p->motor_model != 1) return NMPC_E_INVALID;
  if (!(p->rolling_moment >= 0.0) || !(p->rotor_drag >= 0.0)) return NMPC_E_INVALID;
  if (p->motor_model == 1 && (!rotor_speed || !(p->tau_up > 0.0) || !(p->tau_down > 0.0)))
    return NMPC_E_INVALID;
  return NMPC_OK;
}

int nmpc_reference_setpoint_device(nmpc_handle_t* h, const double* pos, const double* yaw,
                                   double* yref, double* yref_e, void* stream) {
  if (!h || !pos || !yref || !yref_e) return NMPC_E_INVALID;
  CK(cudaSetDevice(h->device));
  const int N = h->cfg.N;
  const size_t total = (size_t)h->batch * ((size_t)N * NZ + NX);
  size_t grid = (total + 255) / 256;
  if (grid > (size_t)h->sm_count * 16) grid = (size_t)h->sm_count * 16;
  const double hover = h->cfg.mass * h->cfg.gravity / 4.0;
  reference_setpoint_kernel<<<(unsigned)grid, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      h->batch, N, pos, yaw, hover, yref, yref_e);
  h->launches++;
  CK(cudaGetLastError());
  return NMPC_OK;
}

int nmpc_reference_figure8_device(nmpc_handle_t* h, const double* t0, const double* amplitude,
                                  const double* period, const double centre[3], double t_offset,
                                  double* yref, double* yref_e, void* stream) {
  if (!h || !t0 || !amplitude || !period || !centre || !yref || !yref_e) return NMPC_E_INVALID;
  CK(cudaSetDevice(h->device));
  const int N = h->cfg.N;
  const size_t total = (size_t)h->batch * (N + 1);
  size_t grid = (total + 127) / 128;
  if (grid > (size_t)h->sm_count * 32) grid = (size_t)h->sm_count * 32;
  const double hover = h->cfg.mass * h->cfg.gravity / 4.0;
  reference_figure8_kernel<<<(unsigned)grid, 128, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      h->batch, N, h->cfg.dt, t_offset, t0, amplitude, period, centre[0], centre[1], centre[2],
      hover, yref, yref_e);
  h->launches++;
  CK(cudaGetLastError());
  return NMPC_OK;
}

int nmpc_odometry_to_state_device(nmpc_handle_t* h, const double* odom, double* x0,
                                  int32_t* zero_quat, void* stream) {
  if (!h || !odom || !x0) return NMPC_E_INVALID;
  CK(cudaSetDevice(h->device));
  odometry_to_state_kernel<<<(h->batch + 127) / 128, 128, 0,
                             reinterpret_cast<cudaStream_t>(stream)>>>(h->batch, odom, x0,
                                                                       zero_quat);
  h->launches++;
  CK(cudaGetLastError());
  return NMPC_OK;
}

static int launch_plant(nmpc_handle* h, const nmpc_plant_t* plant, double* x, double* x0,
                        const double* thrust, double* rotor_speed, const double* pref,
                        size_t pref_stride, const int32_t* status, const int32_t* qp_iters,
                        double* stats, cudaStream_t st) {
  const DevCfg<double>& c = h->dc64;
  const PlantCfg pc = make_plantcfg(h, *plant);
  plant_step_kernel<<<(h->batch + 127) / 128, 128, 0, st>>>(c, pc, h->batch, x, x0, thrust,
                                                            rotor_speed, pref, pref_stride, status,
                                                            qp_iters, stats);
  h->launches++;
  CK(cudaGetLastError());
  mark_enqueued(h, st);
  return NMPC_OK;
}

int nmpc_plant_step_device(nmpc_handle_t* h, const nmpc_plant_t* plant, double* x, double* x0,
                           const double* thrust, double* rotor_speed, void* stream) {
  if (!h || !x) return NMPC_E_INVALID;
  int rc = validate_plant(plant, rotor_speed);
  if (rc != NMPC_OK) return rc;
  CK(cudaSetDevice(h->device));
  return launch_plant(h, plant, x, x0, thrust, rotor_speed, nullptr, 0, nullptr, nullptr, nullptr,
                      reinterpret_cast<cudaStream_t>(stream));
}

int nmpc_rollout_device(nmpc_handle_t* h, const nmpc_rollout_t* run, void* stream) {
  if (!h || !run || !run->x || run->steps < 0) return NMPC_E_INVALID;
  int rc = validate_plant(&run->plant, run->rotor_speed);
  if (rc != NMPC_OK || run->plant.substeps < 1) return NMPC_E_INVALID;
  if (run->ref_mode == NMPC_REF_SETPOINT) {
    if (!run->sp_pos) return NMPC_E_INVALID;
  } else if (run->ref_mode == NMPC_REF_FIGURE8) {
    if (!run->f8_t0 || !run->f8_amplitude || !run->f8_period) return NMPC_E_INVALID;
  } else {
    return NMPC_E_INVALID;
  }
  CK(cudaSetDevice(h->device));
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  h->last_out = h->s_out;
  const int N = h->cfg.N;
  // state handed to the first solve: normalised copy of the plant state (substeps = 0)
  nmpc_plant_t norm_only = run->plant;
  norm_only.substeps = 0;
  rc = launch_plant(h, &norm_only, run->x, h->s_x0, nullptr, run->rotor_speed, nullptr, 0, nullptr,
                    nullptr, nullptr, st);
  if (rc != NMPC_OK) return rc;
  if (run->ref_mode == NMPC_REF_SETPOINT) {  // constant horizon: built once
    rc = nmpc_reference_setpoint_device(h, run->sp_pos, run->sp_yaw, h->s_yref, h->s_yref_e, stream);
    if (rc != NMPC_OK) return rc;
  }
  for (int k = 0; k < run->steps; k++) {
    if (run->ref_mode == NMPC_REF_FIGURE8) {
      rc = nmpc_reference_figure8_device(h, run->f8_t0, run->f8_amplitude, run->f8_period,
                                         run->f8_centre, run->t_start + k * run->plant.dt,
                                         h->s_yref, h->s_yref_e, stream);
      if (rc != NMPC_OK) return rc;
    }
    rc = nmpc_solve_device(h, h->s_x0, h->s_yref, h->s_yref_e, h->s_u0, h->s_status, h->s_thrust,
                           h->s_speed, stream);
    if (rc != NMPC_OK) return rc;
    // tracking error is measured against the reference position at the time of the new state:
    // the set-point itself, or stage 1 of the horizon just used (= p_ref(t + dt))
    const double* pref = (run->ref_mode == NMPC_REF_SETPOINT) ? run->sp_pos : h->s_yref + NZ;
    const size_t stride = (run->ref_mode == NMPC_REF_SETPOINT) ? 3 : (size_t)N * NZ;
    rc = launch_plant(h, &run->plant, run->x, h->s_x0, h->s_thrust, run->rotor_speed, pref, stride,
                      h->s_status, h->qp_iters, run->stats, st);
    if (rc != NMPC_OK) return rc;
  }
  return NMPC_OK;
}

int nmpc_last_outputs(nmpc_handle_t* h, const double** u0, const int32_t** status,
                      const double** thrust, const double** speed) {
  if (!h) return NMPC_E_INVALID;
  const unsigned char* o = h->last_out ? h->last_out : h->s_out;
  if (u0) *u0 = reinterpret_cast<const double*>(o);
  if (status) *status = reinterpret_cast<const int32_t*>(o + h->off_status);
  if (thrust) *thrust = reinterpret_cast<const double*>(o + h->off_thrust);
  if (speed) *speed = reinterpret_cast<const double*>(o + h->off_speed);
  return NMPC_OK;
}

int nmpc_fp64_peak(int32_t device, double* tflops) {
  if (!tflops) return NMPC_E_INVALID;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1) return NMPC_E_NODEVICE;
  CK(cudaSetDevice(device));
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, device));
  const int blocks = prop.multiProcessorCount * 8, threads = 256, iters = 4096;
  double* buf = nullptr;
  CK(cudaMalloc(&buf, (size_t)blocks * threads * sizeof(double)));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  double best = 0.0;
  for (int rep = 0; rep < 5; rep++) {
    CK(cudaEventRecord(e0));
    fp64_peak_kernel<<<blocks, threads>>>(buf, iters, 1.0000001, 1e-9);
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms = 0.f;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    const double flops = 2.0 * 64.0 * iters * (double)blocks * threads;
    if (rep > 0 && flops / (ms * 1e-3) * 1e-12 > best) best = flops / (ms * 1e-3) * 1e-12;
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(buf);
  *tflops = best;
  return NMPC_OK;
}

int nmpc_fp32_peak(int32_t device, double* tflops) {
  if (!tflops) return NMPC_E_INVALID;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1) return NMPC_E_NODEVICE;
  CK(cudaSetDevice(device));
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, device));
  const int blocks = prop.multiProcessorCount * 8, threads = 256, iters = 8192;
  float* buf = nullptr;
  CK(cudaMalloc(&buf, (size_t)blocks * threads * sizeof(float)));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  double best = 0.0;
  for (int rep = 0; rep < 5; rep++) {
    CK(cudaEventRecord(e0));
    fp32_peak_kernel<<<blocks, threads>>>(buf, iters, 1.0000001f, 1e-9f);
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms = 0.f;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    const double flops = 2.0 * 64.0 * iters * (double)blocks * threads;
    if (rep > 0 && flops / (ms * 1e-3) * 1e-12 > best) best = flops / (ms * 1e-3) * 1e-12;
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(buf);
  *tflops = best;
  return NMPC_OK;
}

int nmpc_debug_poison(nmpc_handle_t* h, int32_t on) {
  if (!h) return NMPC_E_INVALID;
  h->poison = on ? 1 : 0;
  return NMPC_OK;
}

#ifdef NMPC_PROFILE
// Instrumented build only (scripts/phase_profile.py): read and clear the per-phase cycle counters.
int nmpc_debug_phase_cycles(double* out, int32_t n) {
  unsigned long long v[PH_COUNT];
  CK(cudaDeviceSynchronize());
  CK(cudaMemcpyFromSymbol(v, g_phase_cycles, sizeof(v)));
  for (int i = 0; i < n && i < PH_COUNT; i++) out[i] = (double)v[i];
  memset(v, 0, sizeof(v));
  CK(cudaMemcpyToSymbol(g_phase_cycles, v, sizeof(v)));
  return NMPC_OK;
}
#endif

int nmpc_batch(const nmpc_handle_t* h) { return h ? h->batch : NMPC_E_INVALID; }
int nmpc_horizon(const nmpc_handle_t* h) { return h ? h->cfg.N : NMPC_E_INVALID; }
int64_t nmpc_workspace_bytes(const nmpc_handle_t* h) { return h ? (int64_t)h->ws_bytes : -1; }
int nmpc_kernel_launches(const nmpc_handle_t* h) { return h ? h->launches : NMPC_E_INVALID; }

const char* nmpc_strerror(int code) {
  switch (code) {
    case NMPC_OK: return "ok";
    case NMPC_E_INVALID: return "invalid argument or unsupported configuration";
    case NMPC_E_CUDA: return "CUDA runtime error";
    case NMPC_E_NOMEM: return "out of memory";
    case NMPC_E_NODEVICE: return "no CUDA device (no CPU fallback exists)";
    default: return "unknown error";
  }
}
const char* nmpc_last_error(void) { return g_err; }
const char* nmpc_version(void) { return "nmpc_b200 0.1 (sm_100a)"; }

}  // extern "C"
