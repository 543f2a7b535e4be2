// nmpc_solver.cuh -- one-warp-per-problem SQP solver (active-set QP method, Riccati interior point
// as the fallback), sm_100a.
//
// Replaces, for the OCP of /root/reference/src/rotors_mpc_controller/controller.py:175-264,
// what `solver.solve()` (controller.py:447) runs inside acados: ERK + forward sensitivities,
// LINEAR_LS Gauss-Newton assembly + Levenberg-Marquardt, the HPIPM QP solve and the full
// step.  Nothing here is derived from acados / HPIPM sources (they are not in the reference
// tree); the QP methods are a primal-dual active-set method on the input box (as_solve) and the
// textbook stage-wise Riccati recursion inside a Mehrotra predictor-corrector (ipm_solve), laid
// out for one warp:
//
//   K1 linearise: phase A, lane k integrates the nominal state of stage k once; phase B, the
//      11 N stored sensitivity columns of M_k = [B_k | A_k] (inputs, q, w; the p and v columns are
//      constants) are dealt out flat over the lanes and evaluate the directional derivative of
//      the model at the stored points (linearize_points; linearize is the column-per-lane form
//      for horizons whose points do not fit the tile).  No Jacobian matrix is formed.
//   K2 factorisation: both half-warps own the 14 computed columns and split the ROWS of
//      W = P M, G = M'W and of the cost-to-go update (see nmpc_riccati.cuh); P, the cp.async-staged
//      M_k, K_k and the 4x4 input block go through a per-warp shared-memory tile.  Active-set
//      re-solves restart the recursion from cost-to-go checkpoints (ck_stage).
//   Roll-outs: lanes 0..12 own the rows of M_k, lanes 16..19 the rows of K_k; the roll-out of an
//      active-set iteration also delivers the multipliers of the fixed inputs.
//   Per-stage records [M_k | b_k | K_k | factor slots], feed-forwards, checkpoints and all
//   horizon-long vectors live in a global workspace slot owned by the warp (sized by N, so
//   N = 600 runs unchanged); per-problem inputs are staged by 1-D TMA bulk copies.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "nmpc_model.cuh"

namespace nmpc {

constexpr unsigned FULL = 0xffffffffu;

enum : int { ST_SUCCESS = 0, ST_NAN = 1, ST_MAXITER = 2, ST_MINSTEP = 3, ST_QPFAIL = 4 };

// Phase timers of the instrumented build (-DNMPC_PROFILE, scripts/phase_profile.py): cycles per
// phase summed over all warps.  Compiled out of the product library.
enum : int { PH_LOAD = 0, PH_LIN, PH_GRAD, PH_FACTOR, PH_FWD, PH_DELTA, PH_ELEM, PH_STEP, PH_EPI,
             PH_TOTAL, PH_COUNT };
#ifdef NMPC_PROFILE
__device__ unsigned long long g_phase_cycles[PH_COUNT];
#define PROF_T0() long long prof_t_ = clock64()
#define PROF_ADD(ph_)                                                                   \
  do {                                                                                  \
    const long long now_ = clock64();                                                   \
    if (lane == 0) atomicAdd(&g_phase_cycles[ph_], (unsigned long long)(now_ - prof_t_)); \
    prof_t_ = now_;                                                                     \
  } while (0)
#else
#define PROF_T0() do { } while (0)
#define PROF_ADD(ph_) do { } while (0)
#endif

// Offsets (in elements of T) inside one workspace slot.
struct WsLayout {
  int N;
  int M, K, Qi, kff, kffd;
  int xb, q, b, dx;                                                // x-type vectors
  int ub, r, lb, ubd, du, ua, tl, tu, ll, lu, rbl0, rbu0, ru0;    // u-type vectors
  int dtl, dtu, dll, dlu, Rt, rt, dr, itl, itu;
  int ck;  // NCK cost-to-go checkpoints (P, p) of the factor sweep, CK_ELEMS each
  int total;
};

// One record per stage keeps what the sweeps stream together contiguous (one 16-byte-chunked
// cp.async range per stage): [ M_k | K_k | LDL' factors of Quu_k ].
// Nothing in the model depends on the position or the velocity (controller.py:286-346), so the
// p and v columns of every A_k are the constants [I; 0; 0; 0] and [dt I; I; 0; 0]: they are never
// stored, every sweep synthesises them in registers.  A row of M_k is therefore
//     [ B_k(i, 0..3) | A_k(i, q0..q3) | A_k(i, w0..w2) | b_k(i) ]        12 elements, 16-byte aligned
// (the defect b_k rides in the last column, so the right-hand side travels with the tile).
constexpr int M_LD = 12;
constexpr int MC_Q = 4, MC_W = 8, MC_B = 11;  // first q / w column and the defect column of a row
constexpr int REC_K = NX * M_LD;              // 156: offset of K_k inside a record
constexpr int K_LD = 14;                      // rows of K_k (4 x 13) padded to 16-byte multiples
constexpr int REC_F = REC_K + NU * K_LD;      // 212: offset of the factors (10 used of 16)
constexpr int REC = REC_F + 16;               // 228 elements per stage
// column of a row of M_k that holds dx+/d(state j), j = 6..12 (q, w)
__host__ __device__ constexpr int mcol_of_state(int j) { return j - 6 + MC_Q; }

// Checkpoints of the backward recursion: the cost-to-go (P_c, p_c) at NCK stages c = ck_stage(N, j).
// P_c depends on the stages >= c only, so when the active set of a QP changes at stages <= kmax the
// next factorisation restarts from the lowest checkpoint above kmax instead of the terminal stage
// (as_solve): the stages above it would be recomputed to the same bits.
constexpr int NCK = 4;
constexpr int CK_ELEMS = (NX + 1) * 14;  // 13 rows of P (14 wide) + p (14)
__host__ __device__ constexpr int ck_stage(int N, int j) { return ((j + 1) * N) / (NCK + 1); }

__host__ __device__ inline WsLayout make_layout(int N) {
  WsLayout L;
  L.N = N;
  int o = 0;
  auto take = [&](int n) {
    int r = o;
    o += (n + 3) & ~3;  // keep 32-byte alignment for T = double
    return r;
  };
  const int nx = (N + 1) * NX, nu = N * NU;
  L.M = take(N * REC);  // stage records [M_k (13 x 12) | K_k (4 x 14) | factors (16)], REC each
  L.K = L.M + REC_K;
  L.Qi = L.M + REC_F;
  L.kff = take(nu);
  L.kffd = take(nu);
  L.xb = take(nx);
  L.q = take(nx);
  L.b = take(nx);
  L.dx = take(nx);
  L.ub = take(nu);
  L.r = take(nu);
  L.lb = take(nu);
  L.ubd = take(nu);
  L.du = take(nu);
  L.ua = take(nu);
  L.tl = take(nu);
  L.tu = take(nu);
  L.ll = take(nu);
  L.lu = take(nu);
  L.rbl0 = take(nu);
  L.rbu0 = take(nu);
  L.ru0 = take(nu);
  L.dtl = take(nu);
  L.dtu = take(nu);
  L.dll = take(nu);
  L.dlu = take(nu);
  L.Rt = take(nu);
  L.rt = take(nu);
  L.dr = take(nu);
  L.itl = take(nu);
  L.itu = take(nu);
  L.ck = take(NCK * CK_ELEMS);
  L.total = o;
  return L;
}

// Horizon known at compile time (NF > 0: the kernel instantiation for the reference's default
// N = 20): every workspace offset and loop bound folds into the instructions -- what acados gets
// from generating code for a fixed horizon.  NF = 0: the layout and N are run-time values.
// (NF = 0 must keep using the caller's layout object -- the kernel parameter in constant memory --
// and not a copy: a copy is address-taken by the sweep calls and would live on the stack.)
#define NMPC_FIXED(c_)                                                   \
  const WsLayout Lfix_ = NF > 0 ? make_layout(NF > 0 ? NF : 1) : WsLayout(); \
  const WsLayout& L = NF > 0 ? Lfix_ : Lr;                               \
  const int N = NF > 0 ? NF : (c_).N;                                    \
  (void)N;

// Per-warp shared-memory tile (elements of T).
constexpr int SP_LD = 14;  // padded leading dimension of P rows (16-byte aligned rows)
constexpr int RING = 4;    // stage records in flight per sweep: cp.async runs 3 stages ahead, which
                           // covers an L2 round trip (~800 cycles) with the ~250 cycles of a roll-out stage
struct SmemLayout {
  static constexpr int P = 0;                    // 13 x 14 cost-to-go + a zero row (dummy W row)
  static constexpr int M = P + (NX + 1) * SP_LD; // stage buffers of the sweeps (cp.async targets):
                                                 //   factor: 2 x 156 + 224 (W), others: RING x 228
  static constexpr int K = M + RING * REC;       // 4 x 14
  static constexpr int p = K + NU * SP_LD;       // 13 (+1)
  static constexpr int wv = p + 14;              // 13 (+1)
  static constexpr int xA = wv + 14;             // 13 (+1) forward-sweep state, ping
  static constexpr int xB = xA + 14;             // pong
  static constexpr int bar = xB + 14;            // mbarrier of the input staging + its parity
  static constexpr int G = bar + 4;              // 16 x 4 exchange tile of factor_sweep + 4 (g_u)
  static constexpr int S = G + 68;               // RING x 8 per-stage scalars of the streaming sweeps
  static constexpr int total = S + RING * 8;     // 1324 elements
};

// ------------------------------------------------------------------ TMA input staging
// Per-problem inputs that are contiguous in HBM -- the reference rows yref[prob] (N x 17), the
// warm-start inputs utraj[prob] (N x 4) and the warm-start states xtraj[prob] ((N+1) x 13) -- are
// brought into shared memory by three bulk-tensor-engine copies (cp.async.bulk, 1-D TMA) that
// complete on one mbarrier: one exposed HBM latency per problem instead of two.  They land in the
// warp's sweep tile, which is idle at the start of a problem, so no shared memory is added (more
// shared memory per CTA shrinks the L1 carve-out: measured -5 % on every workload).  The warm
// start and the gradients of the first SQP iteration are read from there; later SQP iterations
// read yref from L2 as before (the tile belongs to the sweeps by then).
// Bulk copies need 16-byte aligned addresses and sizes: yref rows are N x 136 B (even N only);
// xtraj rows are 8 (mod 16) bytes long when (N+1) is odd, so the enclosing 16-byte aligned range
// is copied (the store is padded by 16 B).  Tile bytes: [0,Y) yref | [Y,Y+U) utraj | [..,+X) xtraj
struct StageLayout {
  int ybytes, ubytes, xbytes, total;  // total = 0: staging off (direct loads)
};
__host__ __device__ inline StageLayout make_stage_layout(int N, int elem_bytes) {
  StageLayout S;
  S.ybytes = N * NZ * 8;
  S.ubytes = N * NU * 8;
  S.xbytes = ((N + 1) * NX * 8 + 15) & ~15;
  S.total = S.ybytes + S.ubytes + S.xbytes;
  if ((N & 1) || S.total > SmemLayout::bar * elem_bytes) S.total = 0;
  return S;
}
constexpr int smem_tile_elems() { return (SmemLayout::total + 3) & ~3; }  // per-warp tile of T
// CTAs of CTA_WARPS independent warps; behind their tiles one block of constants shared by the CTA:
// VBLOCK = the v columns of every A_k, dt e_{p_a} + e_{v_a}, as columns 0..2 of a 13 x M_LD block
#ifndef NMPC_WARPS
#define NMPC_WARPS 4
#endif
constexpr int CTA_WARPS = NMPC_WARPS;
constexpr int VBLOCK_ELEMS = NX * M_LD;
__device__ __forceinline__ unsigned smem_u32(const void* p) {
  return (unsigned)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(void* bar) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(bar)) : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(void* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void tma_load_1d(void* dst, const void* src, unsigned bytes, void* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
          "r"(smem_u32(dst)),
      "l"(src), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}
__device__ __forceinline__ void mbar_wait(void* bar, unsigned parity) {
  unsigned done;
  do {
    asm volatile(
        "{\n.reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n}"
        : "=r"(done)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
  } while (!done);
}

template <typename T>
__device__ __forceinline__ T warp_sum(T v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
  return v;
}
// Min / max over the warp of NON-NEGATIVE, non-NaN values: IEEE bit patterns of such values order
// like unsigned integers, so two redux.sync (high word, then low word among the winners)
// replace five shuffle + fmin rounds (a dozen instructions instead of ~70 for FP64).
__device__ __forceinline__ double warp_min(double v) {
  const unsigned hi = (unsigned)__double2hiint(v), lo = (unsigned)__double2loint(v);
  const unsigned mh = __reduce_min_sync(FULL, hi);
  const unsigned ml = __reduce_min_sync(FULL, hi == mh ? lo : 0xffffffffu);
  return __hiloint2double((int)mh, (int)ml);
}
__device__ __forceinline__ double warp_max(double v) {
  const unsigned hi = (unsigned)__double2hiint(v), lo = (unsigned)__double2loint(v);
  const unsigned mh = __reduce_max_sync(FULL, hi);
  const unsigned ml = __reduce_max_sync(FULL, hi == mh ? lo : 0u);
  return __hiloint2double((int)mh, (int)ml);
}
__device__ __forceinline__ float warp_min(float v) {
  return __uint_as_float(__reduce_min_sync(FULL, __float_as_uint(v)));
}
__device__ __forceinline__ float warp_max(float v) {
  return __uint_as_float(__reduce_max_sync(FULL, __float_as_uint(v)));
}
// Ratio test without a division per candidate: keep the smallest step bound n / (-d) over the
// candidates with d < 0 as the fraction num / den (den > 0), comparing by cross-multiplication.
template <typename T>
__device__ __forceinline__ void ratio_min(T& num, T& den, T n, T d) {
  const T dd = -d;
  if (d < T(0) && n * den < num * dd) {
    num = n;
    den = dd;
  }
}
template <typename T>
__device__ __forceinline__ bool finite_(T v) {
  return (v - v) == T(0);
}

// ------------------------------------------------------------------ K1: linearisation
// For every stage: ERK + forward sensitivities -> the stored columns of M_k = [B_k | A_k] (rows of
// 12 in the stage record) and the defect b_k.  Stages are independent (multiple shooting) and
// nothing in the model depends on the position or the velocity, so the p and v columns of A_k are
// constants that are never formed: 11 columns (u, q, w) per stage are integrated, one per lane,
// as directional derivatives of the ERK map (no Jacobian matrix is built).
// Linearisation with the nominal trajectory integrated ONCE per stage (explicit midpoint, the
// reference's integrator).  The column-per-lane mapping of linearize() below integrates the nominal
// state redundantly in every one of the 11 lanes of a stage -- 43 % of the model arithmetic.  Here
//   phase A: lane k integrates the nominal state of stage k and leaves the points at which the
//            model is evaluated (start and midpoint of every sub-step) in the warp's idle sweep tile,
//            together with the defect b_k; it also writes the LS gradients and the box of its stage;
//   phase B: the 11 N (stage, column) items, dealt out flat over the lanes as before, integrate
//            their sensitivity column only: the directional derivative of the model at the stored points
//            (model_jvp with its nominal outputs unused: the compiler drops them).
// Same expressions in the same order as erk_column, so M_k and b_k are the same to the bit.
// Tile: [x iterate (N+1) x 13 | u iterate N x 4 | points N x (2 steps - 1) x 14 (padded rows)].
constexpr int PT_LD = 14;
template <typename T, int NF>
__device__ __forceinline__ bool linearize_fits_points(const DevCfg<T>& c, int N) {
  const int steps = NF > 0 ? 2 : c.integ_steps;
  const int base = ((N + 1) * NX + N * NU + 1) & ~1;
  return N <= 32 && base + (2 * steps - 1) * N * PT_LD <= SmemLayout::bar;
}
template <typename T, int NF>
__device__ __forceinline__ void linearize_points(const DevCfg<T>& c, T* __restrict__ ws,
                                                 const WsLayout& Lr, T* __restrict__ sm, int lane,
                                                 const double* __restrict__ yref,
                                                 const double* __restrict__ yref_e, bool grads) {
  NMPC_FIXED(c)
  const int steps = NF > 0 ? 2 : c.integ_steps;
  const int npts = 2 * steps - 1;
  const int nxv = (N + 1) * NX, nuv = N * NU;
  T* sx = sm;
  T* su_ = sm + nxv;
  T* pts = sm + ((nxv + nuv + 1) & ~1);
  {
    const T* gx = ws + L.xb;
    const T* gu = ws + L.ub;
    _Pragma("unroll 9") for (int i = lane; i < nxv; i += 32) sx[i] = gx[i];
    _Pragma("unroll 3") for (int i = lane; i < nuv; i += 32) su_[i] = gu[i];
  }
  __syncwarp();
  T* M = ws + L.M;
  T* bv = ws + L.b;
  const T h = c.h;
  // phase A: nominal trajectory of stage `lane`
  if (lane < N) {
    const int k = lane;
    T x[NX], u[NU], f[NX];
#pragma unroll
    for (int i = 0; i < NX; i++) x[i] = sx[k * NX + i];
#pragma unroll
    for (int i = 0; i < NU; i++) u[i] = su_[k * NU + i];
    if (grads) {  // LS gradients and the delta-form box of the stage (what gradients() computes)
      const double* yk = yref + k * NZ;
#pragma unroll
      for (int i = 0; i < NX; i++) ws[L.q + k * NX + i] = c.Wq[i] * (x[i] - (T)yk[i]);
#pragma unroll
      for (int i = 0; i < NU; i++) {
        ws[L.r + k * NU + i] = c.Wr[i] * (u[i] - (T)yk[NX + i]);
        ws[L.lb + k * NU + i] = c.lbu[i] - u[i];
        ws[L.ubd + k * NU + i] = c.ubu[i] - u[i];
        ws[L.Rt + k * NU + i] = c.Rd[i];
      }
    }
    T* pk = pts + k * npts * PT_LD;
    for (int step = 0; step < steps; step++) {
      T xs[NX];
      model_f(c, x, u, f);
#pragma unroll
      for (int i = 0; i < NX; i++) {
        xs[i] = x[i] + (T(0.5) * h) * f[i];
        pk[(2 * step) * PT_LD + i] = xs[i];
      }
      model_f(c, xs, u, f);
#pragma unroll
      for (int i = 0; i < NX; i++) x[i] += h * f[i];
      if (step + 1 < steps) {
#pragma unroll
        for (int i = 0; i < NX; i++) pk[(2 * step + 1) * PT_LD + i] = x[i];
      }
    }
#pragma unroll
    for (int i = 0; i < NX; i++) {  // defect; a copy rides in the last column of M_k's rows
      const T b = x[i] - sx[(k + 1) * NX + i];
      bv[k * NX + i] = b;
      M[k * REC + i * M_LD + MC_B] = b;
    }
  }
  __syncwarp();
  // phase B: sensitivity columns
  constexpr int NCOL = NZ - 6;
  const int total = N * NCOL;
  for (int base = 0; base < total; base += 32) {
    const bool live = base + lane < total;
    const int item = live ? base + lane : total - 1;
    const int k = item / NCOL, j = item - NCOL * k;
    T x[NX], u[NU], s[NX], su[NU], f[NX], dk[NX];
#pragma unroll
    for (int i = 0; i < NX; i++) {
      x[i] = sx[k * NX + i];
      s[i] = (j >= NU && i == j + 2) ? T(1) : T(0);   // column j >= 4 is state j + 2 (q, w = 6..12)
    }
#pragma unroll
    for (int i = 0; i < NU; i++) {
      u[i] = su_[k * NU + i];
      su[i] = (j == i) ? T(1) : T(0);
    }
    const T* pk = pts + k * npts * PT_LD;
    for (int step = 0; step < steps; step++) {
      T ss[NX];
      model_jvp(c, x, u, s, su, f, dk);               // at the start of the sub-step
#pragma unroll
      for (int i = 0; i < NX; i++) ss[i] = s[i] + (T(0.5) * h) * dk[i];
      {
        const T* pm = pk + (2 * step) * PT_LD;        // midpoint
#pragma unroll
        for (int cc = 0; cc < 6; cc++) {
          const Pair<T> v = ld2(pm + 2 * cc);
          x[2 * cc] = v.a;
          x[2 * cc + 1] = v.b;
        }
        x[12] = pm[12];
      }
      model_jvp(c, x, u, ss, su, f, dk);
#pragma unroll
      for (int i = 0; i < NX; i++) s[i] += h * dk[i];
      if (step + 1 < steps) {
        const T* pn = pk + (2 * step + 1) * PT_LD;    // start of the next sub-step
#pragma unroll
        for (int cc = 0; cc < 6; cc++) {
          const Pair<T> v = ld2(pn + 2 * cc);
          x[2 * cc] = v.a;
          x[2 * cc + 1] = v.b;
        }
        x[12] = pn[12];
      }
    }
    if (live) {
#pragma unroll
      for (int i = 0; i < NX; i++) M[k * REC + i * M_LD + j] = s[i];
    }
  }
  if (grads && lane < NX) ws[L.q + N * NX + lane] = c.We[lane] * (sx[N * NX + lane] - (T)yref_e[lane]);
}

template <typename T, int NS, int NF>
__device__ __forceinline__ void linearize(const DevCfg<T>& c, T* __restrict__ ws,
                                          const WsLayout& Lr, T* __restrict__ sm, int lane,
                                          const double* __restrict__ yref,
                                          const double* __restrict__ yref_e, bool grads) {
  NMPC_FIXED(c)
  if (NS == 2 && linearize_fits_points<T, NF>(c, N)) {  // nominal trajectory once per stage
    linearize_points<T, NF>(c, ws, Lr, sm, lane, yref, yref_e, grads);
    return;
  }
  // The iterate is read stage by stage below (17 broadcast loads per pass, each pass would wait
  // for L2): when it fits the (idle) stage buffers it is copied to shared memory once, with all
  // loads in flight together.
  const int nxv = (N + 1) * NX, nuv = N * NU;
  const bool staged = nxv + nuv <= SmemLayout::K - SmemLayout::M;
  const T* xb = ws + L.xb;
  const T* ub = ws + L.ub;
  if (staged) {
    T* sx = sm + SmemLayout::M;
    // (fixed horizon: fully unrolled, all twelve loads of a lane in flight together)
    _Pragma("unroll 9") for (int i = lane; i < nxv; i += 32) sx[i] = xb[i];
    _Pragma("unroll 3") for (int i = lane; i < nuv; i += 32) sx[nxv + i] = ub[i];
    xb = sx;
    ub = sx + nxv;
    __syncwarp();
  }
  T* M = ws + L.M;
  T* bv = ws + L.b;
  // The 11 stored columns of every stage (u0..u3, q0..q3, w0..w2 = columns 0..10 of a record row) are
  // dealt out flat over the lanes: item = 11 * stage + column, 32 items per pass, so N = 20 takes
  // 7 passes with every lane busy (a half-warp per stage with its 14 "computed" columns took 10).
  constexpr int NCOL = NZ - 6;
  const int total = N * NCOL;
  for (int base = 0; base < total; base += 32) {
    const bool live = base + lane < total;
    const int item = live ? base + lane : total - 1;
    const int k = item / NCOL, j = item - NCOL * k;
    T x[NX], u[NU], s[NX], su[NU];
#pragma unroll
    for (int i = 0; i < NX; i++) {
      x[i] = xb[k * NX + i];
      s[i] = (j >= NU && i == j + 2) ? T(1) : T(0);   // column j >= 4 is state j + 2 (q, w = 6..12)
    }
#pragma unroll
    for (int i = 0; i < NU; i++) {
      u[i] = ub[k * NU + i];
      su[i] = (j == i) ? T(1) : T(0);
    }
    const int e2 = j + NCOL;                           // second state entry of lanes j = 0, 1 (11, 12)
    if (grads && live) {
      // LINEAR_LS gradients and the delta-form box of the stage (what gradients() computes), in
      // the same pass: the stage's reference row is read while the integrator runs
      ws[L.q + k * NX + j] = c.Wq[j] * (xb[k * NX + j] - (T)yref[k * NZ + j]);
      if (e2 < NX) ws[L.q + k * NX + e2] = c.Wq[e2] * (xb[k * NX + e2] - (T)yref[k * NZ + e2]);
      if (j < NU) {
        const T uu = ub[k * NU + j];
        ws[L.r + k * NU + j] = c.Wr[j] * (uu - (T)yref[k * NZ + NX + j]);
        ws[L.lb + k * NU + j] = c.lbu[j] - uu;
        ws[L.ubd + k * NU + j] = c.ubu[j] - uu;
        ws[L.Rt + k * NU + j] = c.Rd[j];
      }
    }
    erk_column<T, NS>(c, x, u, s, su);
    if (live) {
#pragma unroll
      for (int i = 0; i < NX; i++) M[k * REC + i * M_LD + j] = s[i];
      // defect (every lane of a stage holds the same nominal x): lane j writes component j, lanes
      // 0 and 1 also 11 and 12; a copy rides in the last column of M_k's rows
      T xl = x[0];
#pragma unroll
      for (int i = 1; i < NCOL; i++) xl = (j == i) ? x[i] : xl;
      const T b1 = xl - xb[(k + 1) * NX + j];
      bv[k * NX + j] = b1;
      M[k * REC + j * M_LD + MC_B] = b1;
      if (e2 < NX) {
        const T b2 = ((j == 0) ? x[NCOL] : x[NCOL + 1]) - xb[(k + 1) * NX + e2];
        bv[k * NX + e2] = b2;
        M[k * REC + e2 * M_LD + MC_B] = b2;
      }
    }
  }
  if (grads && lane < NX) ws[L.q + N * NX + lane] = c.We[lane] * (xb[N * NX + lane] - (T)yref_e[lane]);
}

// LINEAR_LS gradients (controller.py:223-245, stage cost scaled by dt) and the delta-form box.
template <typename T, int NF>
__device__ __noinline__ void gradients(const DevCfg<T>& c, T* __restrict__ ws,
                                       const WsLayout& Lr, const double* __restrict__ yref,
                                       const double* __restrict__ yref_e, int lane) {
  NMPC_FIXED(c)
  const T* xb = ws + L.xb;
  const T* ub = ws + L.ub;
  _Pragma("unroll 3") for (int idx = lane; idx < N * NX; idx += 32) {
    const int k = idx / NX, i = idx - k * NX;
    ws[L.q + idx] = c.Wq[i] * (xb[idx] - (T)yref[k * NZ + i]);
  }
  if (lane < NX) ws[L.q + N * NX + lane] = c.We[lane] * (xb[N * NX + lane] - (T)yref_e[lane]);
  _Pragma("unroll 1") for (int idx = lane; idx < N * NU; idx += 32) {
    const int k = idx >> 2, a = idx & 3;
    const T u = ub[idx];
    ws[L.r + idx] = c.Wr[a] * (u - (T)yref[k * NZ + NX + a]);
    ws[L.lb + idx] = c.lbu[a] - u;
    ws[L.ubd + idx] = c.ubu[a] - u;
    ws[L.Rt + idx] = c.Rd[a];
  }
}

}  // namespace nmpc

#include "nmpc_riccati.cuh"

namespace nmpc {
// ------------------------------------------------------------------ K2: interior point
// Mehrotra predictor-corrector for   min 1/2 z'Hz + g'z  s.t. dynamics, lb <= du <= ub.
//  phase 0: equality-constrained solve; if it respects the box it is the optimum (strictly
//           convex QP) -> 1 factorisation, no interior-point iteration.
//  else:    start there (dynamics and x-stationarity hold exactly), slacks t = max(slack, thr0),
//           lam = mu0/t.  One step length for primal and dual, so the remaining linear residuals
//           (box, u-stationarity) are rho = prod(1-alpha) times their initial values rbl0, rbu0,
//           ru0 and every Newton right-hand side is built from those (no cancellation).
//  warm_eps > 0 (QPs after the first of an SQP run, see solve_one): the slacks and multipliers
//           the previous QP left in the workspace are the start instead, clipped to >= warm_eps,
//           and phase 0 is skipped (start at dz = 0, residuals b and q scaled by rho as well).
// Returns the QP status; *iters = number of Riccati factorisations; *ipm_conv = the interior-point
// loop ran and met its tolerance (its iterate is a valid warm start for the next QP).
template <typename T, int NF>
__device__ __forceinline__ int ipm_solve(const DevCfg<T>& c, T* __restrict__ ws,
                                         const WsLayout& Lr, T* __restrict__ sm, int lane,
                                         T warm_eps, bool* ipm_conv, int* iters) {
  *ipm_conv = false;
  NMPC_FIXED(c)
  PROF_T0();
  const int nb = N * NU;
  const int nxv = (N + 1) * NX;
  const T thr0 = T(0.1), mu0 = T(0.1);
  int it = 0;

  // A warm-started QP skips phase 0 (its bounds were active a moment ago): the loop starts at
  // dz = 0, where the dynamics and x-stationarity residuals are the plain data b and q; they
  // shrink with rho like the box and u-stationarity residuals and enter every Riccati solve as
  // (rho b, rho q).
  const bool skip0 = warm_eps > T(0);

  // finite check of the QP data (NaN state / reference -> QP failure, status 4)
  T res0 = T(0);
  {
    bool fin = true;
    _Pragma("unroll 1") for (int i = lane; i < nb; i += 32) {
      fin = fin && finite_(ws[L.r + i]) && finite_(ws[L.lb + i]) && finite_(ws[L.ubd + i]);
      ws[L.rt + i] = ws[L.r + i];
    }
    _Pragma("unroll 3") for (int i = lane; i < nxv; i += 32) {
      const T v = ws[L.q + i];
      fin = fin && finite_(v);
      res0 = fmax(res0, fabs(v));
    }
    _Pragma("unroll 3") for (int i = lane; i < N * NX; i += 32) {
      const T v = ws[L.b + i];
      fin = fin && finite_(v);
      res0 = fmax(res0, fabs(v));
    }
    if (!__all_sync(FULL, fin)) {
      *iters = 0;
      return ST_QPFAIL;
    }
  }
  __syncwarp();
  if (!skip0) {
    res0 = T(0);  // b and q are met exactly by the phase-0 point
    PROF_ADD(PH_ELEM);
    const bool ok0 = factor_sweep<T, NF, false>(c, ws, L, sm, lane, T(1), false, -1);
    PROF_ADD(PH_FACTOR);
    it = 1;
    if (!ok0) {
      *iters = it;
      return ST_QPFAIL;
    }
    forward_sweep<T, NF, false>(c, ws, L, sm, lane, T(1), false, true, false, L.kff, L.dx, L.du);
    PROF_ADD(PH_FWD);
  }
  // box check + interior-point start
  bool feas = !skip0;
  _Pragma("unroll 1") for (int i = lane; i < nb; i += 32) {
    const T du = skip0 ? T(0) : ws[L.du + i];
    const T lb = ws[L.lb + i], ub = ws[L.ubd + i];
    feas = feas && (du >= lb) && (du <= ub);
    T tl = fmax(du - lb, thr0), tu = fmax(ub - du, thr0);
    T ll = mu0 / tl, lu = mu0 / tu;
    T ru = T(0);
    if (skip0) {
      tl = fmax(ws[L.tl + i], warm_eps);
      tu = fmax(ws[L.tu + i], warm_eps);
      ll = fmax(ws[L.ll + i], warm_eps);
      lu = fmax(ws[L.lu + i], warm_eps);
      ru = ws[L.r + i];
      ws[L.du + i] = T(0);
    }
    ru += -ll + lu;
    const T rbl = du - lb - tl, rbu = ub - du - tu;
    ws[L.tl + i] = tl;
    ws[L.tu + i] = tu;
    ws[L.ll + i] = ll;
    ws[L.lu + i] = lu;
    ws[L.rbl0 + i] = rbl;
    ws[L.rbu0 + i] = rbu;
    ws[L.ru0 + i] = ru;
    res0 = fmax(res0, fmax(fmax(fabs(rbl), fabs(rbu)), fabs(ru)));
  }
  if (__all_sync(FULL, feas)) {
    *iters = it;
    PROF_ADD(PH_ELEM);
    return ST_SUCCESS;
  }
  res0 = warp_max(res0);
  T rho = T(1);
  int status = ST_MAXITER;
  const T inv2nb = T(1) / T(2 * nb);

  // Element-wise passes over the nb = 4N box rows.  NSL rows per lane and trip: all loads of a
  // trip are issued before any arithmetic or store.  NSL = 1 keeps the code small (the instruction
  // cache, not the L2 latency of these short loops, is what the passes compete for).
#ifndef NMPC_NSL
#define NMPC_NSL 1
#endif
  constexpr int NSL = NMPC_NSL;
#define NMPC_ROWS(...)                                                \
  _Pragma("unroll 1") for (int base_ = 0; base_ < nb; base_ += 32 * NSL) { \
    int ix[NSL];                                                      \
    bool in[NSL];                                                     \
    _Pragma("unroll") for (int e = 0; e < NSL; e++) {                 \
      const int i_ = base_ + 32 * e + lane;                           \
      in[e] = i_ < nb;                                                \
      ix[e] = in[e] ? i_ : 0;                                         \
    }                                                                 \
    __VA_ARGS__                                                       \
  }
#define NMPC_LD(dst, off) \
  T dst[NSL];             \
  _Pragma("unroll") for (int e = 0; e < NSL; e++) dst[e] = ws[(off) + ix[e]];

  // predictor set-up from the current (tl,tu,ll,lu): complementarity, slack reciprocals (reused
  // by the affine and corrector passes), barrier Hessian Rt and gradient rt
  T mu = T(0);
  NMPC_ROWS({
    NMPC_LD(tl, L.tl) NMPC_LD(tu, L.tu) NMPC_LD(ll, L.ll) NMPC_LD(lu, L.lu)
    NMPC_LD(ru0, L.ru0) NMPC_LD(rbl0, L.rbl0) NMPC_LD(rbu0, L.rbu0)
    _Pragma("unroll") for (int e = 0; e < NSL; e++) {
      const T itl = rcp_(tl[e]), itu = rcp_(tu[e]);
      const T dl = ll[e] * itl, dw = lu[e] * itu;
      if (in[e]) {
        mu += tl[e] * ll[e] + tu[e] * lu[e];
        ws[L.itl + ix[e]] = itl;
        ws[L.itu + ix[e]] = itu;
        ws[L.Rt + ix[e]] = c.Rd[ix[e] & 3] + dl + dw;
        ws[L.rt + ix[e]] = rho * ru0[e] + dl * (rho * rbl0[e]) + ll[e] - dw * (rho * rbu0[e]) - lu[e];
      }
    }
  })

  for (; it < c.qp_iter_max; it++) {
    mu = warp_sum(mu) * inv2nb;
    if (!(mu == mu)) {
      status = ST_QPFAIL;
      break;
    }
    if (mu <= c.ipm_tol && rho * res0 <= c.ipm_tol) {
      status = ST_SUCCESS;
      break;
    }
    __syncwarp();
    const T rq = skip0 ? rho : T(0);
    PROF_ADD(PH_ELEM);
    const bool ok = factor_sweep<T, NF, false>(c, ws, L, sm, lane, rq, false, -1);
    PROF_ADD(PH_FACTOR);
    if (!ok) {
      // FP32 precision floor: once the complementarity is below 1e-6 a lost pivot means the
      // barrier Hessian can no longer be resolved in single precision -> keep the iterate
      status = (sizeof(T) == 4 && mu <= T(1e-6)) ? ST_SUCCESS : ST_QPFAIL;
      break;
    }
    forward_sweep<T, NF, false>(c, ws, L, sm, lane, rq, false, false, false, L.kff, L.dx, L.ua);   // wx = false: no state output
    PROF_ADD(PH_FWD);
    // affine step length (division-free ratio test: the running minimum is kept as a fraction
    // an/ad and candidates are compared by cross-multiplication)
    T an = T(1), ad = T(1);
    NMPC_ROWS({
      NMPC_LD(tl, L.tl) NMPC_LD(tu, L.tu) NMPC_LD(ll, L.ll) NMPC_LD(lu, L.lu)
      NMPC_LD(ua, L.ua) NMPC_LD(rbl0, L.rbl0) NMPC_LD(rbu0, L.rbu0)
      NMPC_LD(itl, L.itl) NMPC_LD(itu, L.itu)
      _Pragma("unroll") for (int e = 0; e < NSL; e++) {
        const T dtl = ua[e] + rho * rbl0[e];
        const T dtu = -ua[e] + rho * rbu0[e];
        const T dll = -(ll[e] * itl[e]) * dtl - ll[e];
        const T dlu = -(lu[e] * itu[e]) * dtu - lu[e];
        if (in[e]) {
          ws[L.dtl + ix[e]] = dtl;
          ws[L.dtu + ix[e]] = dtu;
          ws[L.dll + ix[e]] = dll;
          ws[L.dlu + ix[e]] = dlu;
          ratio_min(an, ad, tl[e], dtl);
          ratio_min(an, ad, tu[e], dtu);
          ratio_min(an, ad, ll[e], dll);
          ratio_min(an, ad, lu[e], dlu);
        }
      }
    })
    const T a_aff = warp_min(an / ad);
    // centering parameter from the complementarity the affine step would reach
    T mu_aff = T(0);
    NMPC_ROWS({
      NMPC_LD(tl, L.tl) NMPC_LD(tu, L.tu) NMPC_LD(ll, L.ll) NMPC_LD(lu, L.lu)
      NMPC_LD(dtl, L.dtl) NMPC_LD(dtu, L.dtu) NMPC_LD(dll, L.dll) NMPC_LD(dlu, L.dlu)
      _Pragma("unroll") for (int e = 0; e < NSL; e++)
        if (in[e])
          mu_aff += (tl[e] + a_aff * dtl[e]) * (ll[e] + a_aff * dll[e]) +
                    (tu[e] + a_aff * dtu[e]) * (lu[e] + a_aff * dlu[e]);
    })
    mu_aff = warp_sum(mu_aff) * inv2nb;
    T sigma = mu_aff / mu;
    sigma = sigma * sigma * sigma;
    const T smu = sigma * mu;
    // weight of the second-order term: min(1, 2 alpha_aff) (breaks Mehrotra 2-cycles)
    const T cw = fmin(T(1), T(2) * a_aff);
    NMPC_ROWS({
      NMPC_LD(dtl, L.dtl) NMPC_LD(dtu, L.dtu) NMPC_LD(dll, L.dll) NMPC_LD(dlu, L.dlu)
      NMPC_LD(itl, L.itl) NMPC_LD(itu, L.itu)
      _Pragma("unroll") for (int e = 0; e < NSL; e++) {
        const T sl = smu - cw * dtl[e] * dll[e];
        const T su = smu - cw * dtu[e] * dlu[e];
        if (in[e]) ws[L.dr + ix[e]] = -sl * itl[e] + su * itu[e];
      }
    })
    __syncwarp();
    // corrector increment (vector-only re-solve), accumulated onto the predictor direction
    PROF_ADD(PH_ELEM);
    delta_backward<T, NF>(c, ws, L, sm, lane);
    PROF_ADD(PH_DELTA);
    forward_sweep<T, NF, false>(c, ws, L, sm, lane, T(0), true, false, false, L.kffd, L.dx, L.ua);
    PROF_ADD(PH_FWD);
    T bn = T(2), bd = T(1);  // ratios >= 2 cannot bind: alpha = min(1, tau * ratio), tau >= 0.995
    bool fin = true;         // NaN anywhere must surface: comparisons drop NaNs, so test explicitly
    NMPC_ROWS({
      NMPC_LD(tl, L.tl) NMPC_LD(tu, L.tu) NMPC_LD(ll, L.ll) NMPC_LD(lu, L.lu)
      NMPC_LD(itl, L.itl) NMPC_LD(itu, L.itu) NMPC_LD(un, L.ua)
      NMPC_LD(adtl, L.dtl) NMPC_LD(adtu, L.dtu) NMPC_LD(adll, L.dll) NMPC_LD(adlu, L.dlu)
      NMPC_LD(rbl0, L.rbl0) NMPC_LD(rbu0, L.rbu0)
      _Pragma("unroll") for (int e = 0; e < NSL; e++) {
        const T sl = smu - cw * adtl[e] * adll[e];
        const T su = smu - cw * adtu[e] * adlu[e];
        const T dtl = un[e] + rho * rbl0[e];
        const T dtu = -un[e] + rho * rbu0[e];
        const T dll = -(ll[e] * itl[e]) * dtl + sl * itl[e] - ll[e];
        const T dlu = -(lu[e] * itu[e]) * dtu + su * itu[e] - lu[e];
        if (in[e]) {
          ws[L.dtl + ix[e]] = dtl;
          ws[L.dtu + ix[e]] = dtu;
          ws[L.dll + ix[e]] = dll;
          ws[L.dlu + ix[e]] = dlu;
          ratio_min(bn, bd, tl[e], dtl);
          ratio_min(bn, bd, tu[e], dtu);
          ratio_min(bn, bd, ll[e], dll);
          ratio_min(bn, bd, lu[e], dlu);
          fin = fin && finite_(dtl) && finite_(dll) && finite_(dlu);
        }
      }
    })
    T alpha = bn / bd;
    if (!__all_sync(FULL, fin)) {
      status = ST_QPFAIL;
      break;
    }
    alpha = warp_min(alpha);
    const T tau = fmax(T(0.995), T(1) - mu);
    alpha = fmin(T(1), tau * alpha);
    if (alpha < T(1e-12)) {
      status = ST_MINSTEP;
      break;
    }
    // step + the predictor set-up of the next iteration (same rows, values still in registers)
    const T rho_n = rho * (T(1) - alpha);
    mu = T(0);
    NMPC_ROWS({
      NMPC_LD(du, L.du) NMPC_LD(ua, L.ua)
      NMPC_LD(tl, L.tl) NMPC_LD(tu, L.tu) NMPC_LD(ll, L.ll) NMPC_LD(lu, L.lu)
      NMPC_LD(dtl, L.dtl) NMPC_LD(dtu, L.dtu) NMPC_LD(dll, L.dll) NMPC_LD(dlu, L.dlu)
      NMPC_LD(ru0, L.ru0) NMPC_LD(rbl0, L.rbl0) NMPC_LD(rbu0, L.rbu0)
      _Pragma("unroll") for (int e = 0; e < NSL; e++) {
        const T ntl = tl[e] + alpha * dtl[e], ntu = tu[e] + alpha * dtu[e];
        const T nll = ll[e] + alpha * dll[e], nlu = lu[e] + alpha * dlu[e];
        const T itl = rcp_(ntl), itu = rcp_(ntu);
        const T dl = nll * itl, dw = nlu * itu;
        if (in[e]) {
          ws[L.du + ix[e]] = du[e] + alpha * ua[e];
          ws[L.tl + ix[e]] = ntl;
          ws[L.tu + ix[e]] = ntu;
          ws[L.ll + ix[e]] = nll;
          ws[L.lu + ix[e]] = nlu;
          mu += ntl * nll + ntu * nlu;
          ws[L.itl + ix[e]] = itl;
          ws[L.itu + ix[e]] = itu;
          ws[L.Rt + ix[e]] = c.Rd[ix[e] & 3] + dl + dw;
          ws[L.rt + ix[e]] =
              rho_n * ru0[e] + dl * (rho_n * rbl0[e]) + nll - dw * (rho_n * rbu0[e]) - nlu;
        }
      }
    })
    rho = rho_n;
    __syncwarp();
  }
#undef NMPC_ROWS
#undef NMPC_LD
  // The state direction is only needed once: roll the linearised dynamics out for the final
  // input direction, dx_{k+1} = A_k dx_k + B_k du_k + b_k (instead of updating dx every iteration).
  __syncwarp();
  PROF_ADD(PH_ELEM);
  forward_sweep<T, NF, false>(c, ws, L, sm, lane, T(1), false, true, true, L.kff, L.dx, L.du);
  PROF_ADD(PH_FWD);
  *iters = it;
  *ipm_conv = (status == ST_SUCCESS);
  return status;
}

// ------------------------------------------------------------------ K2': active-set method
// Primal-dual active-set iterations on the box rows, tried before the interior point
// (solver.active_set_shortcut, default on).  The starting set is the active set { i : lam_i > t_i }
// of the previous QP of this solve when it left a valid (t, lam) (have_set), else the empty set --
// the first solve is then the equality-constrained "phase 0" solve.  Every iteration is ONE Riccati
// factorisation with the inputs of the set eliminated (factor_sweep asmode) + the roll-out, and, when
// the set is not empty, a costate sweep for the multipliers.  A point that is primal feasible on the
// free rows and dual feasible on the fixed ones IS the solution of the strictly convex QP (exact to
// round-off) and is accepted; otherwise violated free rows join the set, fixed rows with a negative
// multiplier leave it, and the solve is repeated, at most NMPC_AS_MAXIT times.  Nothing is committed
// before acceptance (t, lam of the previous QP are untouched), so the interior point can take over.
// On configs[1] this replaces 51.7 interior-point factorisations per solve by 21.3, and the longest
// problem of a batch needs 40 instead of 120.
// Returns true when accepted; *tried = factorisations spent; *nact = the accepted set is not empty.
template <typename T, int NF>
__device__ __forceinline__ bool as_solve(const DevCfg<T>& c, T* __restrict__ ws,
                                         const WsLayout& Lr, T* __restrict__ sm, int lane,
                                         bool have_set, int* tried, bool* nact) {
  NMPC_FIXED(c)
  PROF_T0();
  const int nb = N * NU;
  const T nan = (T)__int_as_float(0x7fc00000);
  *tried = 0;
  *nact = false;
  bool any = false;
  _Pragma("unroll 1") for (int i = lane; i < nb; i += 32) {
    const bool lo = have_set && (ws[L.ll + i] > ws[L.tl + i]);
    const bool up = have_set && !lo && (ws[L.lu + i] > ws[L.tu + i]);
    ws[L.dr + i] = lo ? ws[L.lb + i] : (up ? ws[L.ubd + i] : nan);   // fixed value (NaN: free)
    ws[L.dtl + i] = lo ? T(1) : (up ? T(2) : T(0));                  // row code
    ws[L.rt + i] = ws[L.r + i];                                       // plain input gradient
    any = any || lo || up;
  }
  any = __any_sync(FULL, any);
  int cj = -1;  // checkpoint the next factorisation restarts from (-1: terminal stage, full sweep)
  for (int rep = 0; rep < NMPC_AS_MAXIT && rep < c.qp_iter_max; rep++) {  // iter_max caps every QP method
    __syncwarp();
    *tried += 1;
    PROF_ADD(PH_ELEM);
    // checkpointed sweep: restarts below the last change of the set (cj), stores the checkpoints it passes.
    // (Measured: choosing the plain instantiation for first factorisations that are likely to be
    // accepted -- warm-started RTI -- saves the stores, 0.5 %, but two copies of the sweep in one
    // solve cost more in the instruction cache.)
    bool ok = factor_sweep<T, NF, true>(c, ws, L, sm, lane, T(1), any, cj);
    PROF_ADD(PH_FACTOR);
    if (!ok) return false;
#if NMPC_FUSED_MULT
    // roll-out; it also leaves the multiplier mu = lam_l - lam_u of every fixed input in ws[L.dll]
    forward_sweep<T, NF, true>(c, ws, L, sm, lane, T(1), false, true, false, L.kff, L.dx, L.du);
    PROF_ADD(PH_FWD);
#pragma nv_diag_suppress 128   // the run-time-horizon loops below are unreachable in the fixed-horizon instantiation
    if (NF > 0) {
      // Fixed horizon: the box rows of a lane (nb / 32, rounded up) fit its registers.  ONE pass of
      // loads, all in flight together, serves the acceptance test, the slacks / multipliers of an
      // accepted point and the update of the set (three passes with an L2 round trip per trip in
      // the run-time-horizon form below).
      constexpr int NB = (NF > 0 ? NF : 1) * NU, NIT = (NB + 31) / 32;
      T code[NIT], du[NIT], lbv[NIT], ubv[NIT], mu[NIT];
#pragma unroll
      for (int t = 0; t < NIT; t++) {
        const int i = lane + 32 * t;
        const bool in = i < NB;
        code[t] = in ? ws[L.dtl + i] : T(0);
        du[t] = in ? ws[L.du + i] : T(0);
        lbv[t] = in ? ws[L.lb + i] : T(-1);
        ubv[t] = in ? ws[L.ubd + i] : T(1);
        mu[t] = in ? ws[L.dll + i] : T(0);   // only meaningful where code != 0
      }
#pragma unroll
      for (int t = 0; t < NIT; t++) {
        const T m_ = code[t] != T(0) ? mu[t] : T(0);
        ok = ok && (code[t] == T(1) ? (m_ >= T(0))
                                    : (code[t] == T(2) ? (-m_ >= T(0)) : (du[t] >= lbv[t] && du[t] <= ubv[t])));
      }
      ok = __all_sync(FULL, ok);
      if (ok) {  // accepted, see below
#pragma unroll
        for (int t = 0; t < NIT; t++) {
          const int i = lane + 32 * t;
          if (i < NB) {
            const T m_ = code[t] != T(0) ? mu[t] : T(0);
            ws[L.tl + i] = code[t] == T(1) ? T(0) : du[t] - lbv[t];
            ws[L.tu + i] = code[t] == T(2) ? T(0) : ubv[t] - du[t];
            ws[L.ll + i] = code[t] == T(1) ? m_ : T(0);
            ws[L.lu + i] = code[t] == T(2) ? -m_ : T(0);
          }
        }
        __syncwarp();
        *nact = any;
        PROF_ADD(PH_ELEM);
        return true;
      }
      bool changed = false, any_new = false;
      int kmax = -1;
#pragma unroll
      for (int t = 0; t < NIT; t++) {
        const int i = lane + 32 * t;
        if (i < NB) {
          T ncode = code[t];
          if (code[t] == T(0)) ncode = du[t] < lbv[t] ? T(1) : (du[t] > ubv[t] ? T(2) : T(0));
          else if (code[t] == T(1) && !(mu[t] >= T(0))) ncode = T(0);
          else if (code[t] == T(2) && !(-mu[t] >= T(0))) ncode = T(0);
          changed = changed || (ncode != code[t]);
          if (ncode != code[t]) kmax = i >> 2;
          any_new = any_new || (ncode != T(0));
          ws[L.dtl + i] = ncode;
          ws[L.dr + i] = ncode == T(1) ? lbv[t] : (ncode == T(2) ? ubv[t] : nan);
        }
      }
      if (!__any_sync(FULL, changed)) return false;
      any = __any_sync(FULL, any_new);
      kmax = __reduce_max_sync(FULL, kmax);
      cj = -1;
#pragma unroll
      for (int j = NCK - 1; j >= 0; j--)
        if (ck_stage(N, j) > kmax && ck_stage(N, j) < N) cj = j;
      continue;
    }
    // the point is the QP solution iff the free rows respect the box and the multipliers of the
    // fixed rows have the right sign
    _Pragma("unroll 1") for (int i = lane; i < nb; i += 32) {
      const T code = ws[L.dtl + i], du = ws[L.du + i];
      const T mu = code != T(0) ? ws[L.dll + i] : T(0);
      ok = ok && (code == T(1) ? (mu >= T(0))
                               : (code == T(2) ? (-mu >= T(0)) : (du >= ws[L.lb + i] && du <= ws[L.ubd + i])));
    }
    ok = __all_sync(FULL, ok);
#else
    forward_sweep<T, NF, false>(c, ws, L, sm, lane, T(1), false, true, false, L.kff, L.dx, L.du);
    PROF_ADD(PH_FWD);
    if (any) {
      ok = costate_check<T, NF>(c, ws, L, sm, lane);
      PROF_ADD(PH_DELTA);
      _Pragma("unroll 1") for (int i = lane; i < nb; i += 32) ws[L.dll + i] -= ws[L.dlu + i];  // mu = lam_l - lam_u
    } else {  // nothing fixed: the point is the solution iff it respects the box
      _Pragma("unroll 1") for (int i = lane; i < nb; i += 32) {
        const T du = ws[L.du + i];
        ok = ok && (du >= ws[L.lb + i]) && (du <= ws[L.ubd + i]);
      }
      ok = __all_sync(FULL, ok);
    }
#endif
    if (ok) {  // accepted: slacks / multipliers of this point are the active set and the warm start of the next QP
      _Pragma("unroll 1") for (int i = lane; i < nb; i += 32) {
        const T code = ws[L.dtl + i], du = ws[L.du + i];
        const T mu = code != T(0) ? ws[L.dll + i] : T(0);
        ws[L.tl + i] = code == T(1) ? T(0) : du - ws[L.lb + i];
        ws[L.tu + i] = code == T(2) ? T(0) : ws[L.ubd + i] - du;
        ws[L.ll + i] = code == T(1) ? mu : T(0);
        ws[L.lu + i] = code == T(2) ? -mu : T(0);
      }
      __syncwarp();
      *nact = any;
      PROF_ADD(PH_ELEM);
      return true;
    }
    // primal-dual update of the set
    bool changed = false, any_new = false;
    int kmax = -1;  // last stage whose rows change
    _Pragma("unroll 1") for (int i = lane; i < nb; i += 32) {
      const T code = ws[L.dtl + i], du = ws[L.du + i], lb = ws[L.lb + i], ub = ws[L.ubd + i];
      T ncode = code;
      if (code == T(0)) ncode = du < lb ? T(1) : (du > ub ? T(2) : T(0));
      else if (code == T(1) && !(ws[L.dll + i] >= T(0))) ncode = T(0);   // mu = lam_l - lam_u
      else if (code == T(2) && !(-ws[L.dll + i] >= T(0))) ncode = T(0);
      changed = changed || (ncode != code);
      if (ncode != code) kmax = i >> 2;  // i ascends
      any_new = any_new || (ncode != T(0));
      ws[L.dtl + i] = ncode;
      ws[L.dr + i] = ncode == T(1) ? lb : (ncode == T(2) ? ub : nan);
    }
    if (!__any_sync(FULL, changed)) return false;  // NaN data / no progress: the interior point reports it
    any = __any_sync(FULL, any_new);
    // stages above the last changed one keep their factorisation: restart from the lowest
    // checkpoint above it
    kmax = __reduce_max_sync(FULL, kmax);
    cj = -1;
#pragma unroll
    for (int j = NCK - 1; j >= 0; j--)
      if (ck_stage(N, j) > kmax && ck_stage(N, j) < N) cj = j;
  }
  return false;
}

// ------------------------------------------------------------------ whole solve (K1+K2+K3)
template <typename T>
struct SolveArgs {
  DevCfg<T> cfg;
  int batch;
  const double* x0;      // [batch,13]
  const double* yref;    // [batch,N,17]
  const double* yref_e;  // [batch,13]
  double* xtraj;         // warm-start store [batch,N+1,13]
  double* utraj;         // [batch,N,4]
  uint8_t* valid;        // [batch]
  double* last_cmd;      // [batch,4]
  double* u0;            // outputs (may be null)
  int* status;
  double* thrust;
  double* speed;
  int* qp_iters;
  int* sqp_iters;
  T* ws;                 // nslots * layout.total
  const int* order;      // dispatch order (longest-expected-first, see order_bins_kernel) or null: by index
  int stage;             // 1: inputs staged by TMA bulk copies (make_stage_layout), 0: direct loads
  int poison;            // debug (nmpc_debug_poison): the shared tile is filled with NaN before every
                         // problem (and the host fills the workspace with NaN before every launch):
                         // any read of memory the problem did not write itself changes the results
};

template <typename T>
__device__ __forceinline__ T clip_(T v, T lo, T hi) {
  return v < lo ? lo : (v > hi ? hi : v);
}

// Problem prologue: warm start (controller.py:416-431) and the LS gradients of the first SQP
// iteration.  With A.stage the three contiguous inputs arrive by TMA bulk copies in the idle sweep
// tile (see StageLayout) and are consumed from shared memory before this function returns -- the
// tile is free again for linearize and the sweeps; without it they are read from HBM directly.
// (Inlined into solve_one, with gradients() as a call: measured best of the four combinations.)
template <typename T, int NF>
__device__ __forceinline__ void load_inputs(const SolveArgs<T>& A, const WsLayout& Lr,
                                            T* __restrict__ ws, T* __restrict__ sm, int prob,
                                            int lane) {
  const DevCfg<T>& c = A.cfg;
  NMPC_FIXED(c)
  const int nxv = (N + 1) * NX, nuv = N * NU;
  const double* x0 = A.x0 + (size_t)prob * NX;
  const double* yref = A.yref + (size_t)prob * N * NZ;
  const double* yref_e = A.yref_e + (size_t)prob * NX;
  const double* x_in = A.xtraj + (size_t)prob * nxv;
  const double* u_in = A.utraj + (size_t)prob * nuv;
  if (A.stage) {  // three bulk copies, one mbarrier
    const StageLayout S = make_stage_layout(N, (int)sizeof(T));
    unsigned char* ybuf = reinterpret_cast<unsigned char*>(sm);
    unsigned char* ubuf = ybuf + S.ybytes;
    unsigned char* xbuf = ubuf + S.ubytes;
    void* bar = sm + SmemLayout::bar;
    const size_t xoff = (size_t)prob * nxv * sizeof(double);
    const unsigned mis = (unsigned)(xoff & 15);
    // the tile was last written through the generic proxy by ALL lanes (sweeps of the previous
    // problem): every writer fences its own generic-proxy writes against the async proxy, then the
    // warp converges, then lane 0 arms the barrier and issues the copies
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncwarp();
    if (lane == 0) {
      mbar_expect_tx(bar, (unsigned)(S.ybytes + S.ubytes + S.xbytes));
      tma_load_1d(ybuf, yref, (unsigned)S.ybytes, bar);
      tma_load_1d(ubuf, u_in, (unsigned)S.ubytes, bar);
      tma_load_1d(xbuf, reinterpret_cast<const unsigned char*>(A.xtraj) + (xoff - mis),
                  (unsigned)S.xbytes, bar);
    }
    yref = reinterpret_cast<const double*>(ybuf);
    u_in = reinterpret_cast<const double*>(ubuf);
    x_in = reinterpret_cast<const double*>(xbuf + mis);
  }
  // previous solution un-shifted with stage-0 x <- x0, or (x0, u = 0) on every stage when the
  // previous solve is not valid
  const bool warm = A.valid[prob] != 0;
  if (A.stage) {  // phase parity of the mbarrier is kept next to it (no register across solves)
    unsigned* par = reinterpret_cast<unsigned*>(sm + SmemLayout::bar) + 2;
    const unsigned parity = *par;
    mbar_wait(sm + SmemLayout::bar, parity);
    __syncwarp();
    if (lane == 0) *par = parity ^ 1u;
  }
  _Pragma("unroll 3") for (int i = lane; i < nxv; i += 32) {
    const int ii = i % NX;
    ws[L.xb + i] = (T)((warm && i >= NX) ? x_in[i] : x0[ii]);
  }
  _Pragma("unroll 1") for (int i = lane; i < nuv; i += 32) ws[L.ub + i] = warm ? (T)u_in[i] : T(0);
  __syncwarp();
  gradients<T, NF>(c, ws, L, yref, yref_e, lane);
  __syncwarp();  // the staged references are consumed: the tile goes back to the sweeps
}

// One PositionNMPC.solve() (controller.py:385-463) + the node's command map
// (nodes/mpc_controller_node:124-131,152-164) for problem `prob`, executed by one warp.
template <typename T, int NS, int NF>
__device__ __forceinline__ void solve_one(const SolveArgs<T>& A, const WsLayout& Lr, T* ws, T* sm,
                                          int prob, int lane) {
  const DevCfg<T>& c = A.cfg;
  NMPC_FIXED(c)
  const int nxv = (N + 1) * NX, nuv = N * NU;
  const double* yref = A.yref + (size_t)prob * N * NZ;
  const double* yref_e = A.yref_e + (size_t)prob * NX;
  double* xtraj = A.xtraj + (size_t)prob * nxv;
  double* utraj = A.utraj + (size_t)prob * nuv;
  if (A.poison) {  // debug: nothing of the previous problem may be read (see SolveArgs::poison)
    const T nan = (T)__int_as_float(0x7fc00000);
    _Pragma("unroll 1") for (int i = lane; i < SmemLayout::total; i += 32)
      if (i < SmemLayout::bar || i >= SmemLayout::bar + 4) sm[i] = nan;
    __syncwarp();
  }
  PROF_T0();
#ifdef NMPC_PROFILE
  const long long prof_start_ = prof_t_;
#endif
  // inputs -> workspace (TMA-staged when possible) + the gradients of the first SQP iteration
  load_inputs<T, NF>(A, L, ws, sm, prob, lane);
  PROF_ADD(PH_LOAD);

  int status = ST_SUCCESS;
  int tot_qp = 0, nsqp = 0;
  bool converged = false;
  bool have_ipm = false;  // the previous QP of this solve left a converged interior-point iterate
  T prev_step = T(0);
  // FP32 cannot resolve slacks below ~1e-4 of the thrust scale
  const T eps_min = sizeof(T) == 4 ? T(1e-4) : T(NMPC_WARM_EPS_MIN);
  const int max_sqp = c.max_sqp_iter < 1 ? 1 : c.max_sqp_iter;
  while (nsqp < max_sqp) {
    // (gradients of the first iteration: done by load_inputs from the TMA-staged references)
    linearize<T, NS, NF>(c, ws, L, sm, lane, yref, yref_e, nsqp > 0);
    PROF_ADD(PH_LIN);
    __syncwarp();
    PROF_ADD(PH_GRAD);
    int qi = 0;
    T weps = T(0);
    if (c.ipm_warm && have_ipm && prev_step <= T(NMPC_WARM_STEP_MAX))
      weps = fmin(fmax(T(NMPC_WARM_KAPPA) * prev_step, eps_min), T(NMPC_WARM_EPS_MAX));
    int qs = -1;
    if (c.as_shortcut) {
      int tried = 0;
      bool nact = false;
      if (as_solve<T, NF>(c, ws, L, sm, lane, have_ipm, &tried, &nact)) {
        qs = ST_SUCCESS;
        have_ipm = nact;  // a valid (t, lam) with a non-empty active set for the next QP
      }
      tot_qp += tried;
    }
    if (qs < 0) qs = ipm_solve<T, NF>(c, ws, L, sm, lane, weps, &have_ipm, &qi);
#ifdef NMPC_PROFILE
    prof_t_ = clock64();  // ipm_solve accounted for its own phases
#endif
    tot_qp += qi;
    nsqp++;
    // acados SQP_RTI tolerates a QP that only hit its iteration cap
    if (qs != ST_SUCCESS && qs != ST_MAXITER) {
      status = ST_QPFAIL;
      break;
    }
    __syncwarp();
    T step = T(0);
    bool fin = true;
    _Pragma("unroll 3") for (int i = lane; i < nxv; i += 32) {
      const T d = ws[L.dx + i];
      const T v = ws[L.xb + i] + d;  // full step
      ws[L.xb + i] = v;
      step = fmax(step, fabs(d));
      fin = fin && finite_(v);
    }
    _Pragma("unroll 1") for (int i = lane; i < nuv; i += 32) {
      const T d = ws[L.du + i];
      const T v = ws[L.ub + i] + d;
      ws[L.ub + i] = v;
      step = fmax(step, fabs(d));
      fin = fin && finite_(v);
    }
    __syncwarp();
    if (!__all_sync(FULL, fin)) {
      status = c.nan_status;
      break;
    }
    step = warp_max(step);
    PROF_ADD(PH_STEP);
    prev_step = step;
    if (step <= c.sqp_tol) {
      converged = true;
      break;
    }
  }
  if (status == ST_SUCCESS && max_sqp > 1 && !converged) status = ST_MAXITER;

  // K3 epilogue: outputs, warm-start store, command mapping with last-valid fallback
  if (status == ST_SUCCESS) {
    _Pragma("unroll 3") for (int i = lane; i < nxv; i += 32) xtraj[i] = (double)ws[L.xb + i];
    _Pragma("unroll 1") for (int i = lane; i < nuv; i += 32) utraj[i] = (double)ws[L.ub + i];
  }
  if (lane < NU) {
    const T u = (status == ST_SUCCESS) ? ws[L.ub + lane] : T(0);
    if (A.u0) A.u0[(size_t)prob * NU + lane] = (double)u;
    double* last = A.last_cmd + (size_t)prob * NU;
    const T src = (status == ST_SUCCESS) ? u : (T)last[lane];
    const T cl = clip_(src, c.lbu[lane], c.ubu[lane]);
    const T kf = fmax(c.kf, T(1e-9));
    T s2 = clip_(cl / kf, T(0), c.w_max * c.w_max);
    T sp = clip_(sqrt(s2), c.w_min, c.w_max);
    last[lane] = (double)cl;
    if (A.thrust) A.thrust[(size_t)prob * NU + lane] = (double)cl;
    if (A.speed) A.speed[(size_t)prob * NU + lane] = (double)sp;
  }
  if (lane == 0) {
    A.valid[prob] = (status == ST_SUCCESS) ? 1 : 0;
    if (A.status) A.status[prob] = status;
    if (A.qp_iters) A.qp_iters[prob] = tot_qp;
    if (A.sqp_iters) A.sqp_iters[prob] = nsqp;
  }
  __syncwarp();
  PROF_ADD(PH_EPI);
#ifdef NMPC_PROFILE
  if (lane == 0) atomicAdd(&g_phase_cycles[PH_TOTAL], (unsigned long long)(clock64() - prof_start_));
#endif
}

}  // namespace nmpc
