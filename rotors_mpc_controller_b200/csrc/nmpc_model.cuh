// nmpc_model.cuh -- hummingbird rotor-thrust model and its forward-mode derivative, sm_100a.
//
// Replaces the CasADi-generated `*_expl_vde_forw` that acados' ERK integrator calls for the
// model of /root/reference/src/rotors_mpc_controller/controller.py:267-355.  Hand-written:
// one lane carries ONE column of the sensitivity matrix S = [dx/du | dx/dx0] (13 x 17), so the
// Jacobian-times-column product is a directional derivative of f evaluated in registers.
#pragma once
#include <cuda_runtime.h>

namespace nmpc {

constexpr int NX = 13;
constexpr int NU = 4;
constexpr int NZ = 17;  // per-stage variable order used on the device: z = [u(4); x(13)]

// Device-side solver constants (scalar type T = double or float).
template <typename T>
struct DevCfg {
  int N, qp_iter_max, integ_steps, max_sqp_iter, nan_status;
  bool ipm_warm;      // interior-point warm start between the QPs of one SQP run
  bool as_shortcut;   // active-set shortcut between the QPs of one SQP run (see as_solve)
  T dt, h;            // shooting interval and integrator sub-step
  T Wq[NX], Wr[NU];   // scaled stage weights (sc * W)
  T Qd[NX], Rd[NU];   // stage Hessian diagonals incl. LM
  T We[NX], QN[NX];   // terminal weights, terminal Hessian diagonal incl. LM
  T inv_mass, gravity;
  T J[3], invJ[3];
  T rx[NU], ry[NU], rz[NU];
  T lbu[NU], ubu[NU];
  T kf, w_min, w_max, hover;
  T sqp_tol, ipm_tol;
};

// Two consecutive elements moved by one 16-byte (FP64) / 8-byte (FP32) access.
template <typename T>
struct alignas(2 * sizeof(T)) Pair {
  T a, b;
};
template <typename T>
__device__ __forceinline__ Pair<T> ld2(const T* p) {
  return *reinterpret_cast<const Pair<T>*>(p);
}

// f(x,u) and its directional derivative df = F_x sx + F_u su   (controller.py:286-346).
template <typename T>
__device__ __forceinline__ void model_jvp(const DevCfg<T>& c, const T (&x)[NX], const T (&u)[NU],
                                          const T (&sx)[NX], const T (&su)[NU], T (&f)[NX],
                                          T (&df)[NX]) {
  const T qw = x[6], qx = x[7], qy = x[8], qz = x[9];
  const T wx = x[10], wy = x[11], wz = x[12];
  const T dqw = sx[6], dqx = sx[7], dqy = sx[8], dqz = sx[9];
  const T dwx = sx[10], dwy = sx[11], dwz = sx[12];
  const T two = T(2), half = T(0.5);
  // pdot = v                                                     (:286, :343)
  f[0] = x[3];
  f[1] = x[4];
  f[2] = x[5];
  df[0] = sx[3];
  df[1] = sx[4];
  df[2] = sx[5];
  // third column of R(q) from the UN-normalised quaternion       (:294-302)
  const T r13 = two * (qx * qz + qw * qy);
  const T r23 = two * (qy * qz - qw * qx);
  const T r33 = T(1) - two * (qx * qx + qy * qy);
  const T dr13 = two * (dqx * qz + qx * dqz + dqw * qy + qw * dqy);
  const T dr23 = two * (dqy * qz + qy * dqz - dqw * qx - qw * dqx);
  const T dr33 = -T(4) * (qx * dqx + qy * dqy);
  const T a = (u[0] + u[1] + u[2] + u[3]) * c.inv_mass;  // T/m           (:310-313)
  const T da = (su[0] + su[1] + su[2] + su[3]) * c.inv_mass;
  f[3] = r13 * a;
  f[4] = r23 * a;
  f[5] = r33 * a - c.gravity;  //                                           (:314)
  df[3] = dr13 * a + r13 * da;
  df[4] = dr23 * a + r23 * da;
  df[5] = dr33 * a + r33 * da;
  // qdot = 1/2 q (x) [0, w]                                       (:316-319)
  f[6] = half * (-qx * wx - qy * wy - qz * wz);
  f[7] = half * (qw * wx + qy * wz - qz * wy);
  f[8] = half * (qw * wy + qz * wx - qx * wz);
  f[9] = half * (qw * wz + qx * wy - qy * wx);
  df[6] = half * (-dqx * wx - qx * dwx - dqy * wy - qy * dwy - dqz * wz - qz * dwz);
  df[7] = half * (dqw * wx + qw * dwx + dqy * wz + qy * dwz - dqz * wy - qz * dwy);
  df[8] = half * (dqw * wy + qw * dwy + dqz * wx + qz * dwx - dqx * wz - qx * dwz);
  df[9] = half * (dqw * wz + qw * dwz + dqx * wy + qx * dwy - dqy * wx - qy * dwx);
  // torques                                                       (:326-328)
  T tx = T(0), ty = T(0), tz = T(0), dtx = T(0), dty = T(0), dtz = T(0);
#pragma unroll
  for (int i = 0; i < NU; i++) {
    tx += u[i] * c.ry[i];
    ty -= u[i] * c.rx[i];
    tz += u[i] * c.rz[i];
    dtx += su[i] * c.ry[i];
    dty -= su[i] * c.rx[i];
    dtz += su[i] * c.rz[i];
  }
  // wdot = J^-1 (tau - w x J w)                                   (:331-341)
  const T Jx = c.J[0], Jy = c.J[1], Jz = c.J[2];
  const T cx = wy * (Jz * wz) - wz * (Jy * wy);
  const T cy = wz * (Jx * wx) - wx * (Jz * wz);
  const T cz = wx * (Jy * wy) - wy * (Jx * wx);
  const T dcx = (Jz - Jy) * (dwy * wz + wy * dwz);
  const T dcy = (Jx - Jz) * (dwz * wx + wz * dwx);
  const T dcz = (Jy - Jx) * (dwx * wy + wx * dwy);
  f[10] = (tx - cx) * c.invJ[0];
  f[11] = (ty - cy) * c.invJ[1];
  f[12] = (tz - cz) * c.invJ[2];
  df[10] = (dtx - dcx) * c.invJ[0];
  df[11] = (dty - dcy) * c.invJ[1];
  df[12] = (dtz - dcz) * c.invJ[2];
}

// f(x,u) only (plant roll-outs).
template <typename T>
__device__ __forceinline__ void model_f(const DevCfg<T>& c, const T (&x)[NX], const T (&u)[NU],
                                        T (&f)[NX]) {
  T z13[NX], z4[NU], df[NX];
#pragma unroll
  for (int i = 0; i < NX; i++) z13[i] = T(0);
#pragma unroll
  for (int i = 0; i < NU; i++) z4[i] = T(0);
  model_jvp(c, x, u, z13, z4, f, df);  // df is dead code and is eliminated
}

// One shooting interval of the explicit RK scheme (NS stages, c.integ_steps sub-steps) for the
// nominal state x and ONE sensitivity column s (seeded by the caller: e_i for a state column,
// 0 for an input column whose seed lives in su).  On return x = phi(x,u), s = column of [B|A].
// Tableaus (acados sim_erk): NS=1 Euler, NS=2 explicit midpoint, NS=4 classic RK4; in all of
// them stage i only depends on stage i-1, so a low-storage form is used.
template <typename T, int NS>
__device__ __forceinline__ void erk_column(const DevCfg<T>& c, T (&x)[NX], const T (&u)[NU],
                                           T (&s)[NX], const T (&su)[NU]) {
  const T h = c.h;
  for (int step = 0; step < c.integ_steps; step++) {
    T k[NX], dk[NX];
    if (NS == 1) {
      model_jvp(c, x, u, s, su, k, dk);
#pragma unroll
      for (int i = 0; i < NX; i++) {
        x[i] += h * k[i];
        s[i] += h * dk[i];
      }
    } else if (NS == 2) {
      T xs[NX], ss[NX];
      model_jvp(c, x, u, s, su, k, dk);
#pragma unroll
      for (int i = 0; i < NX; i++) {
        xs[i] = x[i] + (T(0.5) * h) * k[i];
        ss[i] = s[i] + (T(0.5) * h) * dk[i];
      }
      model_jvp(c, xs, u, ss, su, k, dk);
#pragma unroll
      for (int i = 0; i < NX; i++) {
        x[i] += h * k[i];
        s[i] += h * dk[i];
      }
    } else {
      T xs[NX], ss[NX], xa[NX], sa[NX];
      const T a_next[4] = {T(0.5), T(0.5), T(1), T(0)};
      const T bw[4] = {T(1) / T(6), T(1) / T(3), T(1) / T(3), T(1) / T(6)};
#pragma unroll
      for (int i = 0; i < NX; i++) {
        xs[i] = x[i];
        ss[i] = s[i];
        xa[i] = x[i];
        sa[i] = s[i];
      }
#pragma unroll
      for (int st = 0; st < 4; st++) {
        model_jvp(c, xs, u, ss, su, k, dk);
#pragma unroll
        for (int i = 0; i < NX; i++) {
          xa[i] += (h * bw[st]) * k[i];
          sa[i] += (h * bw[st]) * dk[i];
          xs[i] = x[i] + (h * a_next[st]) * k[i];
          ss[i] = s[i] + (h * a_next[st]) * dk[i];
        }
      }
#pragma unroll
      for (int i = 0; i < NX; i++) {
        x[i] = xa[i];
        s[i] = sa[i];
      }
    }
  }
}

}  // namespace nmpc
