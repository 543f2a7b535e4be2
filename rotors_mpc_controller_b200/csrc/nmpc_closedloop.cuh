// nmpc_closedloop.cuh -- device-side pieces AROUND the solve:
// reference horizons, odometry -> state, plant step and per-problem run statistics, so that a
// closed-loop Monte-Carlo run never leaves HBM.  All FP64, one thread per problem (or per
// output element); these kernels are trivially HBM/latency bound (a few hundred bytes and a
// few hundred flops per problem) next to the solve kernel.
//
// Reference anchors (under /root/reference/):
//   src/rotors_mpc_controller/reference.py:75-91        set-point tiling
//   src/rotors_mpc_controller/controller.py:406-411     quaternion normalisation of the state
//   src/rotors_mpc_controller/controller.py:286-346     model integrated by the nominal plant
//   nodes/mpc_controller_node:25-44,88-113,139-150      odometry -> state (euler, R(rpy) v_body)
//   third_party/rotors_simulator/rotors_gazebo_plugins/src/gazebo_motor_model.cpp:415-483
//       + include/rotors_gazebo_plugins/common.h:147-183  first-order rotor-speed filter
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "nmpc_model.cuh"

namespace nmpc {

// yref[b,k,:] = [p, 0, q(yaw), 0, hover x4], yref_e[b,:] = [p, 0, q(yaw), 0]: what B
// ReferenceGenerators (after update_hover_thrust) + controller.py:433-445 produce.
__global__ void __launch_bounds__(256)
    reference_setpoint_kernel(int B, int N, const double* __restrict__ pos,
                              const double* __restrict__ yaw, double hover,
                              double* __restrict__ yref, double* __restrict__ yref_e) {
  const size_t total = (size_t)B * (N * NZ + NX);
  for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (size_t)gridDim.x * blockDim.x) {
    const size_t per = (size_t)N * NZ + NX;
    const int b = (int)(idx / per);
    const int r = (int)(idx - (size_t)b * per);
    const bool term = r >= N * NZ;
    const int j = term ? r - N * NZ : r % NZ;
    double v = 0.0;
    if (j < 3) {
      v = pos[(size_t)b * 3 + j];
    } else if (j >= 6 && j < 10) {
      const double half = 0.5 * (yaw ? yaw[b] : 0.0);  // reference.py:11-13
      v = (j == 6) ? cos(half) : ((j == 9) ? sin(half) : 0.0);
    } else if (j >= NX) {
      v = hover;
    }
    if (term) yref_e[(size_t)b * NX + j] = v;
    else yref[(size_t)b * N * NZ + r] = v;
  }
}

// Figure-8 track sampled at the stage times t0 + k dt:
//   p(t) = c + (A sin wt, A/2 sin 2wt, 0),  v(t) = (A w cos wt, A w cos 2wt, 0),  q = identity,
// t = t0[b] + t_off + k dt.
__global__ void __launch_bounds__(128)
    reference_figure8_kernel(int B, int N, double dt, double t_off, const double* __restrict__ t0,
                             const double* __restrict__ amp, const double* __restrict__ period,
                             double cx, double cy, double cz, double hover,
                             double* __restrict__ yref, double* __restrict__ yref_e) {
  const size_t total = (size_t)B * (N + 1);
  for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (size_t)gridDim.x * blockDim.x) {
    const int b = (int)(idx / (N + 1));
    const int k = (int)(idx - (size_t)b * (N + 1));
    const double A = amp[b];
    const double w = 2.0 * 3.14159265358979323846 / period[b];
    const double wt = w * (t0[b] + t_off + dt * (double)k);
    double row[NX];
    row[0] = cx + A * sin(wt);
    row[1] = cy + 0.5 * A * sin(2.0 * wt);
    row[2] = cz;
    row[3] = A * w * cos(wt);
    row[4] = A * w * cos(2.0 * wt);
    row[5] = 0.0;
    row[6] = 1.0;
#pragma unroll
    for (int j = 7; j < NX; j++) row[j] = 0.0;
    if (k < N) {
      double* dst = yref + ((size_t)b * N + k) * NZ;
#pragma unroll
      for (int j = 0; j < NX; j++) dst[j] = row[j];
#pragma unroll
      for (int j = NX; j < NZ; j++) dst[j] = hover;
    } else {
      double* dst = yref_e + (size_t)b * NX;
#pragma unroll
      for (int j = 0; j < NX; j++) dst[j] = row[j];
    }
  }
}

// nav_msgs/Odometry fields -> solver state, exactly as the node does it
// (nodes/mpc_controller_node:88-113): orientation (x,y,z,w) -> roll/pitch/yaw (:25-44) ->
// R(rpy) (:139-150) -> world velocity = R v_body; quaternion re-ordered to (w,x,y,z) and
// normalised (controller.py:406-411).  odom[b,:] = [p(3), q_xyzw(4), v_body(3), w(3)].
__global__ void __launch_bounds__(128)
    odometry_to_state_kernel(int B, const double* __restrict__ odom, double* __restrict__ x0,
                             int32_t* __restrict__ bad) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  const double* o = odom + (size_t)b * 13;
  double qx = o[3], qy = o[4], qz = o[5], qw = o[6];
  const double norm = sqrt(qx * qx + qy * qy + qz * qz + qw * qw);
  double roll = 0.0, pitch = 0.0, yw = 0.0;
  if (norm != 0.0) {
    const double ex = qx / norm, ey = qy / norm, ez = qz / norm, ew = qw / norm;
    roll = atan2(2.0 * (ew * ex + ey * ez), 1.0 - 2.0 * (ex * ex + ey * ey));
    const double sinp = 2.0 * (ew * ey - ez * ex);
    pitch = (fabs(sinp) >= 1.0) ? copysign(3.14159265358979323846 / 2.0, sinp) : asin(sinp);
    yw = atan2(2.0 * (ew * ez + ex * ey), 1.0 - 2.0 * (ey * ey + ez * ez));
  }
  const double cr = cos(roll), sr = sin(roll), cp = cos(pitch), sp = sin(pitch);
  const double cyw = cos(yw), syw = sin(yw);
  const double vx = o[7], vy = o[8], vz = o[9];
  double* x = x0 + (size_t)b * NX;
  x[0] = o[0];
  x[1] = o[1];
  x[2] = o[2];
  x[3] = (cp * cyw) * vx + (cyw * sp * sr - syw * cr) * vy + (cyw * sp * cr + syw * sr) * vz;
  x[4] = (cp * syw) * vx + (syw * sp * sr + cyw * cr) * vy + (syw * sp * cr - cyw * sr) * vz;
  x[5] = (-sp) * vx + (cp * sr) * vy + (cp * cr) * vz;
  // controller.py:407-408 raises ValueError on a zero quaternion: flagged per problem here
  if (bad) bad[b] = (norm == 0.0) ? 1 : 0;
  const double den = (norm != 0.0) ? norm : 1.0;  // divisions, as numpy does (controller.py:409)
  x[6] = qw / den;
  x[7] = qx / den;
  x[8] = qy / den;
  x[9] = qz / den;
  x[10] = o[10];
  x[11] = o[11];
  x[12] = o[12];
}

// Plant parameters (host-filled; mirrors nmpc_plant_t of include/nmpc_b200.h).
struct PlantCfg {
  int substeps, motor_model, rotors_yaw_sign;
  double dt, tau_up, tau_down, rotor_drag, inv_mass, rolling_moment;
};

// Controller model + optional rotor drag (gazebo_motor_model.cpp:443-453):
//   F_drag = -sum_i |w_i| * c_drag * v_perp,   v_perp = v - (v . z_b) z_b   (no wind).
__device__ __forceinline__ void plant_f(const DevCfg<double>& c, const PlantCfg& pc,
                                        const double (&x)[NX], const double (&u)[NU],
                                        double sum_abs_w, double (&f)[NX]) {
  model_f(c, x, u, f);
  if (pc.rotor_drag != 0.0) {
    const double qw = x[6], qx = x[7], qy = x[8], qz = x[9];
    const double zx = 2.0 * (qx * qz + qw * qy), zy = 2.0 * (qy * qz - qw * qx),
                 zz = 1.0 - 2.0 * (qx * qx + qy * qy);
    const double vz = x[3] * zx + x[4] * zy + x[5] * zz;
    const double k = sum_abs_w * pc.rotor_drag * pc.inv_mass;
    f[3] -= k * (x[3] - vz * zx);
    f[4] -= k * (x[4] - vz * zy);
    f[5] -= k * (x[5] - vz * zz);
  }
  if (pc.rolling_moment != 0.0) {
    // rolling moment (gazebo_motor_model.cpp:470-474): world-frame torque -sum|w_i| mu_1 v_perp on
    // the body; into the body frame with R(q)' (rotation of controller.py:294-302), then J^-1
    const double qw = x[6], qx = x[7], qy = x[8], qz = x[9];
    const double r11 = 1.0 - 2.0 * (qy * qy + qz * qz), r12 = 2.0 * (qx * qy - qw * qz),
                 r13 = 2.0 * (qx * qz + qw * qy);
    const double r21 = 2.0 * (qx * qy + qw * qz), r22 = 1.0 - 2.0 * (qx * qx + qz * qz),
                 r23 = 2.0 * (qy * qz - qw * qx);
    const double r31 = 2.0 * (qx * qz - qw * qy), r32 = 2.0 * (qy * qz + qw * qx),
                 r33 = 1.0 - 2.0 * (qx * qx + qy * qy);
    const double vz = x[3] * r13 + x[4] * r23 + x[5] * r33;
    const double k = -sum_abs_w * pc.rolling_moment;
    const double tx = k * (x[3] - vz * r13), ty = k * (x[4] - vz * r23), tz = k * (x[5] - vz * r33);
    f[10] += (r11 * tx + r21 * ty + r31 * tz) * c.invJ[0];
    f[11] += (r12 * tx + r22 * ty + r32 * tz) * c.invJ[1];
    f[12] += (r13 * tx + r23 * ty + r33 * tz) * c.invJ[2];
  }
}

// One control interval of the plant for every problem + statistics.
//   x        [B,13] plant state, in/out (quaternion NOT re-normalised, like a simulator state)
//   x0       [B,13] out: state handed to the solver (quaternion normalised, controller.py:406-411)
//   thrust   [B,4]  clipped command of this step (nmpc_solve_* output, fail-safe applied)
//   rotor    [B,4]  rotor speeds in/out when motor_model = 1 (first-order lag towards
//                   w_cmd = sqrt(thrust / kf); thrust applied = kf w^2, common.h:147-183)
//   pref     reference position of problem b at pref + b * pref_stride (3 doubles)
//   acc      [B,8]  accumulators: 0 sum |p-pref|^2, 1 last |p-pref|, 2 #status != 0,
//                   3 sum qp_iters, 4 #steps with a rotor at a thrust bound, 5 max |p-pref|,
//                   6 sum of thrusts, 7 #non-finite states
__global__ void __launch_bounds__(128)
    plant_step_kernel(const __grid_constant__ DevCfg<double> c, const __grid_constant__ PlantCfg pc,
                      int B, double* __restrict__ xs, double* __restrict__ x0,
                      const double* __restrict__ thrust, double* __restrict__ rotor,
                      const double* __restrict__ pref, size_t pref_stride,
                      const int32_t* __restrict__ status, const int32_t* __restrict__ qp_iters,
                      double* __restrict__ acc) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  double x[NX], u[NU], w[NU];
#pragma unroll
  for (int i = 0; i < NX; i++) x[i] = xs[(size_t)b * NX + i];
  bool sat = false;
  double usum = 0.0;
  if (thrust) {
#pragma unroll
    for (int i = 0; i < NU; i++) {
      u[i] = thrust[(size_t)b * NU + i];
      const double tol = 1e-9 * (c.ubu[i] - c.lbu[i]);  // "at the bound" up to solver accuracy
      sat = sat || (u[i] <= c.lbu[i] + tol) || (u[i] >= c.ubu[i] - tol);
      usum += u[i];
    }
  } else {
#pragma unroll
    for (int i = 0; i < NU; i++) u[i] = 0.0;
  }
  DevCfg<double> cp = c;  // plant mass may differ from the controller's
  cp.inv_mass = pc.inv_mass;
  if (pc.rotors_yaw_sign) {  // RotorS' drag-torque sign (gazebo_motor_model.cpp:458-459)
#pragma unroll
    for (int i = 0; i < NU; i++) cp.rz[i] = -c.rz[i];
  }
  const double h = pc.dt / (double)(pc.substeps > 0 ? pc.substeps : 1);
  double wcmd[NU];
  if (pc.motor_model) {
#pragma unroll
    for (int i = 0; i < NU; i++) {
      w[i] = rotor[(size_t)b * NU + i];
      wcmd[i] = sqrt(fmax(u[i], 0.0) / fmax(c.kf, 1e-9));
    }
  }
  const double a_up = pc.motor_model ? exp(-h / pc.tau_up) : 0.0;
  const double a_dn = pc.motor_model ? exp(-h / pc.tau_down) : 0.0;
  for (int s = 0; s < pc.substeps; s++) {
    double ua[NU], saw = 0.0;
#pragma unroll
    for (int i = 0; i < NU; i++) {
      ua[i] = pc.motor_model ? c.kf * w[i] * w[i] : u[i];
      saw += pc.motor_model ? fabs(w[i]) : sqrt(fmax(u[i], 0.0) / fmax(c.kf, 1e-9));
    }
    double k1[NX], k2[NX], k3[NX], k4[NX], t[NX];
    plant_f(cp, pc, x, ua, saw, k1);
#pragma unroll
    for (int i = 0; i < NX; i++) t[i] = x[i] + 0.5 * h * k1[i];
    plant_f(cp, pc, t, ua, saw, k2);
#pragma unroll
    for (int i = 0; i < NX; i++) t[i] = x[i] + 0.5 * h * k2[i];
    plant_f(cp, pc, t, ua, saw, k3);
#pragma unroll
    for (int i = 0; i < NX; i++) t[i] = x[i] + h * k3[i];
    plant_f(cp, pc, t, ua, saw, k4);
#pragma unroll
    for (int i = 0; i < NX; i++) x[i] += h / 6.0 * (k1[i] + 2.0 * k2[i] + 2.0 * k3[i] + k4[i]);
    if (pc.motor_model) {
#pragma unroll
      for (int i = 0; i < NU; i++) {
        const double a = (wcmd[i] > w[i]) ? a_up : a_dn;
        w[i] = a * w[i] + (1.0 - a) * wcmd[i];
      }
    }
  }
  bool fin = true;
#pragma unroll
  for (int i = 0; i < NX; i++) {
    xs[(size_t)b * NX + i] = x[i];
    fin = fin && ((x[i] - x[i]) == 0.0);
  }
  if (pc.motor_model) {
#pragma unroll
    for (int i = 0; i < NU; i++) rotor[(size_t)b * NU + i] = w[i];
  }
  if (x0) {
    const double nq = sqrt(x[6] * x[6] + x[7] * x[7] + x[8] * x[8] + x[9] * x[9]);
    const double den = (nq != 0.0) ? nq : 1.0;
#pragma unroll
    for (int i = 0; i < NX; i++)
      x0[(size_t)b * NX + i] = (i >= 6 && i < 10) ? x[i] / den : x[i];
  }
  if (acc && pc.substeps > 0) {
    double* a = acc + (size_t)b * 8;
    double e2 = 0.0;
    if (pref) {
      const double* pr = pref + (size_t)b * pref_stride;
      const double ex = x[0] - pr[0], ey = x[1] - pr[1], ez = x[2] - pr[2];
      e2 = ex * ex + ey * ey + ez * ez;
    }
    const double e = sqrt(e2);
    a[0] += e2;
    a[1] = e;
    if (status) a[2] += (status[b] != 0) ? 1.0 : 0.0;
    if (qp_iters) a[3] += (double)qp_iters[b];
    a[4] += sat ? 1.0 : 0.0;
    a[5] = fmax(a[5], e);
    a[6] += usum;
    a[7] += fin ? 0.0 : 1.0;
  }
}

}  // namespace nmpc
