/*
 * nmpc_oracle.c -- CPU ORACLE (test infrastructure, NOT the product).  See nmpc_oracle.h.
 * PARITY UNPINNED (no reference goldens exist; acados is not installable here).
 *
 * Formulation restated (file:line are into /root/reference/src/rotors_mpc_controller/):
 *   model            controller.py:267-355
 *   options          controller.py:179-190  (ERK ns=2 x 2 steps, GAUSS_NEWTON, LM, SQP_RTI default)
 *   cost / bounds    controller.py:223-261
 *   solve semantics  controller.py:401-463
 */
#include "nmpc_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

#define NX ORC_NX
#define NU ORC_NU
#define NZ ORC_NZ

/* ------------------------------------------------------------------ model */
/* controller.py:286-346 : f_expl = [v; R(q) e3 T/m - g e3; 0.5 q (x) w; J^-1(tau - w x Jw)] */
void orc_f(const orc_config *c, const double *x, const double *u, double *f) {
  const double qw = x[6], qx = x[7], qy = x[8], qz = x[9];
  const double wx = x[10], wy = x[11], wz = x[12];
  /* third column of the rotation built from the un-normalised quaternion (:294-302) */
  const double r13 = 2.0 * (qx * qz + qw * qy);
  const double r23 = 2.0 * (qy * qz - qw * qx);
  const double r33 = 1.0 - 2.0 * (qx * qx + qy * qy);
  const double T = u[0] + u[1] + u[2] + u[3]; /* :310 */
  const double a = T / c->mass;
  f[0] = x[3];
  f[1] = x[4];
  f[2] = x[5];
  f[3] = r13 * a;
  f[4] = r23 * a;
  f[5] = r33 * a - c->gravity; /* :313-314 */
  f[6] = 0.5 * (-qx * wx - qy * wy - qz * wz); /* :316-319 */
  f[7] = 0.5 * (qw * wx + qy * wz - qz * wy);
  f[8] = 0.5 * (qw * wy + qz * wx - qx * wz);
  f[9] = 0.5 * (qw * wz + qx * wy - qy * wx);
  double tx = 0, ty = 0, tz = 0; /* :326-328 */
  for (int i = 0; i < 4; i++) {
    tx += u[i] * c->rotor_y[i];
    ty += u[i] * (-c->rotor_x[i]);
    tz += u[i] * c->rotor_z[i];
  }
  const double Jx = c->J[0], Jy = c->J[1], Jz = c->J[2];
  /* omega x (J omega), :331-337 */
  const double cx = wy * (Jz * wz) - wz * (Jy * wy);
  const double cy = wz * (Jx * wx) - wx * (Jz * wz);
  const double cz = wx * (Jy * wy) - wy * (Jx * wx);
  f[10] = (tx - cx) / Jx; /* :338-341 */
  f[11] = (ty - cy) / Jy;
  f[12] = (tz - cz) / Jz;
}

/* Dense analytic Jacobians of orc_f (hand-derived, checked by complex step in tests). */
void orc_jac(const orc_config *c, const double *x, const double *u, double *Fx, double *Fu) {
  memset(Fx, 0, sizeof(double) * NX * NX);
  memset(Fu, 0, sizeof(double) * NX * NU);
  const double qw = x[6], qx = x[7], qy = x[8], qz = x[9];
  const double wx = x[10], wy = x[11], wz = x[12];
  const double T = u[0] + u[1] + u[2] + u[3];
  const double a = T / c->mass;
  const double im = 1.0 / c->mass;
#define FX(i, j) Fx[(i)*NX + (j)]
#define FU(i, j) Fu[(i)*NU + (j)]
  FX(0, 3) = 1.0;
  FX(1, 4) = 1.0;
  FX(2, 5) = 1.0;
  /* d(vdot)/dq */
  FX(3, 6) = 2.0 * qy * a;
  FX(3, 7) = 2.0 * qz * a;
  FX(3, 8) = 2.0 * qw * a;
  FX(3, 9) = 2.0 * qx * a;
  FX(4, 6) = -2.0 * qx * a;
  FX(4, 7) = -2.0 * qw * a;
  FX(4, 8) = 2.0 * qz * a;
  FX(4, 9) = 2.0 * qy * a;
  FX(5, 7) = -4.0 * qx * a;
  FX(5, 8) = -4.0 * qy * a;
  const double r13 = 2.0 * (qx * qz + qw * qy);
  const double r23 = 2.0 * (qy * qz - qw * qx);
  const double r33 = 1.0 - 2.0 * (qx * qx + qy * qy);
  for (int i = 0; i < 4; i++) {
    FU(3, i) = r13 * im;
    FU(4, i) = r23 * im;
    FU(5, i) = r33 * im;
  }
  /* d(qdot)/dq */
  FX(6, 7) = -0.5 * wx;
  FX(6, 8) = -0.5 * wy;
  FX(6, 9) = -0.5 * wz;
  FX(7, 6) = 0.5 * wx;
  FX(7, 8) = 0.5 * wz;
  FX(7, 9) = -0.5 * wy;
  FX(8, 6) = 0.5 * wy;
  FX(8, 7) = -0.5 * wz;
  FX(8, 9) = 0.5 * wx;
  FX(9, 6) = 0.5 * wz;
  FX(9, 7) = 0.5 * wy;
  FX(9, 8) = -0.5 * wx;
  /* d(qdot)/dw */
  FX(6, 10) = -0.5 * qx;
  FX(6, 11) = -0.5 * qy;
  FX(6, 12) = -0.5 * qz;
  FX(7, 10) = 0.5 * qw;
  FX(7, 11) = -0.5 * qz;
  FX(7, 12) = 0.5 * qy;
  FX(8, 10) = 0.5 * qz;
  FX(8, 11) = 0.5 * qw;
  FX(8, 12) = -0.5 * qx;
  FX(9, 10) = -0.5 * qy;
  FX(9, 11) = 0.5 * qx;
  FX(9, 12) = 0.5 * qw;
  /* d(wdot)/dw : wdot_x = (tx - (Jz-Jy) wy wz)/Jx etc. */
  const double Jx = c->J[0], Jy = c->J[1], Jz = c->J[2];
  FX(10, 11) = -(Jz - Jy) * wz / Jx;
  FX(10, 12) = -(Jz - Jy) * wy / Jx;
  FX(11, 10) = -(Jx - Jz) * wz / Jy;
  FX(11, 12) = -(Jx - Jz) * wx / Jy;
  FX(12, 10) = -(Jy - Jx) * wy / Jz;
  FX(12, 11) = -(Jy - Jx) * wx / Jz;
  for (int i = 0; i < 4; i++) {
    FU(10, i) = c->rotor_y[i] / Jx;
    FU(11, i) = -c->rotor_x[i] / Jy;
    FU(12, i) = c->rotor_z[i] / Jz;
  }
#undef FX
#undef FU
}

/* Butcher tableaus of acados' ERK (UPSTREAM: sim_erk: ns=1 Euler, 2 midpoint, 4 classic RK4) */
static void tableau(int ns, double *A, double *bvec) {
  memset(A, 0, sizeof(double) * 16);
  memset(bvec, 0, sizeof(double) * 4);
  if (ns == 1) {
    bvec[0] = 1.0;
  } else if (ns == 2) {
    A[1 * 4 + 0] = 0.5;
    bvec[1] = 1.0;
  } else { /* 4 */
    A[1 * 4 + 0] = 0.5;
    A[2 * 4 + 1] = 0.5;
    A[3 * 4 + 2] = 1.0;
    bvec[0] = 1.0 / 6.0;
    bvec[1] = 1.0 / 3.0;
    bvec[2] = 1.0 / 3.0;
    bvec[3] = 1.0 / 6.0;
  }
}

/* S <- Fx * S + [0 | Fu]  evaluated into K (13x17, columns [x(13) | u(4)]) */
static void vde(const double *Fx, const double *Fu, const double *S, double *K) {
  for (int i = 0; i < NX; i++)
    for (int j = 0; j < NZ; j++) {
      double s = (j >= NX) ? Fu[i * NU + (j - NX)] : 0.0;
      for (int k = 0; k < NX; k++) s += Fx[i * NX + k] * S[k * NZ + j];
      K[i * NZ + j] = s;
    }
}

void orc_discretize(const orc_config *c, const double *x, const double *u, double *xn, double *A,
                    double *B) {
  double At[16], bt[4];
  const int ns = c->integ_ns;
  tableau(ns, At, bt);
  const double h = c->dt / (double)c->integ_steps;
  double xc[NX], S[NX * NZ];
  memcpy(xc, x, sizeof(xc));
  memset(S, 0, sizeof(S));
  for (int i = 0; i < NX; i++) S[i * NZ + i] = 1.0;
  double kx[4][NX], kS[4][NX * NZ];
  for (int step = 0; step < c->integ_steps; step++) {
    for (int s = 0; s < ns; s++) {
      double xs[NX], Ss[NX * NZ];
      memcpy(xs, xc, sizeof(xs));
      memcpy(Ss, S, sizeof(Ss));
      for (int j = 0; j < s; j++) {
        const double a = At[s * 4 + j];
        if (a == 0.0) continue;
        for (int i = 0; i < NX; i++) xs[i] += h * a * kx[j][i];
        for (int i = 0; i < NX * NZ; i++) Ss[i] += h * a * kS[j][i];
      }
      double Fx[NX * NX], Fu[NX * NU];
      orc_f(c, xs, u, kx[s]);
      orc_jac(c, xs, u, Fx, Fu);
      vde(Fx, Fu, Ss, kS[s]);
    }
    for (int s = 0; s < ns; s++) {
      if (bt[s] == 0.0) continue;
      for (int i = 0; i < NX; i++) xc[i] += h * bt[s] * kx[s][i];
      for (int i = 0; i < NX * NZ; i++) S[i] += h * bt[s] * kS[s][i];
    }
  }
  memcpy(xn, xc, sizeof(xc));
  for (int i = 0; i < NX; i++) {
    for (int j = 0; j < NX; j++) A[i * NX + j] = S[i * NZ + j];
    for (int j = 0; j < NU; j++) B[i * NU + j] = S[i * NZ + NX + j];
  }
}

void orc_plant_step(const orc_config *c, const double *x, const double *u, double dt, int substeps,
                    double *xn) {
  double xc[NX];
  memcpy(xc, x, sizeof(xc));
  const double h = dt / (double)substeps;
  for (int s = 0; s < substeps; s++) {
    double k1[NX], k2[NX], k3[NX], k4[NX], t[NX];
    orc_f(c, xc, u, k1);
    for (int i = 0; i < NX; i++) t[i] = xc[i] + 0.5 * h * k1[i];
    orc_f(c, t, u, k2);
    for (int i = 0; i < NX; i++) t[i] = xc[i] + 0.5 * h * k2[i];
    orc_f(c, t, u, k3);
    for (int i = 0; i < NX; i++) t[i] = xc[i] + h * k3[i];
    orc_f(c, t, u, k4);
    for (int i = 0; i < NX; i++) xc[i] += h / 6.0 * (k1[i] + 2.0 * k2[i] + 2.0 * k3[i] + k4[i]);
  }
  memcpy(xn, xc, sizeof(xc));
}

/* Plant with the optional RotorS motor model: first-order rotor-speed filter
 * (third_party/rotors_simulator/rotors_gazebo_plugins/include/rotors_gazebo_plugins/common.h:147-183,
 * applied once per integration step like gazebo_motor_model.cpp:476-477, thrust = kf w^2 from the
 * speed BEFORE the filter update, :432-434) and rotor drag -sum|w_i| c_drag v_perp (:443-453).
 * rotor[4] is in/out when motor_model = 1.  mass <= 0 selects the controller's mass. */
void orc_plant_step_ext(const orc_config *c, const double *x, const double *u, double dt,
                        int substeps, int motor_model, double tau_up, double tau_down,
                        double rotor_drag, double mass, double *rotor, double *xn) {
  orc_plant_step_full(c, x, u, dt, substeps, motor_model, tau_up, tau_down, rotor_drag, mass, 0.0, 0,
                      rotor, xn);
}

/* + rolling moment -sum|w_i| mu_1 v_perp as a world-frame torque (gazebo_motor_model.cpp:470-474,
 * coefficient hummingbird.xacro:42) and the plant-side yaw-torque sign (:458-459; opposite to the
 * controller's, controller.py:102-103): the reference's real model mismatch, opt-in. */
void orc_plant_step_full(const orc_config *c, const double *x, const double *u, double dt,
                         int substeps, int motor_model, double tau_up, double tau_down,
                         double rotor_drag, double mass, double rolling_moment, int rotors_yaw_sign,
                         double *rotor, double *xn) {
  orc_config cp = *c;
  if (mass > 0.0) cp.mass = mass;
  if (rotors_yaw_sign)
    for (int i = 0; i < NU; i++) cp.rotor_z[i] = -c->rotor_z[i];
  double xc[NX], w[NU], wcmd[NU];
  memcpy(xc, x, sizeof(xc));
  const double h = dt / (double)(substeps > 0 ? substeps : 1);
  const double kf = c->kf > 1e-9 ? c->kf : 1e-9;
  for (int i = 0; i < NU; i++) {
    wcmd[i] = sqrt((u[i] > 0.0 ? u[i] : 0.0) / kf);
    w[i] = motor_model ? rotor[i] : wcmd[i];
  }
  const double a_up = motor_model ? exp(-h / tau_up) : 0.0;
  const double a_dn = motor_model ? exp(-h / tau_down) : 0.0;
  for (int s = 0; s < substeps; s++) {
    double ua[NU], saw = 0.0;
    for (int i = 0; i < NU; i++) {
      ua[i] = motor_model ? c->kf * w[i] * w[i] : u[i];
      saw += fabs(w[i]);
    }
    double k[4][NX], t[NX];
    for (int st = 0; st < 4; st++) {
      const double a = (st == 0) ? 0.0 : (st == 3 ? 1.0 : 0.5);
      for (int i = 0; i < NX; i++) t[i] = (st == 0) ? xc[i] : xc[i] + a * h * k[st - 1][i];
      orc_f(&cp, t, ua, k[st]);
      if (rotor_drag != 0.0) {
        const double qw = t[6], qx = t[7], qy = t[8], qz = t[9];
        const double zx = 2.0 * (qx * qz + qw * qy), zy = 2.0 * (qy * qz - qw * qx),
                     zz = 1.0 - 2.0 * (qx * qx + qy * qy);
        const double vz = t[3] * zx + t[4] * zy + t[5] * zz;
        const double kd = saw * rotor_drag * (1.0 / cp.mass);
        k[st][3] -= kd * (t[3] - vz * zx);
        k[st][4] -= kd * (t[4] - vz * zy);
        k[st][5] -= kd * (t[5] - vz * zz);
      }
      if (rolling_moment != 0.0) {
        const double qw = t[6], qx = t[7], qy = t[8], qz = t[9];
        const double R[3][3] = {
            {1.0 - 2.0 * (qy * qy + qz * qz), 2.0 * (qx * qy - qw * qz), 2.0 * (qx * qz + qw * qy)},
            {2.0 * (qx * qy + qw * qz), 1.0 - 2.0 * (qx * qx + qz * qz), 2.0 * (qy * qz - qw * qx)},
            {2.0 * (qx * qz - qw * qy), 2.0 * (qy * qz + qw * qx), 1.0 - 2.0 * (qx * qx + qy * qy)}};
        const double vz = t[3] * R[0][2] + t[4] * R[1][2] + t[5] * R[2][2];
        const double kk = -saw * rolling_moment;
        const double tw[3] = {kk * (t[3] - vz * R[0][2]), kk * (t[4] - vz * R[1][2]),
                              kk * (t[5] - vz * R[2][2])};
        for (int a = 0; a < 3; a++)
          k[st][10 + a] += (R[0][a] * tw[0] + R[1][a] * tw[1] + R[2][a] * tw[2]) / cp.J[a];
      }
    }
    for (int i = 0; i < NX; i++)
      xc[i] += h / 6.0 * (k[0][i] + 2.0 * k[1][i] + 2.0 * k[2][i] + k[3][i]);
    if (motor_model)
      for (int i = 0; i < NU; i++) {
        const double a = (wcmd[i] > w[i]) ? a_up : a_dn;
        w[i] = a * w[i] + (1.0 - a) * wcmd[i];
      }
  }
  memcpy(xn, xc, sizeof(xc));
  if (motor_model) memcpy(rotor, w, sizeof(w));
}
