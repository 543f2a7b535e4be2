/*
 * nmpc_oracle_qp.c -- CPU ORACLE (test infrastructure, NOT the product).  See nmpc_oracle.h.
 * PARITY UNPINNED.  QP assembly (acados LINEAR_LS + Gauss-Newton + LM, delta form) and a
 * box-constrained Riccati interior-point solver standing in for HPIPM
 * (controller.py:181-185; HPIPM itself is un-vendored).  The QP is strictly convex, so
 * its primal optimum is unique and any tightly converged solver pins it.
 */
#include <math.h>
#ifdef ORC_TRACE
#include <stdio.h>
#endif
#include <stdlib.h>
#include <string.h>

#include "nmpc_oracle.h"

#define NX ORC_NX
#define NU ORC_NU
#define NZ ORC_NZ

orc_qp *orc_qp_alloc(int N) {
  orc_qp *qp = (orc_qp *)calloc(1, sizeof(orc_qp));
  qp->N = N;
  qp->A = (double *)calloc((size_t)N * NX * NX, sizeof(double));
  qp->B = (double *)calloc((size_t)N * NX * NU, sizeof(double));
  qp->b = (double *)calloc((size_t)N * NX, sizeof(double));
  qp->Q = (double *)calloc((size_t)(N + 1) * NX, sizeof(double));
  qp->R = (double *)calloc((size_t)N * NU, sizeof(double));
  qp->q = (double *)calloc((size_t)(N + 1) * NX, sizeof(double));
  qp->r = (double *)calloc((size_t)N * NU, sizeof(double));
  qp->lb = (double *)calloc((size_t)N * NU, sizeof(double));
  qp->ub = (double *)calloc((size_t)N * NU, sizeof(double));
  qp->lbx = (double *)calloc((size_t)(N + 1) * NX, sizeof(double));
  qp->ubx = (double *)calloc((size_t)(N + 1) * NX, sizeof(double));
  qp->xbox = 0;
  return qp;
}

void orc_qp_free(orc_qp *qp) {
  if (!qp) return;
  free(qp->A);
  free(qp->B);
  free(qp->b);
  free(qp->Q);
  free(qp->R);
  free(qp->q);
  free(qp->r);
  free(qp->lb);
  free(qp->ub);
  free(qp->lbx);
  free(qp->ubx);
  free(qp);
}

/* controller.py:223-245 (LINEAR_LS, y=[x;u]), :189-190 (LM), :248-250 (u box);
 * UPSTREAM acados: stage cost scaled by the time step, terminal by 1. */
void orc_build_qp(const orc_config *c, const double *xtraj, const double *utraj,
                  const double *yref, const double *yref_e, orc_qp *qp) {
  const int N = c->N;
  const double sc = c->cost_scale_dt ? c->dt : 1.0;
  const double lm_s = c->lm_scaled_by_dt ? c->lm * c->dt : c->lm;
  for (int k = 0; k < N; k++) {
    const double *x = xtraj + (size_t)k * NX;
    const double *u = utraj + (size_t)k * NU;
    double xn[NX];
    orc_discretize(c, x, u, xn, qp->A + (size_t)k * NX * NX, qp->B + (size_t)k * NX * NU);
    for (int i = 0; i < NX; i++) {
      qp->b[k * NX + i] = xn[i] - xtraj[(size_t)(k + 1) * NX + i];
      qp->Q[k * NX + i] = sc * c->W[i] + lm_s;
      qp->q[k * NX + i] = sc * c->W[i] * (x[i] - yref[(size_t)k * NZ + i]);
    }
    for (int i = 0; i < NU; i++) {
      qp->R[k * NU + i] = sc * c->W[NX + i] + lm_s;
      qp->r[k * NU + i] = sc * c->W[NX + i] * (u[i] - yref[(size_t)k * NZ + NX + i]);
      qp->lb[k * NU + i] = c->lbu[i] - u[i];
      qp->ub[k * NU + i] = c->ubu[i] - u[i];
    }
  }
  const double *xN = xtraj + (size_t)N * NX;
  for (int i = 0; i < NX; i++) {
    qp->Q[N * NX + i] = c->We[i] + c->lm;
    qp->q[N * NX + i] = c->We[i] * (xN[i] - yref_e[i]);
  }
  /* controller.py:257-261: lbx = -1e6, ubx = +1e6 on every state, stages 1..N-1 */
  qp->xbox = c->keep_state_box ? 1 : 0;
  if (qp->xbox)
    for (int k = 1; k < N; k++)
      for (int i = 0; i < NX; i++) {
        qp->lbx[k * NX + i] = -c->state_box - xtraj[(size_t)k * NX + i];
        qp->ubx[k * NX + i] = c->state_box - xtraj[(size_t)k * NX + i];
      }
}

/* ---------------------------------------------------------------- Riccati */
typedef struct {
  int N;
  double *K;    /* N*4*13 feedback */
  double *Qi;   /* N*16  inverse of Quu */
  double *kff;  /* N*4 */
} ric_ws;

static int chol4_inv(const double *G, double *Ginv) {
  /* inverse of a symmetric positive definite 4x4 via Cholesky */
  double L[16] = {0};
  for (int j = 0; j < 4; j++) {
    double d = G[j * 4 + j];
    for (int k = 0; k < j; k++) d -= L[j * 4 + k] * L[j * 4 + k];
    if (!(d > 0.0)) return 1;
    L[j * 4 + j] = sqrt(d);
    for (int i = j + 1; i < 4; i++) {
      double s = G[i * 4 + j];
      for (int k = 0; k < j; k++) s -= L[i * 4 + k] * L[j * 4 + k];
      L[i * 4 + j] = s / L[j * 4 + j];
    }
  }
  /* Linv */
  double Li[16] = {0};
  for (int j = 0; j < 4; j++) {
    Li[j * 4 + j] = 1.0 / L[j * 4 + j];
    for (int i = j + 1; i < 4; i++) {
      double s = 0.0;
      for (int k = j; k < i; k++) s -= L[i * 4 + k] * Li[k * 4 + j];
      Li[i * 4 + j] = s / L[i * 4 + i];
    }
  }
  for (int i = 0; i < 4; i++)
    for (int j = 0; j < 4; j++) {
      double s = 0.0;
      for (int k = 0; k < 4; k++) s += Li[k * 4 + i] * Li[k * 4 + j];
      Ginv[i * 4 + j] = s;
    }
  return 0;
}

/* Backward factorisation with modified input Hessian Rt and gradient rt, fused with the
 * right-hand side (rho*b, rho*q, rt); forward roll-out from dx_0 = 0.  Returns non-zero on breakdown. */
static int riccati_factor_solve(const orc_qp *qp, double rho, const double *Qt, const double *qt,
                                const double *Rt, const double *rt, ric_ws *ws, double *dx,
                                double *du) {
  /* Qt / qt: state Hessian diagonal and state gradient ((N+1)*13 each).  Without state-box rows
   * they are the plain data Q and rho*q; with them (keep_state_box) they carry the barrier terms
   * of those rows exactly like Rt / rt do for the input box. */
  const int N = qp->N;
  double P[NX * NX] = {0}, p[NX];
  for (int i = 0; i < NX; i++) {
    P[i * NX + i] = Qt[N * NX + i];
    p[i] = qt[N * NX + i];
  }
  for (int k = N - 1; k >= 0; k--) {
    const double *A = qp->A + (size_t)k * NX * NX;
    const double *B = qp->B + (size_t)k * NX * NU;
    const double *b = qp->b + (size_t)k * NX;
    double PA[NX * NX], PB[NX * NU], w[NX];
    for (int i = 0; i < NX; i++) {
      for (int j = 0; j < NX; j++) {
        double s = 0;
        for (int l = 0; l < NX; l++) s += P[i * NX + l] * A[l * NX + j];
        PA[i * NX + j] = s;
      }
      for (int j = 0; j < NU; j++) {
        double s = 0;
        for (int l = 0; l < NX; l++) s += P[i * NX + l] * B[l * NU + j];
        PB[i * NU + j] = s;
      }
      double s = p[i];
      for (int l = 0; l < NX; l++) s += P[i * NX + l] * (rho * b[l]);
      w[i] = s;
    }
    double Quu[16], Qux[NU * NX], Qxx[NX * NX], gu[NU], gx[NX];
    for (int i = 0; i < NU; i++) {
      for (int j = 0; j < NU; j++) {
        double s = (i == j) ? Rt[k * NU + i] : 0.0;
        for (int l = 0; l < NX; l++) s += B[l * NU + i] * PB[l * NU + j];
        Quu[i * 4 + j] = s;
      }
      for (int j = 0; j < NX; j++) {
        double s = 0;
        for (int l = 0; l < NX; l++) s += B[l * NU + i] * PA[l * NX + j];
        Qux[i * NX + j] = s;
      }
      double s = rt[k * NU + i];
      for (int l = 0; l < NX; l++) s += B[l * NU + i] * w[l];
      gu[i] = s;
    }
    for (int i = 0; i < NX; i++) {
      for (int j = 0; j < NX; j++) {
        double s = (i == j) ? Qt[k * NX + i] : 0.0;
        for (int l = 0; l < NX; l++) s += A[l * NX + i] * PA[l * NX + j];
        Qxx[i * NX + j] = s;
      }
      double s = qt[k * NX + i];
      for (int l = 0; l < NX; l++) s += A[l * NX + i] * w[l];
      gx[i] = s;
    }
    double *Qi = ws->Qi + (size_t)k * 16;
    if (chol4_inv(Quu, Qi)) return 1;
    double *K = ws->K + (size_t)k * NU * NX;
    double *kff = ws->kff + (size_t)k * NU;
    for (int i = 0; i < NU; i++) {
      for (int j = 0; j < NX; j++) {
        double s = 0;
        for (int l = 0; l < NU; l++) s -= Qi[i * 4 + l] * Qux[l * NX + j];
        K[i * NX + j] = s;
      }
      double s = 0;
      for (int l = 0; l < NU; l++) s -= Qi[i * 4 + l] * gu[l];
      kff[i] = s;
    }
    for (int i = 0; i < NX; i++) {
      for (int j = 0; j < NX; j++) {
        double s = Qxx[i * NX + j];
        for (int l = 0; l < NU; l++) s += Qux[l * NX + i] * K[l * NX + j];
        P[i * NX + j] = s;
      }
      double s = gx[i];
      for (int l = 0; l < NU; l++) s += Qux[l * NX + i] * kff[l];
      p[i] = s;
    }
    /* symmetrise */
    for (int i = 0; i < NX; i++)
      for (int j = i + 1; j < NX; j++) {
        const double m = 0.5 * (P[i * NX + j] + P[j * NX + i]);
        P[i * NX + j] = P[j * NX + i] = m;
      }
  }
  memset(dx, 0, sizeof(double) * NX);
  for (int k = 0; k < N; k++) {
    const double *A = qp->A + (size_t)k * NX * NX;
    const double *B = qp->B + (size_t)k * NX * NU;
    const double *K = ws->K + (size_t)k * NU * NX;
    const double *x = dx + (size_t)k * NX;
    double *u = du + (size_t)k * NU;
    double *xn = dx + (size_t)(k + 1) * NX;
    for (int i = 0; i < NU; i++) {
      double s = ws->kff[k * NU + i];
      for (int j = 0; j < NX; j++) s += K[i * NX + j] * x[j];
      u[i] = s;
    }
    for (int i = 0; i < NX; i++) {
      double s = rho * qp->b[k * NX + i];
      for (int j = 0; j < NX; j++) s += A[i * NX + j] * x[j];
      for (int j = 0; j < NU; j++) s += B[i * NU + j] * u[j];
      xn[i] = s;
    }
  }
  return 0;
}

/* Re-solve with the same factorisation for a right-hand side that differs only in the
 * input gradient (b = 0, q = 0, r = dr): the corrector increment of Mehrotra's method. */
static void riccati_delta_solve(const orc_qp *qp, const double *dr, const double *dq,
                                const ric_ws *ws, double *dx, double *du, double *kffd) {
  const int N = qp->N;
  double p[NX] = {0};
  for (int k = N - 1; k >= 0; k--) {
    const double *A = qp->A + (size_t)k * NX * NX;
    const double *B = qp->B + (size_t)k * NX * NU;
    const double *K = ws->K + (size_t)k * NU * NX;
    const double *Qi = ws->Qi + (size_t)k * 16;
    double gu[NU], gx[NX];
    for (int i = 0; i < NU; i++) {
      double s = dr[k * NU + i];
      for (int l = 0; l < NX; l++) s += B[l * NU + i] * p[l];
      gu[i] = s;
    }
    for (int i = 0; i < NX; i++) {
      double s = dq ? dq[k * NX + i] : 0.0; /* state-box rows (keep_state_box) */
      for (int l = 0; l < NX; l++) s += A[l * NX + i] * p[l];
      gx[i] = s;
    }
    for (int i = 0; i < NU; i++) {
      double s = 0;
      for (int l = 0; l < NU; l++) s -= Qi[i * 4 + l] * gu[l];
      kffd[k * NU + i] = s;
    }
    for (int i = 0; i < NX; i++) {
      double s = gx[i];
      for (int l = 0; l < NU; l++) s += K[l * NX + i] * gu[l];
      p[i] = s;
    }
  }
  memset(dx, 0, sizeof(double) * NX);
  for (int k = 0; k < N; k++) {
    const double *A = qp->A + (size_t)k * NX * NX;
    const double *B = qp->B + (size_t)k * NX * NU;
    const double *K = ws->K + (size_t)k * NU * NX;
    const double *x = dx + (size_t)k * NX;
    double *u = du + (size_t)k * NU;
    double *xn = dx + (size_t)(k + 1) * NX;
    for (int i = 0; i < NU; i++) {
      double s = kffd[k * NU + i];
      for (int j = 0; j < NX; j++) s += K[i * NX + j] * x[j];
      u[i] = s;
    }
    for (int i = 0; i < NX; i++) {
      double s = 0;
      for (int j = 0; j < NX; j++) s += A[i * NX + j] * x[j];
      for (int j = 0; j < NU; j++) s += B[i * NU + j] * u[j];
      xn[i] = s;
    }
  }
}

/* ------------------------------------------------------- active-set shortcut */
/* Exact solve of the QP for a GIVEN active set (act[i]: 0 free, 1 on the lower, 2 on the upper
 * bound) in ONE Riccati factorisation: the active inputs are eliminated stage by stage
 * (du_i = bound, their row / column of Quu replaced by the identity, their coupling moved to the
 * gradients), the forward roll-out gives (dx, du), a costate sweep gives the multipliers
 *     s = R du + r + B' pi_{k+1} = lam_l - lam_u .
 * The point is THE solution of the QP (strictly convex) iff it is primal feasible on the free rows
 * and the multipliers of the active rows are non-negative; then 0 is returned, otherwise 1 and the
 * caller falls back to the interior point.  Used between the QPs of one SQP run once the active set
 * has settled (orc_solve): same solutions, one factorisation instead of several. */
int orc_qp_solve_active_set(const orc_qp *qp, const unsigned char *act, double *dx, double *du,
                            double *lam_l, double *lam_u) {
  const int N = qp->N;
  const int nb = N * NU;
  double *Ks = (double *)malloc(sizeof(double) * (size_t)N * NU * NX);
  double *kffs = (double *)malloc(sizeof(double) * (size_t)N * NU);
  int bad = 0;
  double P[NX * NX] = {0}, p[NX];
  for (int i = 0; i < NX; i++) {
    P[i * NX + i] = qp->Q[N * NX + i];
    p[i] = qp->q[N * NX + i];
  }
  for (int k = N - 1; k >= 0 && !bad; k--) {
    const double *A = qp->A + (size_t)k * NX * NX;
    const double *B = qp->B + (size_t)k * NX * NU;
    const double *b = qp->b + (size_t)k * NX;
    double PA[NX * NX], PB[NX * NU], w[NX];
    for (int i = 0; i < NX; i++) {
      for (int j = 0; j < NX; j++) {
        double s = 0;
        for (int l = 0; l < NX; l++) s += P[i * NX + l] * A[l * NX + j];
        PA[i * NX + j] = s;
      }
      for (int j = 0; j < NU; j++) {
        double s = 0;
        for (int l = 0; l < NX; l++) s += P[i * NX + l] * B[l * NU + j];
        PB[i * NU + j] = s;
      }
      double s = p[i];
      for (int l = 0; l < NX; l++) s += P[i * NX + l] * b[l];
      w[i] = s;
    }
    double Quu[16], Qux[NU * NX], Qxx[NX * NX], gu[NU], gx[NX];
    for (int i = 0; i < NU; i++) {
      for (int j = 0; j < NU; j++) {
        double s = (i == j) ? qp->R[k * NU + i] : 0.0;
        for (int l = 0; l < NX; l++) s += B[l * NU + i] * PB[l * NU + j];
        Quu[i * 4 + j] = s;
      }
      for (int j = 0; j < NX; j++) {
        double s = 0;
        for (int l = 0; l < NX; l++) s += B[l * NU + i] * PA[l * NX + j];
        Qux[i * NX + j] = s;
      }
      double s = qp->r[k * NU + i];
      for (int l = 0; l < NX; l++) s += B[l * NU + i] * w[l];
      gu[i] = s;
    }
    for (int i = 0; i < NX; i++) {
      for (int j = 0; j < NX; j++) {
        double s = (i == j) ? qp->Q[k * NX + i] : 0.0;
        for (int l = 0; l < NX; l++) s += A[l * NX + i] * PA[l * NX + j];
        Qxx[i * NX + j] = s;
      }
      double s = qp->q[k * NX + i];
      for (int l = 0; l < NX; l++) s += A[l * NX + i] * w[l];
      gx[i] = s;
    }
    /* eliminate the active inputs: du_i = d_i */
    for (int i = 0; i < NU; i++) {
      const int a = act[k * NU + i];
      if (!a) continue;
      const double d = (a == 1) ? qp->lb[k * NU + i] : qp->ub[k * NU + i];
      for (int j = 0; j < NU; j++)
        if (j != i && !act[k * NU + j]) gu[j] += Quu[j * 4 + i] * d;
      for (int j = 0; j < NX; j++) gx[j] += Qux[i * NX + j] * d;
    }
    for (int i = 0; i < NU; i++) {
      if (!act[k * NU + i]) continue;
      const int a = act[k * NU + i];
      const double d = (a == 1) ? qp->lb[k * NU + i] : qp->ub[k * NU + i];
      for (int j = 0; j < NU; j++) Quu[i * 4 + j] = Quu[j * 4 + i] = 0.0;
      Quu[i * 4 + i] = 1.0;
      for (int j = 0; j < NX; j++) Qux[i * NX + j] = 0.0;
      gu[i] = -d;
    }
    double Qi[16];
    if (chol4_inv(Quu, Qi)) {
      bad = 1;
      break;
    }
    double *K = Ks + (size_t)k * NU * NX;
    double *kff = kffs + (size_t)k * NU;
    for (int i = 0; i < NU; i++) {
      for (int j = 0; j < NX; j++) {
        double s = 0;
        for (int l = 0; l < NU; l++) s -= Qi[i * 4 + l] * Qux[l * NX + j];
        K[i * NX + j] = s;
      }
      double s = 0;
      for (int l = 0; l < NU; l++) s -= Qi[i * 4 + l] * gu[l];
      kff[i] = s;
    }
    for (int i = 0; i < NX; i++) {
      for (int j = 0; j < NX; j++) {
        double s = Qxx[i * NX + j];
        for (int l = 0; l < NU; l++) s += Qux[l * NX + i] * K[l * NX + j];
        P[i * NX + j] = s;
      }
      double s = gx[i];
      for (int l = 0; l < NU; l++) s += Qux[l * NX + i] * kff[l];
      p[i] = s;
    }
    for (int i = 0; i < NX; i++)
      for (int j = i + 1; j < NX; j++) {
        const double m = 0.5 * (P[i * NX + j] + P[j * NX + i]);
        P[i * NX + j] = P[j * NX + i] = m;
      }
  }
  if (!bad) {
    memset(dx, 0, sizeof(double) * NX);
    for (int k = 0; k < N; k++) {
      const double *A = qp->A + (size_t)k * NX * NX;
      const double *B = qp->B + (size_t)k * NX * NU;
      const double *K = Ks + (size_t)k * NU * NX;
      const double *x = dx + (size_t)k * NX;
      double *u = du + (size_t)k * NU;
      double *xn = dx + (size_t)(k + 1) * NX;
      for (int i = 0; i < NU; i++) {
        double s = kffs[k * NU + i];
        for (int j = 0; j < NX; j++) s += K[i * NX + j] * x[j];
        const int a = act[k * NU + i];
        u[i] = a ? ((a == 1) ? qp->lb[k * NU + i] : qp->ub[k * NU + i]) : s; /* exactly on the bound */
      }
      for (int i = 0; i < NX; i++) {
        double s = qp->b[k * NX + i];
        for (int j = 0; j < NX; j++) s += A[i * NX + j] * x[j];
        for (int j = 0; j < NU; j++) s += B[i * NU + j] * u[j];
        xn[i] = s;
      }
    }
    /* costate sweep: multipliers of the active rows, primal feasibility of the free ones */
    double pi[NX];
    for (int i = 0; i < NX; i++) pi[i] = qp->Q[N * NX + i] * dx[N * NX + i] + qp->q[N * NX + i];
    for (int k = N - 1; k >= 0; k--) {
      const double *A = qp->A + (size_t)k * NX * NX;
      const double *B = qp->B + (size_t)k * NX * NU;
      for (int i = 0; i < NU; i++) {
        double s = qp->R[k * NU + i] * du[k * NU + i] + qp->r[k * NU + i];
        for (int l = 0; l < NX; l++) s += B[l * NU + i] * pi[l];
        const int a = act[k * NU + i];
        lam_l[k * NU + i] = (a == 1) ? s : 0.0;
        lam_u[k * NU + i] = (a == 2) ? -s : 0.0;
        if (a == 1 && !(s >= 0.0)) bad = 1;
        if (a == 2 && !(-s >= 0.0)) bad = 1;
        if (!a && !(du[k * NU + i] >= qp->lb[k * NU + i] && du[k * NU + i] <= qp->ub[k * NU + i]))
          bad = 1;
      }
      double pn[NX];
      for (int i = 0; i < NX; i++) {
        double s = qp->Q[k * NX + i] * dx[k * NX + i] + qp->q[k * NX + i];
        for (int l = 0; l < NX; l++) s += A[l * NX + i] * pi[l];
        pn[i] = s;
      }
      memcpy(pi, pn, sizeof(pn));
    }
    for (int i = 0; i < nb; i++)
      if (!(du[i] == du[i])) bad = 1;
  }
  free(Ks);
  free(kffs);
  return bad;
}

/* ------------------------------------------------------- interior point */
/* Weight of Mehrotra's second-order term.  With a short affine step the full term
 * dt_aff*dlam_aff over-corrects and the iteration can fall into a 2-cycle (seen on warm
 * re-solves with many active bounds); scaling it by min(1, 2 alpha_aff) removes the cycle at
 * no measurable cost in iterations. */
static double corr_weight(double a) { return fmin(1.0, 2.0 * a); }

/* Mehrotra predictor-corrector on
 *     min 1/2 z'Hz + g'z   s.t. dynamics,  lb <= du <= ub
 * with slacks t_l = du - lb, t_u = ub - du and multipliers lam_l, lam_u >= 0.
 *
 * One step length is used for primal and dual variables, so every *linear* residual
 * (dynamics, box, stationarity) of the iterate equals rho = prod(1 - alpha) times its value
 * at the starting point (dz = 0, pi = 0), which is plain problem data:
 *     dynamics b,  x-stationarity q,  u-stationarity ru0 = r - lam_l0 + lam_u0,
 *     box rbl0 = -lb - t_l0,  rbu0 = ub - t_u0.
 * The Newton step is therefore computed from right-hand sides that shrink with the
 * iterate (no cancellation, no costates needed):
 *     Rt = R + lam_l/t_l + lam_u/t_u
 *     rt = rho ru0 + d_l rho rbl0 + lam_l - d_u rho rbu0 - lam_u - s_l/t_l + s_u/t_u ,
 *     s  = sigma*mu - (dt_aff * dlam_aff)          (0 for the predictor)
 *     Riccati solve with (rho b, rho q, rt)  ->  (Dx, Du)
 * Predictor and corrector share the factorisation and differ only in rt, so the
 * corrector is a vector-only re-solve added to the predictor direction. */
int orc_qp_solve(const orc_qp *qp, int iter_max, double tol, double *dx, double *du, double *lam_l,
                 double *lam_u, int *iters) {
  return orc_qp_solve_warm(qp, iter_max, tol, dx, du, lam_l, lam_u, NULL, NULL, 0.0, NULL, iters);
}

/* The same solver with an optional warm start of the interior-point iterate (used between the QPs
 * of one SQP run, see orc_solve): with warm_eps > 0 the loop starts from the previous QP's slacks
 * (t_l_io, t_u_io) and multipliers (lam_l, lam_u), pushed away from the boundary,
 *     t = max(t_prev, eps),  lam = max(lam_prev, eps),
 * instead of t = max(slack, 0.1), lam = 0.1 / t, and phase 0 is skipped (see below).  On return t_l_io / t_u_io (when given) hold the
 * slacks of the final iterate and *ipm_converged tells whether the interior-point loop ran and
 * met its tolerance (only then is the iterate a valid warm start for the next QP). */
int orc_qp_solve_warm(const orc_qp *qp, int iter_max, double tol, double *dx, double *du,
                      double *lam_l, double *lam_u, double *t_l_io, double *t_u_io, double warm_eps,
                      int *ipm_converged, int *iters) {
  const int N = qp->N;
  const int nb = N * NU;
  const int nxv = (N + 1) * NX;
  /* state-box rows (keep_state_box): rows k*13+i for k = 1..N-1; the others are masked out */
  const int xbox = qp->xbox;
  const int xlo = NX, xhi = N * NX; /* [xlo, xhi) = stages 1..N-1 */
  const int nbx = (xbox && xhi > xlo) ? (xhi - xlo) : 0;
  ric_ws ws;
  ws.N = N;
  ws.K = (double *)malloc(sizeof(double) * (size_t)N * NU * NX);
  ws.Qi = (double *)malloc(sizeof(double) * (size_t)N * 16);
  ws.kff = (double *)malloc(sizeof(double) * (size_t)N * NU);
  double *buf = (double *)calloc((size_t)nb * 15 + (size_t)nxv * 16, sizeof(double));
  double *t_l = buf, *t_u = t_l + nb, *Rt = t_u + nb, *rt = Rt + nb;
  double *dtl = rt + nb, *dtu = dtl + nb, *dll = dtu + nb, *dlu = dll + nb;
  double *dr = dlu + nb, *ua = dr + nb, *uc = ua + nb, *kffd = uc + nb;
  double *rbl0 = kffd + nb, *rbu0 = rbl0 + nb, *ru0 = rbu0 + nb;
  double *xa = ru0 + nb, *xc = xa + nxv;
  /* state rows: Hessian diagonal / gradient handed to the Riccati solve, slacks, multipliers,
   * their directions, corrector gradient, initial residuals */
  double *Qt = xc + nxv, *qt = Qt + nxv, *dq = qt + nxv;
  double *s_l = dq + nxv, *s_u = s_l + nxv, *m_l = s_u + nxv, *m_u = m_l + nxv;
  double *dsl = m_u + nxv, *dsu = dsl + nxv, *dml = dsu + nxv, *dmu = dml + nxv;
  double *rxl0 = dmu + nxv, *rxu0 = rxl0 + nxv, *rx0 = rxu0 + nxv;
  int status = ORC_MAXITER;
  int it = 0;
  if (ipm_converged) *ipm_converged = 0;

  /* Phase 0: equality-constrained QP (box ignored).  If its solution respects the box it
   * is the QP solution (strict convexity) and no interior-point iteration is needed. */
  double res0 = 0.0;
  for (int i = 0; i < nb; i++) res0 = fmax(res0, fabs(qp->r[i]) + fabs(qp->lb[i]) + fabs(qp->ub[i]));
  for (int i = 0; i < N * NX; i++) res0 = fmax(res0, fabs(qp->b[i]));
  for (int i = 0; i < nxv; i++) res0 = fmax(res0, fabs(qp->q[i]));
  if (!(res0 == res0) || isinf(res0)) { /* NaN/Inf in the QP data */
    status = ORC_QP_FAILURE;
    goto done;
  }
  /* A warm-started QP skips phase 0 (its bounds were active a moment ago): the loop starts at
   * dz = 0, where the dynamics and x-stationarity residuals are the plain data b and q; they
   * shrink with rho like the others and enter every Riccati solve as (rho b, rho q). */
  const int skip0 = warm_eps > 0.0;
  const double rq = skip0 ? 1.0 : 0.0;
  if (skip0) {
    for (int i = 0; i < nb; i++) du[i] = 0.0;
    for (int i = 0; i < nxv; i++) dx[i] = 0.0;
  } else {
    if (riccati_factor_solve(qp, 1.0, qp->Q, qp->q, qp->R, qp->r, &ws, dx, du)) {
      status = ORC_QP_FAILURE;
      goto done;
    }
    it = 1; /* iterations are counted in Riccati factorisations */
  }
  if (!skip0) {
    int feasible = 1;
    for (int i = 0; i < nb; i++)
      if (!(du[i] >= qp->lb[i] && du[i] <= qp->ub[i])) feasible = 0;
    for (int i = xlo; i < xlo + nbx; i++)
      if (!(dx[i] >= qp->lbx[i] && dx[i] <= qp->ubx[i])) feasible = 0;
    if (feasible) {
      for (int i = 0; i < nb; i++) lam_l[i] = lam_u[i] = 0.0;
      status = ORC_SUCCESS;
      goto done;
    }
  }
  /* Interior-point start: the unconstrained solution (dynamics and x-stationarity hold
   * exactly there), slacks clipped away from zero, lam = mu0 / t. */
  const double thr0 = 0.1, mu0 = 0.1;
  res0 = 0.0;
  for (int i = 0; i < nb; i++) {
    if (warm_eps > 0.0) {
      t_l[i] = fmax(t_l_io[i], warm_eps);
      t_u[i] = fmax(t_u_io[i], warm_eps);
      lam_l[i] = fmax(lam_l[i], warm_eps);
      lam_u[i] = fmax(lam_u[i], warm_eps);
    } else {
      t_l[i] = fmax(du[i] - qp->lb[i], thr0);
      t_u[i] = fmax(qp->ub[i] - du[i], thr0);
      lam_l[i] = mu0 / t_l[i];
      lam_u[i] = mu0 / t_u[i];
    }
    rbl0[i] = du[i] - qp->lb[i] - t_l[i];
    rbu0[i] = qp->ub[i] - du[i] - t_u[i];
    ru0[i] = -lam_l[i] + lam_u[i] + rq * qp->r[i];
    res0 = fmax(res0, fmax(fabs(rbl0[i]), fabs(rbu0[i])));
    res0 = fmax(res0, fabs(ru0[i]));
  }
  for (int i = 0; i < nxv; i++) rx0[i] = rq * qp->q[i];
  for (int i = xlo; i < xlo + nbx; i++) { /* same start rule for the state rows (never warm) */
    s_l[i] = fmax(dx[i] - qp->lbx[i], thr0);
    s_u[i] = fmax(qp->ubx[i] - dx[i], thr0);
    m_l[i] = mu0 / s_l[i];
    m_u[i] = mu0 / s_u[i];
    rxl0[i] = dx[i] - qp->lbx[i] - s_l[i];
    rxu0[i] = qp->ubx[i] - dx[i] - s_u[i];
    rx0[i] += -m_l[i] + m_u[i];
    res0 = fmax(res0, fmax(fabs(rxl0[i]), fabs(rxu0[i])));
    res0 = fmax(res0, fabs(rx0[i]));
  }
  double rho = 1.0;
  if (skip0) {
    for (int i = 0; i < N * NX; i++) res0 = fmax(res0, fabs(qp->b[i]));
    for (int i = 0; i < nxv; i++) res0 = fmax(res0, fabs(qp->q[i]));
  }
  const double inv_rows = 1.0 / (double)(2 * nb + 2 * nbx);

  for (; it < iter_max; it++) {
    double mu = 0.0;
    for (int i = 0; i < nb; i++) mu += t_l[i] * lam_l[i] + t_u[i] * lam_u[i];
    for (int i = xlo; i < xlo + nbx; i++) mu += s_l[i] * m_l[i] + s_u[i] * m_u[i];
    mu *= inv_rows;
    if (!(mu == mu)) {
      status = ORC_QP_FAILURE;
      break;
    }
    if (mu <= tol && rho * res0 <= tol) {
      status = ORC_SUCCESS;
      break;
    }
    /* predictor */
    for (int i = 0; i < nb; i++) {
      const double dl = lam_l[i] / t_l[i], dw = lam_u[i] / t_u[i];
      Rt[i] = qp->R[i] + dl + dw;
      rt[i] = rho * ru0[i] + dl * (rho * rbl0[i]) + lam_l[i] - dw * (rho * rbu0[i]) - lam_u[i];
    }
    for (int i = 0; i < nxv; i++) {
      Qt[i] = qp->Q[i];
      qt[i] = rho * rx0[i];
    }
    for (int i = xlo; i < xlo + nbx; i++) {
      const double dl = m_l[i] / s_l[i], dw = m_u[i] / s_u[i];
      Qt[i] += dl + dw;
      qt[i] += dl * (rho * rxl0[i]) + m_l[i] - dw * (rho * rxu0[i]) - m_u[i];
    }
    if (riccati_factor_solve(qp, rq * rho, Qt, qt, Rt, rt, &ws, xa, ua)) {
      status = ORC_QP_FAILURE;
      break;
    }
    double alpha_aff = 1.0;
    for (int i = 0; i < nb; i++) {
      dtl[i] = ua[i] + rho * rbl0[i];
      dtu[i] = -ua[i] + rho * rbu0[i];
      dll[i] = -(lam_l[i] / t_l[i]) * dtl[i] - lam_l[i];
      dlu[i] = -(lam_u[i] / t_u[i]) * dtu[i] - lam_u[i];
      if (dtl[i] < 0) alpha_aff = fmin(alpha_aff, -t_l[i] / dtl[i]);
      if (dtu[i] < 0) alpha_aff = fmin(alpha_aff, -t_u[i] / dtu[i]);
      if (dll[i] < 0) alpha_aff = fmin(alpha_aff, -lam_l[i] / dll[i]);
      if (dlu[i] < 0) alpha_aff = fmin(alpha_aff, -lam_u[i] / dlu[i]);
    }
    for (int i = xlo; i < xlo + nbx; i++) {
      dsl[i] = xa[i] + rho * rxl0[i];
      dsu[i] = -xa[i] + rho * rxu0[i];
      dml[i] = -(m_l[i] / s_l[i]) * dsl[i] - m_l[i];
      dmu[i] = -(m_u[i] / s_u[i]) * dsu[i] - m_u[i];
      if (dsl[i] < 0) alpha_aff = fmin(alpha_aff, -s_l[i] / dsl[i]);
      if (dsu[i] < 0) alpha_aff = fmin(alpha_aff, -s_u[i] / dsu[i]);
      if (dml[i] < 0) alpha_aff = fmin(alpha_aff, -m_l[i] / dml[i]);
      if (dmu[i] < 0) alpha_aff = fmin(alpha_aff, -m_u[i] / dmu[i]);
    }
    double mu_aff = 0.0;
    for (int i = 0; i < nb; i++)
      mu_aff += (t_l[i] + alpha_aff * dtl[i]) * (lam_l[i] + alpha_aff * dll[i]) +
                (t_u[i] + alpha_aff * dtu[i]) * (lam_u[i] + alpha_aff * dlu[i]);
    for (int i = xlo; i < xlo + nbx; i++)
      mu_aff += (s_l[i] + alpha_aff * dsl[i]) * (m_l[i] + alpha_aff * dml[i]) +
                (s_u[i] + alpha_aff * dsu[i]) * (m_u[i] + alpha_aff * dmu[i]);
    mu_aff *= inv_rows;
    double sigma = mu_aff / mu;
    sigma = sigma * sigma * sigma;
    const double CORRW = corr_weight(alpha_aff);
    /* corrector increment: only the box-row gradients change */
    for (int i = 0; i < nb; i++) {
      const double sl = sigma * mu - CORRW * dtl[i] * dll[i];
      const double su = sigma * mu - CORRW * dtu[i] * dlu[i];
      dr[i] = -sl / t_l[i] + su / t_u[i];
    }
    for (int i = xlo; i < xlo + nbx; i++) {
      const double sl = sigma * mu - CORRW * dsl[i] * dml[i];
      const double su = sigma * mu - CORRW * dsu[i] * dmu[i];
      dq[i] = -sl / s_l[i] + su / s_u[i];
    }
    riccati_delta_solve(qp, dr, nbx ? dq : NULL, &ws, xc, uc, kffd);
    double alpha = 1e300; /* longest step keeping t, lam >= 0 (not capped at 1 yet) */
    for (int i = 0; i < nb; i++) {
      const double un = ua[i] + uc[i];
      const double sl = sigma * mu - CORRW * dtl[i] * dll[i];
      const double su = sigma * mu - CORRW * dtu[i] * dlu[i];
      dtl[i] = un + rho * rbl0[i];
      dtu[i] = -un + rho * rbu0[i];
      dll[i] = -(lam_l[i] / t_l[i]) * dtl[i] + sl / t_l[i] - lam_l[i];
      dlu[i] = -(lam_u[i] / t_u[i]) * dtu[i] + su / t_u[i] - lam_u[i];
      if (dtl[i] < 0) alpha = fmin(alpha, -t_l[i] / dtl[i]);
      if (dtu[i] < 0) alpha = fmin(alpha, -t_u[i] / dtu[i]);
      if (dll[i] < 0) alpha = fmin(alpha, -lam_l[i] / dll[i]);
      if (dlu[i] < 0) alpha = fmin(alpha, -lam_u[i] / dlu[i]);
    }
    for (int i = xlo; i < xlo + nbx; i++) {
      const double xn = xa[i] + xc[i];
      const double sl = sigma * mu - CORRW * dsl[i] * dml[i];
      const double su = sigma * mu - CORRW * dsu[i] * dmu[i];
      dsl[i] = xn + rho * rxl0[i];
      dsu[i] = -xn + rho * rxu0[i];
      dml[i] = -(m_l[i] / s_l[i]) * dsl[i] + sl / s_l[i] - m_l[i];
      dmu[i] = -(m_u[i] / s_u[i]) * dsu[i] + su / s_u[i] - m_u[i];
      if (dsl[i] < 0) alpha = fmin(alpha, -s_l[i] / dsl[i]);
      if (dsu[i] < 0) alpha = fmin(alpha, -s_u[i] / dsu[i]);
      if (dml[i] < 0) alpha = fmin(alpha, -m_l[i] / dml[i]);
      if (dmu[i] < 0) alpha = fmin(alpha, -m_u[i] / dmu[i]);
    }
#ifdef ORC_TRACE
    fprintf(stderr, "it %d mu %.3e mu_aff %.3e sig %.2e a_aff %.2e amax %.3e rho %.2e\n", it, mu,
            mu_aff, sigma, alpha_aff, alpha, rho);
#endif
    if (!(alpha == alpha)) {
      status = ORC_QP_FAILURE;
      break;
    }
    /* fraction to the boundary, then cap at the full Newton step */
    const double tau = fmax(0.995, 1.0 - mu);
    alpha = fmin(1.0, tau * alpha);
    if (alpha < 1e-12) {
      status = ORC_MINSTEP;
      break;
    }
    for (int i = 0; i < nb; i++) {
      du[i] += alpha * (ua[i] + uc[i]);
      t_l[i] += alpha * dtl[i];
      t_u[i] += alpha * dtu[i];
      lam_l[i] += alpha * dll[i];
      lam_u[i] += alpha * dlu[i];
    }
    for (int i = xlo; i < xlo + nbx; i++) {
      s_l[i] += alpha * dsl[i];
      s_u[i] += alpha * dsu[i];
      m_l[i] += alpha * dml[i];
      m_u[i] += alpha * dmu[i];
    }
    for (int i = 0; i < nxv; i++) dx[i] += alpha * (xa[i] + xc[i]);
    rho *= (1.0 - alpha);
  }
  if (t_l_io && t_u_io) {
    memcpy(t_l_io, t_l, sizeof(double) * (size_t)nb);
    memcpy(t_u_io, t_u, sizeof(double) * (size_t)nb);
  }
  if (ipm_converged) *ipm_converged = (status == ORC_SUCCESS);
done:
  if (iters) *iters = it;
  free(buf);
  free(ws.K);
  free(ws.Qi);
  free(ws.kff);
  return status;
}
