/*
 * nmpc_oracle_solve.c -- CPU ORACLE (test infrastructure, NOT the product).  See nmpc_oracle.h.
 * PARITY UNPINNED.  The per-call semantics of PositionNMPC.solve (controller.py:385-463)
 * and of the node's command mapping (nodes/mpc_controller_node:124-131,152-164).
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "nmpc_oracle.h"

#define NX ORC_NX
#define NU ORC_NU
#define NZ ORC_NZ

/* diagnostics (not thread-safe, per process): active-set shortcut attempts */
long orc_as_accepted = 0, orc_as_rejected = 0;

static int has_nan(const double *v, int n) {
  for (int i = 0; i < n; i++)
    if (!(v[i] == v[i]) || isinf(v[i])) return 1;
  return 0;
}

int orc_solve(const orc_config *c, const double *x0, const double *yref, const double *yref_e,
              double *xtraj, double *utraj, unsigned char *valid, double *u0, int *qp_iters,
              int *sqp_iters) {
  const int N = c->N;
  /* working iterate: the warm start is only committed on success (controller.py:452-461) */
  double *xb = (double *)malloc(sizeof(double) * (size_t)(N + 1) * NX);
  double *ub = (double *)malloc(sizeof(double) * (size_t)N * NU);
  double *dx = (double *)malloc(sizeof(double) * (size_t)(N + 1) * NX);
  double *du = (double *)malloc(sizeof(double) * (size_t)N * NU);
  double *ll = (double *)malloc(sizeof(double) * (size_t)N * NU);
  double *lu = (double *)malloc(sizeof(double) * (size_t)N * NU);
  double *tl = (double *)calloc((size_t)N * NU, sizeof(double));
  double *tu = (double *)calloc((size_t)N * NU, sizeof(double));
  double *ll2 = (double *)malloc(sizeof(double) * (size_t)N * NU);
  double *lu2 = (double *)malloc(sizeof(double) * (size_t)N * NU);
  unsigned char *act = (unsigned char *)malloc((size_t)N * NU);
  int n_as_ok = 0, n_as_fail = 0;
  orc_qp *qp = orc_qp_alloc(N);

  if (*valid) { /* controller.py:419-424 : previous solution, NOT shifted; stage 0 x <- x0 (:416) */
    memcpy(xb, xtraj, sizeof(double) * (size_t)(N + 1) * NX);
    memcpy(ub, utraj, sizeof(double) * (size_t)N * NU);
    memcpy(xb, x0, sizeof(double) * NX);
  } else { /* controller.py:425-431 : every x <- x0, every u <- 0 */
    for (int k = 0; k <= N; k++) memcpy(xb + (size_t)k * NX, x0, sizeof(double) * NX);
    memset(ub, 0, sizeof(double) * (size_t)N * NU);
  }

  int status = ORC_SUCCESS;
  int tot_qp = 0, nsqp = 0, converged = 0;
  int have_ipm = 0;       /* previous QP of this solve left a converged interior-point iterate */
  double prev_step = 0.0; /* inf-norm of the previous SQP step */
  const int max_sqp = c->max_sqp_iter < 1 ? 1 : c->max_sqp_iter;
  while (nsqp < max_sqp) {
    orc_build_qp(c, xb, ub, yref, yref_e, qp);
    int qi = 0;
    double weps = 0.0;
    if (c->ipm_warm && !c->keep_state_box && have_ipm && prev_step <= ORC_WARM_STEP_MAX)
      weps = fmin(fmax(ORC_WARM_KAPPA * prev_step, ORC_WARM_EPS_MIN), ORC_WARM_EPS_MAX);
    int qs = -1;
    /* Active-set method first (c->as_shortcut): primal-dual active-set iterations on the box rows.
     * The starting set is the active set of the previous QP of this solve when it left a valid
     * (t, lam), else the empty set (the first solve is then the equality-constrained "phase 0").
     * Every iteration is ONE Riccati factorisation with the inputs of the set eliminated + roll-out
     * (+ a costate sweep for the multipliers when the set is not empty).  A point that is primal
     * feasible on the free rows with non-negative multipliers on the fixed ones IS the solution of
     * the strictly convex QP and is accepted; otherwise violated free rows join the set, fixed rows
     * with a negative multiplier leave it, and the solve is repeated -- at most ORC_AS_MAXIT times,
     * then the interior point takes over with everything as the previous QP left it. */
    if (c->as_shortcut && !c->keep_state_box) {
      int nact = 0;
      for (int i = 0; i < N * NU; i++) {
        act[i] = !have_ipm ? 0 : ((ll[i] > tl[i]) ? 1 : ((lu[i] > tu[i]) ? 2 : 0));
        nact += act[i] != 0;
      }
      for (int rep = 0; rep < ORC_AS_MAXIT && rep < c->qp_iter_max; rep++) { /* iter_max caps every QP method */
        tot_qp++;
        if (orc_qp_solve_active_set(qp, act, dx, du, ll2, lu2) == 0) {
          qs = ORC_SUCCESS;
          memcpy(ll, ll2, sizeof(double) * (size_t)N * NU);
          memcpy(lu, lu2, sizeof(double) * (size_t)N * NU);
          for (int i = 0; i < N * NU; i++) { /* slacks of the accepted point (0 on active rows) */
            tl[i] = (act[i] == 1) ? 0.0 : du[i] - qp->lb[i];
            tu[i] = (act[i] == 2) ? 0.0 : qp->ub[i] - du[i];
          }
          have_ipm = nact > 0; /* a valid (t, lam) with a non-empty active set for the next QP */
          n_as_ok++;
          break;
        }
        n_as_fail++;
        int changed = 0;
        for (int i = 0; i < N * NU; i++) {
          const int a0 = act[i];
          if (a0 == 0 && du[i] < qp->lb[i]) act[i] = 1;
          else if (a0 == 0 && du[i] > qp->ub[i]) act[i] = 2;
          else if (a0 == 1 && !(ll2[i] >= 0.0)) act[i] = 0;
          else if (a0 == 2 && !(lu2[i] >= 0.0)) act[i] = 0;
          changed += act[i] != a0;
        }
        if (!changed) break; /* NaN data or a broken factorisation: let the interior point report it */
        nact = 0;
        for (int i = 0; i < N * NU; i++) nact += act[i] != 0;
      }
    }
    if (qs < 0)
      qs = orc_qp_solve_warm(qp, c->qp_iter_max, c->ipm_tol, dx, du, ll, lu, tl, tu, weps, &have_ipm,
                             &qi);
    tot_qp += qi;
    nsqp++;
    /* UPSTREAM acados SQP_RTI: a QP that only hit its iteration cap is tolerated. */
    if (qs != ORC_SUCCESS && qs != ORC_MAXITER) {
      status = ORC_QP_FAILURE;
      break;
    }
    double step = 0.0;
    for (int i = 0; i < (N + 1) * NX; i++) {
      xb[i] += dx[i]; /* full step (FIXED_STEP, alpha = 1) */
      step = fmax(step, fabs(dx[i]));
    }
    for (int i = 0; i < N * NU; i++) {
      ub[i] += du[i];
      step = fmax(step, fabs(du[i]));
    }
    if (has_nan(xb, (N + 1) * NX) || has_nan(ub, N * NU)) {
      status = c->nan_status;
      break;
    }
    prev_step = step;
    if (step <= c->sqp_tol) {
      converged = 1;
      break;
    }
  }
  /* extension (max_sqp_iter > 1 only): cap reached without convergence -> ACADOS_MAXITER */
  if (status == ORC_SUCCESS && max_sqp > 1 && !converged) status = ORC_MAXITER;
  if (qp_iters) *qp_iters = tot_qp;
  if (sqp_iters) *sqp_iters = nsqp;

  if (status != ORC_SUCCESS) { /* controller.py:448-450 */
    *valid = 0;
    for (int i = 0; i < NU; i++) u0[i] = 0.0;
  } else { /* controller.py:452-461 */
    memcpy(u0, ub, sizeof(double) * NU);
    memcpy(xtraj, xb, sizeof(double) * (size_t)(N + 1) * NX);
    memcpy(utraj, ub, sizeof(double) * (size_t)N * NU);
    *valid = 1;
  }
  orc_qp_free(qp);
  free(xb);
  free(ub);
  free(dx);
  free(du);
  free(ll);
  free(lu);
  free(tl);
  free(tu);
  free(ll2);
  free(lu2);
  free(act);
  orc_as_accepted += n_as_ok;
  orc_as_rejected += n_as_fail;
  return status;
}

static double clipd(double v, double lo, double hi) { return v < lo ? lo : (v > hi ? hi : v); }

/* nodes/mpc_controller_node:124-129 (status != 0 -> republish _last_command) and :152-164 */
void orc_command(const orc_config *c, const double *u0, int status, double *last_cmd,
                 double *thrust, double *speed) {
  const double kf = fmax(c->kf, 1e-9);
  for (int i = 0; i < NU; i++) {
    const double src = (status != 0) ? last_cmd[i] : u0[i];
    const double cl = clipd(src, c->lbu[i], c->ubu[i]);
    double s2 = clipd(cl / kf, 0.0, c->w_max * c->w_max);
    double s = sqrt(s2);
    s = clipd(s, c->w_min, c->w_max);
    thrust[i] = cl;
    speed[i] = s;
    last_cmd[i] = cl;
  }
}

void orc_solve_batch(const orc_config *c, int batch, const double *x0, const double *yref,
                     const double *yref_e, double *xtraj, double *utraj, unsigned char *valid,
                     double *last_cmd, double *u0, int *status, int *qp_iters, double *thrust,
                     double *speed) {
  const int N = c->N;
  for (int b = 0; b < batch; b++) {
    int qi = 0, si = 0;
    status[b] = orc_solve(c, x0 + (size_t)b * NX, yref + (size_t)b * N * NZ,
                          yref_e + (size_t)b * NX, xtraj + (size_t)b * (N + 1) * NX,
                          utraj + (size_t)b * N * NU, valid + b, u0 + (size_t)b * NU, &qi, &si);
    if (qp_iters) qp_iters[b] = qi;
    if (thrust && speed && last_cmd)
      orc_command(c, u0 + (size_t)b * NU, status[b], last_cmd + (size_t)b * NU,
                  thrust + (size_t)b * NU, speed + (size_t)b * NU);
  }
}
