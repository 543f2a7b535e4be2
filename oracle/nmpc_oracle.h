/*
 * nmpc_oracle.h -- CPU ORACLE (test infrastructure, NOT the product).
 *
 * Plain-C restatement of the NMPC solve that the reference obtains from acados
 * (`AcadosOcpSolver.set/solve/get`) in
 *   /root/reference/src/rotors_mpc_controller/controller.py
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may call this code.  The product path (rotors_mpc_controller_b200) never does.
 *
 * PARITY UNPINNED: the reference ships no tests, golden vectors or known-answer
 * fixtures for this path, and acados / HPIPM / BLASFEO / CasADi are neither vendored
 * under /root/reference nor installed here (no version is pinned anywhere in the
 * reference).  This oracle therefore restates the *published* acados formulation for
 * the options that controller.py sets (SQP_RTI default, ERK ns=2 x steps=2,
 * LINEAR_LS cost scaled by dt, Levenberg-Marquardt, box on u) and is pinned by
 * invariants instead (complex-step Jacobians, KKT residuals, an independent dense QP
 * solver, fixed points, closed-loop convergence) -- see tests/.
 */
#ifndef NMPC_ORACLE_H
#define NMPC_ORACLE_H

#ifdef __cplusplus
extern "C" {
#endif

#define ORC_NX 13
#define ORC_NU 4
#define ORC_NZ 17 /* [u;x] per stage */

/* acados-style status codes (controller.py:447-450 only tests `status != 0`). */
#define ORC_SUCCESS 0
#define ORC_NAN_DETECTED 1
#define ORC_MAXITER 2
#define ORC_MINSTEP 3
#define ORC_QP_FAILURE 4

/* All solver constants: the reference's ControllerParams (controller.py:24-49) plus the
 * acados options of controller.py:179-190 made explicit. */
typedef struct {
  int N;              /* horizon_steps                       controller.py:121 */
  double dt;          /* dt                                  controller.py:122 */
  double W[17];       /* stage weights [p,v,q,w,u]           controller.py:234-240 */
  double We[13];      /* terminal weights                    controller.py:241 */
  double lm;          /* levenberg_marquardt                 controller.py:190 */
  int qp_iter_max;    /* qp_solver_iter_max                  controller.py:185 */
  double mass;        /* controller.py:73 */
  double J[3];        /* inertia diagonal                    controller.py:81-83 */
  double gravity;     /* controller.py:74 */
  double rotor_x[4];  /* controller.py:100 */
  double rotor_y[4];  /* controller.py:101 */
  double rotor_z[4];  /* spin*km                             controller.py:102-103 */
  double lbu[4];      /* controller.py:105-110 */
  double ubu[4];
  double kf;          /* rotor_force_constant */
  double w_min;       /* motor_min_speed */
  double w_max;       /* motor_max_speed */
  int integ_ns;       /* sim_method_num_stages (1,2,4)       controller.py:187 */
  int integ_steps;    /* sim_method_num_steps                controller.py:188 */
  int cost_scale_dt;  /* acados cost scaling by time step (default 1) */
  int lm_scaled_by_dt;/* LM term of stages 0..N-1 scaled by dt (default 1; terminal unscaled) */
  int max_sqp_iter;   /* 1 = SQP_RTI (reference), >1 = SQP extension */
  double sqp_tol;     /* inf-norm step tolerance for max_sqp_iter>1 */
  double ipm_tol;     /* complementarity / residual tolerance of the QP IPM */
  int nan_status;     /* status reported when NaNs appear (4 = QP failure, or 1) */
  int ipm_warm;       /* 1: QPs after the first of an SQP run (max_sqp_iter > 1) warm-start the
                         interior point from the previous QP's (t, lam), see ORC_WARM_* */
  int as_shortcut;    /* 1: every QP is first attacked by primal-dual active-set iterations
                         (orc_qp_solve_active_set, one factorisation each, see orc_solve); the interior
                         point is the fallback.  0: interior point only */
  int keep_state_box; /* 1: the +-state_box rows on all 13 states at stages 1..N-1
                         (controller.py:257-261) are real inequality rows of the QP, as in the
                         reference's HPIPM problem.  ORACLE-ONLY sensitivity flag (default 0): the
                         rows are never active, they only perturb the interior-point iterates --
                         scripts/flag_sensitivity.py measures by how much.  Disables ipm_warm. */
  double state_box;   /* 1e6 (controller.py:258) */
} orc_config;

/* Warm start between the QPs of one SQP run: used when the previous QP's interior-point loop
 * converged and the previous SQP step was <= ORC_WARM_STEP_MAX (the active set is settling);
 * the iterate is clipped to eps = clamp(ORC_WARM_KAPPA * step, ORC_WARM_EPS_MIN, ORC_WARM_EPS_MAX). */
#ifndef ORC_WARM_KAPPA
#define ORC_WARM_KAPPA 0.1
#endif
#ifndef ORC_WARM_STEP_MAX
#define ORC_WARM_STEP_MAX 0.3
#endif
#ifndef ORC_WARM_EPS_MIN
#define ORC_WARM_EPS_MIN 1e-8
#endif
#ifndef ORC_WARM_EPS_MAX
#define ORC_WARM_EPS_MAX 0.1
#endif
/* active-set iterations tried per QP before the interior point takes over */
#ifndef ORC_AS_MAXIT
#define ORC_AS_MAXIT 32
#endif

/* continuous model, controller.py:267-355 */
void orc_f(const orc_config *c, const double *x, const double *u, double *f);
/* dense analytic Jacobians: Fx row-major 13x13, Fu row-major 13x4 */
void orc_jac(const orc_config *c, const double *x, const double *u, double *Fx, double *Fu);
/* one shooting interval of ERK(ns) x steps with forward sensitivities.
 * A row-major 13x13, B row-major 13x4 */
void orc_discretize(const orc_config *c, const double *x, const double *u, double *xn, double *A,
                    double *B);

/* QP data in delta form for one linearisation (all row-major, stage-major). */
typedef struct {
  int N;
  double *A;   /* N*169 */
  double *B;   /* N*52  */
  double *b;   /* N*13   phi_k - xbar_{k+1} */
  double *Q;   /* (N+1)*13 diagonal of x-Hessian (incl. LM) */
  double *R;   /* N*4 */
  double *q;   /* (N+1)*13 */
  double *r;   /* N*4 */
  double *lb;  /* N*4   lbu - ubar */
  double *ub;  /* N*4 */
  double *lbx; /* (N+1)*13  -state_box - xbar (rows of stages 1..N-1 used when xbox) */
  double *ubx; /* (N+1)*13 */
  int xbox;    /* state box rows present (orc_config.keep_state_box) */
} orc_qp;

orc_qp *orc_qp_alloc(int N);
void orc_qp_free(orc_qp *qp);

/* Linearise at (xtraj[(N+1)*13], utraj[N*4]); yref[N*17], yref_e[13]. */
void orc_build_qp(const orc_config *c, const double *xtraj, const double *utraj,
                  const double *yref, const double *yref_e, orc_qp *qp);

/* Box-constrained Riccati interior point (Mehrotra predictor-corrector).
 * dx[(N+1)*13] (dx_0 = 0), du[N*4], lam_l/lam_u[N*4] multipliers of the u-box.
 * Returns 0 ok, 2 iteration cap (tolerated by RTI), 3 min step, 4 NaN. */
int orc_qp_solve(const orc_qp *qp, int iter_max, double tol, double *dx, double *du, double *lam_l,
                 double *lam_u, int *iters);
int orc_qp_solve_warm(const orc_qp *qp, int iter_max, double tol, double *dx, double *du,
                      double *lam_l, double *lam_u, double *t_l_io, double *t_u_io, double warm_eps,
                      int *ipm_converged, int *iters);

/* Exact solve for a given active set (act[N*4]: 0 free, 1 lower, 2 upper) in one factorisation;
 * returns 0 when the point is the QP solution (primal feasible, multipliers >= 0), else 1. */
int orc_qp_solve_active_set(const orc_qp *qp, const unsigned char *act, double *dx, double *du,
                            double *lam_l, double *lam_u);

/* One PositionNMPC.solve() (controller.py:385-463) for a single problem.
 * xtraj/utraj/valid are the warm-start store (in/out).  x0 must already hold the
 * normalised quaternion.  u0[4] = zeros and valid=0 when status != 0. */
int orc_solve(const orc_config *c, const double *x0, const double *yref, const double *yref_e,
              double *xtraj, double *utraj, unsigned char *valid, double *u0, int *qp_iters,
              int *sqp_iters);

/* nodes/mpc_controller_node:124-131,152-164: clip, omega = sqrt(f/kf), last-valid fallback.
 * last_cmd[4] in/out. */
void orc_command(const orc_config *c, const double *u0, int status, double *last_cmd,
                 double *thrust, double *speed);

/* Loop of orc_solve + orc_command over a batch (same memory layout as the GPU handle). */
void orc_solve_batch(const orc_config *c, int batch, const double *x0, const double *yref,
                     const double *yref_e, double *xtraj, double *utraj, unsigned char *valid,
                     double *last_cmd, double *u0, int *status, int *qp_iters, double *thrust,
                     double *speed);

/* Nominal plant step (controller model, RK4 with `substeps` sub-steps over `dt`). */
void orc_plant_step(const orc_config *c, const double *x, const double *u, double dt,
                    int substeps, double *xn);

/* Plant with the optional RotorS first-order motor lag + rotor drag (see nmpc_oracle.c). */
void orc_plant_step_ext(const orc_config *c, const double *x, const double *u, double dt,
                        int substeps, int motor_model, double tau_up, double tau_down,
                        double rotor_drag, double mass, double *rotor, double *xn);

void orc_plant_step_full(const orc_config *c, const double *x, const double *u, double dt,
                         int substeps, int motor_model, double tau_up, double tau_down,
                         double rotor_drag, double mass, double rolling_moment, int rotors_yaw_sign,
                         double *rotor, double *xn);

#ifdef __cplusplus
}
#endif
#endif
