/*
 * nmpc_b200.h -- C ABI of the B200-native batched NMPC solver (libnmpc_b200.so).
 *
 * Drop-in boundary for what the reference's PositionNMPC obtains from acados'
 * generated solver through ctypes (`AcadosOcpSolver` create / set / solve / get):
 *   /root/reference/src/rotors_mpc_controller/controller.py
 *     :175-264  _create_solver  (OCP description -> codegen -> .so -> *_acados_create)
 *     :414-445  solver.set(stage, 'lbx'|'ubx'|'x'|'u'|'yref', ...)   64 calls / solve
 *     :447      solver.solve()
 *     :452-460  solver.get(stage, 'u'|'x')                             42 calls / solve
 *   /root/reference/nodes/mpc_controller_node:124-131,152-164  (command map + fail-safe)
 *
 * Plain pointers and sizes only; no torch / CUDA types in the signatures (a stream is
 * passed as void*).  All functions return 0 on success or a negative NMPC_E_* code and
 * never throw.  Per-problem solver status is DATA (int32[batch], acados-style codes).
 * A handle is bound to one CUDA device and is not thread-safe (the reference serialises
 * solve/reconfigure with a lock: nodes/mpc_controller_node:58,122,193).
 */
#ifndef NMPC_B200_H
#define NMPC_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NMPC_NX 13
#define NMPC_NU 4
#define NMPC_NY 17

/* API error codes (function return values) */
#define NMPC_OK 0
#define NMPC_E_INVALID (-1)   /* bad argument / unsupported configuration */
#define NMPC_E_CUDA (-2)      /* CUDA runtime error (see nmpc_last_error) */
#define NMPC_E_NOMEM (-3)
#define NMPC_E_NODEVICE (-4)  /* no CUDA device: there is NO CPU fallback */

/* per-problem solver status (acados enum; controller.py:448 tests `status != 0`) */
#define NMPC_STATUS_SUCCESS 0
#define NMPC_STATUS_NAN 1
#define NMPC_STATUS_MAXITER 2
#define NMPC_STATUS_MINSTEP 3
#define NMPC_STATUS_QP_FAILURE 4

#define NMPC_DISPATCH_AUTO 0
#define NMPC_DISPATCH_FIFO 1
#define NMPC_DISPATCH_LPT 2

#define NMPC_PRECISION_F64 0
#define NMPC_PRECISION_F32 1

/* Solver constants: the reference's ControllerParams (controller.py:24-49) and the acados
 * options of controller.py:179-190 made explicit.  Replaces the AcadosOcp description +
 * generated C code: no code generation or compiler is involved at run time. */
typedef struct nmpc_config {
  int32_t N;               /* horizon_steps (1..600, cfg/rotors_mpc.cfg:9) */
  int32_t qp_iter_max;     /* iter_max -> qp_solver_iter_max (controller.py:185) */
  int32_t integ_ns;        /* sim_method_num_stages: 1 Euler, 2 midpoint (ref), 4 RK4 */
  int32_t integ_steps;     /* sim_method_num_steps (ref: 2) */
  int32_t cost_scale_dt;   /* stage cost scaled by dt (acados default: 1) */
  int32_t lm_scaled_by_dt; /* LM term of stages 0..N-1 scaled by dt like the stage cost (acados: Ts[i] *
                              levenberg_marquardt; terminal stage unscaled).  Default 1; 0 = unscaled */
  int32_t max_sqp_iter;    /* 1 = SQP_RTI (reference); >1 = SQP until step <= sqp_tol */
  int32_t nan_status;      /* status reported when NaNs appear in the iterate (4 or 1) */
  int32_t precision;       /* NMPC_PRECISION_F64 | NMPC_PRECISION_F32 */
  int32_t ipm_warm_start;  /* 1: inside one solve with max_sqp_iter > 1, QPs after the first start the
                              interior point from the previous QP's slacks / multipliers (see
                              NMPC_WARM_*); 0: every QP starts cold.  RTI solves are not affected */
  int32_t active_set_shortcut; /* 1: every QP is first attacked by primal-dual active-set iterations:
                              the equality-constrained problem with the inputs of the current set fixed
                              on their bounds is solved exactly in ONE factorisation (the set starts
                              from the previous QP's active set inside one solve, else empty) and the
                              point is accepted iff it is primal and dual feasible -- then it IS the QP
                              solution; otherwise the set is updated, at most NMPC_AS_MAXIT times, and
                              finally the interior point takes over.  0: interior point only */
  int32_t dispatch;        /* order in which the problems of a batch are handed to the warps:
                              NMPC_DISPATCH_AUTO (0): longest-expected-first when max_sqp_iter > 1 and
                              the batch is larger than the number of resident warps, else by index;
                              NMPC_DISPATCH_FIFO (1): by index; NMPC_DISPATCH_LPT (2): always sorted.
                              Results do not depend on it (problems are independent) */
  double dt;
  double W[17];            /* stage weights [p(3) v(3) q(4) w(3) u(4)] (controller.py:234-240) */
  double We[13];           /* terminal weights (controller.py:241) */
  double lm;               /* levenberg_marquardt (controller.py:190) */
  double mass;
  double J[3];             /* inertia diagonal (controller.py:81-83) */
  double gravity;
  double rotor_x[4];       /* controller.py:100 */
  double rotor_y[4];       /* controller.py:101 */
  double rotor_z[4];       /* spin * km (controller.py:102-103) */
  double lbu[4];           /* controller.py:105-110, 248-250 */
  double ubu[4];
  double kf;               /* rotor_force_constant */
  double w_min;            /* motor_min_speed */
  double w_max;            /* motor_max_speed */
  double sqp_tol;          /* inf-norm step tolerance when max_sqp_iter > 1 */
  double ipm_tol;          /* interior-point complementarity / residual tolerance */
} nmpc_config_t;

/* Interior-point warm start between the QPs of one SQP run: used when the previous QP's
 * interior-point loop converged and the previous SQP step (inf-norm) was <= NMPC_WARM_STEP_MAX;
 * slacks and multipliers are clipped to eps = clamp(NMPC_WARM_KAPPA * step, NMPC_WARM_EPS_MIN,
 * NMPC_WARM_EPS_MAX).  The QP solutions are the same to the interior-point tolerance. */
#define NMPC_WARM_KAPPA 0.1
#define NMPC_WARM_STEP_MAX 0.3
#define NMPC_WARM_EPS_MIN 1e-8
#define NMPC_WARM_EPS_MAX 0.1
/* active-set iterations tried per QP before the interior point takes over */
#define NMPC_AS_MAXIT 32

typedef struct nmpc_handle nmpc_handle_t;

/* Replaces AcadosOcpSolver(ocp, json_file=...) (controller.py:263): allocates the device
 * workspace, warm-start store (x:(N+1,13), u:(N,4) per problem, controller.py:163-167), the
 * valid mask (False) and the last-valid command (hover thrust, mpc_controller_node:64). */
int nmpc_create(const nmpc_config_t *cfg, int32_t batch, int32_t device, nmpc_handle_t **out);
int nmpc_destroy(nmpc_handle_t *h);

/* Swap constants without re-allocating (same N / precision): the counterpart of building a new
 * PositionNMPC on every dynamic-reconfigure event (mpc_controller_node:196-198).  Invalidates
 * the warm start like controller.py:163-167. */
int nmpc_reconfigure(nmpc_handle_t *h, const nmpc_config_t *cfg);

/* Override the input box (default: controller.py:105-110). */
int nmpc_set_bounds(nmpc_handle_t *h, const double lbu[4], const double ubu[4]);

/* _prev_solution_valid = False (controller.py:167, :449) for problems with mask[i] != 0 (all when
 * mask == NULL).  `mask` is a HOST pointer; the call returns when the flags are cleared, so a
 * solve enqueued afterwards on any stream sees them.  It waits for this handle's last kernel only
 * (stream-scoped: other handles and streams of the process keep running). */
int nmpc_invalidate(nmpc_handle_t *h, const uint8_t *mask);
/* Same with a DEVICE mask (or NULL), asynchronous on `stream` (a cudaStream_t): one tiny kernel,
 * no host synchronisation -- for closed-loop runs that reset problems between control steps. */
int nmpc_invalidate_device(nmpc_handle_t *h, const uint8_t *mask, void *stream);
/* Reset the last-valid command to hover thrust (all problems). */
int nmpc_reset_last_command(nmpc_handle_t *h);

/* One PositionNMPC.solve() per problem + the node's command mapping, inputs and outputs in
 * DEVICE memory (FP64 regardless of compute precision):
 *   x0[batch,13] (quaternion already normalised, controller.py:406-411)
 *   yref[batch,N,17], yref_e[batch,13]                   (controller.py:433-445)
 *   u0[batch,4], status[batch] (int32), thrust[batch,4], speed[batch,4]; any output may be NULL.
 * Asynchronous on `stream` (a cudaStream_t, NULL = default stream).
 * Any 8-byte aligned pointers work; with a 16-byte aligned `yref` (cudaMalloc / torch allocations
 * are) and an even horizon the kernel stages the per-problem rows with TMA bulk copies, otherwise
 * it reads them with plain loads (same results).  Environment: NMPC_B200_NO_TMA=1 at
 * nmpc_create time forces the plain loads. */
int nmpc_solve_device(nmpc_handle_t *h, const double *x0, const double *yref, const double *yref_e,
                      double *u0, int32_t *status, double *thrust, double *speed, void *stream);

/* Allocation-free solve.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid allocations).  Same kernel and same results as nmpc_create +
 * nmpc_solve_device, but there is no handle: the caller owns the warm-start store, and the solver's
 * scratch memory (per-warp workspaces + the work-queue counter) is a file-scope static __device__
 * array of NMPC_STATIC_SCRATCH_BYTES that every GPU that loads the library holds until exit.  The
 * call makes no cudaMalloc / cudaFree / cudaHostAlloc and calls no allocating library routine.
 *   cfg                       as for nmpc_create (any precision / integrator / horizon within the limit)
 *   x0, yref, yref_e          inputs, device, as for nmpc_solve_device
 *   xtraj[batch,N+1,13], utraj[batch,N,4], valid[batch], last_cmd[batch,4]
 *                             the caller's warm-start store and last valid command, device, IN/OUT,
 *                             all required (start with valid = 0 and last_cmd = m g / 4)
 *   u0, status, thrust, speed, qp_iters, sqp_iters   outputs, device, any may be NULL
 * Limits (NMPC_E_INVALID + nmpc_last_error() beyond them): 1 <= batch <= NMPC_STATIC_MAX_BATCH, and
 * a horizon whose four per-warp workspaces (one CTA) fit the scratch array (N <= 600 in FP64 does).
 * The number of problems in flight is what fits the array (N = 20, FP64: 156 CTAs x 4 warps);
 * problems are handed out in index order.  The per-problem rows are read with plain loads, so the
 * caller's buffers need no padding and only 8-byte alignment.
 * ALL calls on one GPU share the array: calls that overlap on the GPU (different streams or
 * host threads of one process) race -- serialise them on one stream.  The array
 * keeps the previous call's contents; the only part that must start at zero (the work-queue
 * counter) is cleared by a memset node on `stream` in front of every launch.  The first call on a
 * device resolves the array's device address (cudaGetSymbolAddress) and sets the kernel's
 * shared-memory attribute: make that call outside a stream capture. */
#define NMPC_STATIC_SCRATCH_BYTES (40u << 20)
#define NMPC_STATIC_MAX_BATCH 65536
int nmpc_solve_static(const nmpc_config_t *cfg, int32_t batch, int32_t device, const double *x0,
                      const double *yref, const double *yref_e, double *xtraj, double *utraj,
                      uint8_t *valid, double *last_cmd, double *u0, int32_t *status, double *thrust,
                      double *speed, int32_t *qp_iters, int32_t *sqp_iters, void *stream);

/* Same with HOST buffers (pinned or pageable): H2D copies, solve, D2H copies, stream sync. */
int nmpc_solve_host(nmpc_handle_t *h, const double *x0, const double *yref, const double *yref_e,
                    double *u0, int32_t *status, double *thrust, double *speed);

/* Copy out the warm-start store / statistics (dst_is_device: 0 host, 1 device).
 * xtraj[batch,N+1,13], utraj[batch,N,4], valid[batch] (uint8), qp_iters[batch] (int32, number of
 * Riccati factorisations), sqp_iters[batch] (int32).  NULL pointers are skipped.  Synchronous for
 * the caller, stream-scoped on the device: the copies queue behind this handle's last kernel on
 * the handle's own stream (no cudaDeviceSynchronize; other streams keep running). */
int nmpc_get_trajectories(nmpc_handle_t *h, double *xtraj, double *utraj, uint8_t *valid,
                          int32_t dst_is_device);
int nmpc_get_stats(nmpc_handle_t *h, int32_t *qp_iters, int32_t *sqp_iters, int32_t dst_is_device);
/* Overwrite the warm-start store (host pointers), e.g. to resume a Monte-Carlo run. */
int nmpc_set_trajectories(nmpc_handle_t *h, const double *xtraj, const double *utraj,
                          const uint8_t *valid);

/* Test hooks on the linearisation stage: discretise (x[n,13],u[n,4]) -> xn[n,13], A[n,13,13],
 * B[n,13,4] (row-major, device pointers), i.e. what acados' ERK + expl_vde_forw returns. */
int nmpc_discretize_device(nmpc_handle_t *h, int32_t n, const double *x, const double *u,
                           double *xn, double *A, double *B, void *stream);

/* ---------------------------------------------------------------------------------------------
 * Closed loop on the device: everything a Monte-Carlo run needs
 * around the solve, so that states, references and commands never leave HBM.  All pointers are
 * DEVICE pointers, all calls asynchronous on `stream`.
 * ------------------------------------------------------------------------------------------- */

/* ReferenceGenerator.build_horizon (reference.py:75-91) + update_hover_thrust (:70-72) stacked as
 * controller.py:433-445 does, for every problem: pos[batch,3], yaw[batch] (NULL = 0) ->
 * yref[batch,N,17], yref_e[batch,13]. */
int nmpc_reference_setpoint_device(nmpc_handle_t *h, const double *pos, const double *yaw,
                                   double *yref, double *yref_e, void *stream);
/* Figure-8 track p(t) = c + (A sin wt, A/2 sin 2wt, 0), w = 2 pi / period, sampled (with its
 * velocity) at t0[b] + t_offset + k dt, k = 0..N. */
int nmpc_reference_figure8_device(nmpc_handle_t *h, const double *t0, const double *amplitude,
                                  const double *period, const double centre[3], double t_offset,
                                  double *yref, double *yref_e, void *stream);

/* nav_msgs/Odometry -> solver state exactly as nodes/mpc_controller_node:88-113 does (euler
 * angles :25-44, R(rpy) v_body :139-150, xyzw -> wxyz) followed by the quaternion normalisation of
 * controller.py:406-411.  odom[batch,13] = [p(3), q_xyzw(4), v_body(3), w(3)] -> x0[batch,13];
 * zero_quat[batch] (may be NULL) flags the inputs for which controller.py:407-408 raises. */
int nmpc_odometry_to_state_device(nmpc_handle_t *h, const double *odom, double *x0,
                                  int32_t *zero_quat, void *stream);

/* Plant used by closed-loop runs (RK4 of the controller model, controller.py:286-346). */
typedef struct nmpc_plant {
  int32_t substeps;    /* RK4 sub-steps per control interval (>= 1) */
  int32_t motor_model; /* 0: thrust = clipped command; 1: first-order rotor-speed lag
                          (rotors_gazebo_plugins common.h:147-183), thrust = kf w^2 */
  double dt;           /* control interval */
  double tau_up;       /* rotor time constants (hummingbird.xacro: 0.0125 / 0.025 s) */
  double tau_down;
  double rotor_drag;   /* rotor_drag_coefficient (gazebo_motor_model.cpp:443-453), 0 = off */
  double mass;         /* plant mass; <= 0 means the controller's */
  double rolling_moment; /* rolling_moment_coefficient (gazebo_motor_model.cpp:470-474,
                          hummingbird.xacro:42: 1e-6): world-frame torque -sum|w_i| mu v_perp; 0 = off */
  int32_t rotors_yaw_sign; /* 0: yaw torque with the CONTROLLER's sign (controller.py:102-103,328:
                          nominal plant, the loop is well posed); 1: RotorS' sign, drag torque
                          -turning_direction * thrust * km (gazebo_motor_model.cpp:458-459,
                          rotors_control common.h:92-93) -- the reference's real model mismatch */
  int32_t reserved0;   /* must be 0 */
} nmpc_plant_t;

/* One control interval for every problem: x[batch,13] (in/out, quaternion left as integrated),
 * x0[batch,13] (out, may be NULL: the state with the quaternion normalised, i.e. what solve()
 * is fed), thrust[batch,4] (clipped command; NULL = zero), rotor_speed[batch,4] (in/out, only for
 * motor_model 1).  With substeps = 0 only x0 is produced. */
int nmpc_plant_step_device(nmpc_handle_t *h, const nmpc_plant_t *plant, double *x, double *x0,
                           const double *thrust, double *rotor_speed, void *stream);

#define NMPC_REF_SETPOINT 0
#define NMPC_REF_FIGURE8 1
#define NMPC_NSTATS 8
/* Whole closed-loop run: `steps` times (reference -> solve with warm start -> command with
 * last-valid fallback -> plant step -> statistics).  2 or 3 kernel launches per step, no host
 * synchronisation inside.  stats[batch,8] (device, zeroed by the caller) accumulates per
 * problem: 0 sum |p-p_ref|^2, 1 last |p-p_ref|, 2 #solves with status != 0, 3 sum of Riccati
 * factorisations, 4 #steps with a rotor at a thrust bound, 5 max |p-p_ref|, 6 sum of thrusts,
 * 7 #non-finite states. */
typedef struct nmpc_rollout {
  int32_t steps;
  int32_t ref_mode;        /* NMPC_REF_SETPOINT | NMPC_REF_FIGURE8 */
  nmpc_plant_t plant;
  double *x;               /* [batch,13] plant state, in/out */
  double *rotor_speed;     /* [batch,4] in/out (motor_model 1) or NULL */
  const double *sp_pos;    /* set-point mode: [batch,3] */
  const double *sp_yaw;    /* [batch] or NULL */
  const double *f8_t0;     /* figure-8 mode: [batch] each */
  const double *f8_amplitude;
  const double *f8_period;
  double f8_centre[3];
  double t_start;          /* time of the first control step (figure-8 phase) */
  double *stats;           /* [batch,8] or NULL */
} nmpc_rollout_t;
int nmpc_rollout_device(nmpc_handle_t *h, const nmpc_rollout_t *run, void *stream);
/* Outputs of the last solve made by nmpc_rollout_device / nmpc_solve_host (device pointers owned
 * by the handle): u0[batch,4], status[batch], thrust[batch,4], speed[batch,4]. */
int nmpc_last_outputs(nmpc_handle_t *h, const double **u0, const int32_t **status,
                      const double **thrust, const double **speed);

/* Measured DFMA throughput of `device` in TFLOP/s (8 independent FMA chains per thread,
 * best of 4): the FP64 roofline denominator used by bench.py. */
int nmpc_fp64_peak(int32_t device, double *tflops);
/* FFMA twin: the FP32 roofline denominator of the FP32 variant (bench.py --dtype f32). */
int nmpc_fp32_peak(int32_t device, double *tflops);

/* Debug aid standing in for compute-sanitizer's initcheck (the tool is closed on this GPU pool):
 * with `on`, every launch first fills the whole workspace with NaN and the kernel fills its
 * shared-memory tile with NaN before every problem, so a read of anything the problem has not
 * written itself poisons its results.  Results must be bit-identical with and without it
 * (tests/test_gpu_api.py::test_poisoned_memory_changes_nothing). */
int nmpc_debug_poison(nmpc_handle_t *h, int32_t on);

/* Introspection */
int nmpc_batch(const nmpc_handle_t *h);
int nmpc_horizon(const nmpc_handle_t *h);
int64_t nmpc_workspace_bytes(const nmpc_handle_t *h);
int nmpc_kernel_launches(const nmpc_handle_t *h); /* kernels launched so far by this handle */
const char *nmpc_strerror(int code);
const char *nmpc_last_error(void); /* text of the last CUDA error seen by this library */
const char *nmpc_version(void);

#ifdef __cplusplus
}
#endif
#endif /* NMPC_B200_H */
