/* ctr_b200.h -- C ABI of libctr_b200.so: the B200-native batched CTR-Reach environment step.
 *
 * The reference (keshaviyengar/gym-ctr-reach) is pure Python and has no FFI layer; the boundary it
 * exposes for this path is its gym API (ctr_reach_envs/envs/ctr_reach_env.py).  Each entry point below
 * names the reference method it replaces (file:line relative to /root/reference/ctr_reach_envs/).
 * INTEGRATION.md shows the ctypes binding a maintainer of the reference would add.
 *
 * Conventions
 *  - plain C types only; every function returns 0 on success, non-zero on error; ctr_last_error()
 *    returns a thread-local message.  No exceptions cross the ABI.
 *  - a handle is bound to ONE CUDA device and is not thread-safe; handles are independent.
 *  - "dev" pointers are device pointers on the handle's device, "host" pointers are host memory
 *    (pinned or pageable).  Batched arrays are SoA: [field][n] (field-major, n contiguous).
 *  - `stream` is a cudaStream_t passed as void* (NULL = the legacy default stream).  Device-pointer
 *    calls are asynchronous on that stream.  *_host calls run on the handle's own non-blocking
 *    streams and synchronise before returning; they are ordered AFTER every earlier device-pointer
 *    call that writes env state (ctr_reset, ctr_step*, ctr_set_tube_params), whatever stream that
 *    call was given, so `ctr_reset(h, .., stream); ctr_step_host(h, ..)` needs no synchronisation
 *    by the caller.  Mixing streams between two device-pointer calls is the caller's to order.
 *  - there is NO CPU fallback: without a CUDA device ctr_create fails.
 */
#ifndef CTR_B200_H
#define CTR_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CTR_ABI_VERSION 2
#define CTR_MAX_SYSTEMS 8
#define CTR_NUM_TUBES 3
#define CTR_TUBE_FIELDS 8   /* length, length_curved, diameter_inner, diameter_outer, stiffness,
                               torsional_stiffness, x_curvature, y_curvature  (Tube.py:7-8) */
#define CTR_ENV_PARAMS 18   /* per-env derived constants: L[3], Lc[3], EI[3], GJ[3], Ux[3], Uy[3] */

enum { CTR_REP_EGOCENTRIC = 0, CTR_REP_PROPRIOCEPTIVE = 1 };   /* obs.py:147-151 */
enum { CTR_INTEGRATOR_DP45 = 0,   /* scipy RK45 restated: adaptive Dormand-Prince 5(4), the parity mode */
       CTR_INTEGRATOR_RK4 = 1 };  /* fixed-step classical RK4, rk4_steps per section (throughput mode)   */
enum { CTR_PRECISION_F64 = 0,     /* ODE state / RHS in float64 (reference arithmetic)                   */
       CTR_PRECISION_F32 = 1 };   /* ODE state / RHS in float32; segmentation stays float64              */
enum { CTR_REWARD_SPARSE = 0,     /* -(d > tol)  (ctr_reach_env.py:168-170)                              */
       CTR_REWARD_DENSE = 1 };    /* -d          (extension; BASELINE.json config 2)                     */

/* Mirrors the kwargs of CtrReachEnv.__init__ (ctr_reach_env.py:14-17) and register() (__init__.py:6-94). */
typedef struct ctr_config {
    int32_t abi_version;                 /* must be CTR_ABI_VERSION */
    int32_t n_systems;                   /* number of selected tube systems, 1..CTR_MAX_SYSTEMS */
    double tube[CTR_MAX_SYSTEMS][CTR_NUM_TUBES][CTR_TUBE_FIELDS];
    int32_t n_substeps;                  /* 10  (__init__.py:75)  */
    int32_t max_steps_per_episode;       /* 150 (__init__.py:74)  */
    int32_t joint_representation;        /* CTR_REP_*             */
    int32_t constrain_alpha;             /* clip alpha to +-pi (obs.py:64-69) */
    double extension_action_limit;       /* 0.001 m  (__init__.py:72) */
    double rotation_action_limit_deg;    /* 5 deg    (__init__.py:73) */
    int32_t resample_joints;             /* reset re-samples the starting joints (ctr_reach_env.py:104-112) */
    int32_t auto_reset;                  /* VecEnv semantics: finished envs are reset inside ctr_step */
    int32_t integrator;                  /* CTR_INTEGRATOR_* */
    int32_t precision;                   /* CTR_PRECISION_*  */
    int32_t rk4_steps;                   /* steps per section for CTR_INTEGRATOR_RK4 */
    int32_t reward_mode;                 /* CTR_REWARD_*     */
    double rtol, atol;                   /* 1e-3, 1e-6: solve_ivp defaults (model.py:149-151 passes none) */
    double initial_tol;                  /* goal tolerance until ctr_set_tolerance is called */
    double domain_rand;                  /* +-fraction on d_i, d_o, E, G, Ux at every reset (model_utils.py:15-19) */
    uint64_t seed;                       /* Philox key */
    int64_t env_offset;                  /* global index of this handle's env 0 (multi-GPU shards) */
    /* --- extensions beyond the reference (all 0 = reference behaviour) --- */
    double domain_rand_length;           /* +-fraction on L and L_c at every reset, needs domain_rand > 0; the joint limits,
                                            ordering constants and goal-sampling box then follow each env's own lengths
                                            (the reference forbids length randomisation, model_utils.py:12-14) */
    int32_t action_scale;                /* 1: actions arrive in [-1, 1] (a policy's output) and are multiplied by
                                            extension_action_limit / rotation_action_limit (ctr_reach_env.py:55-62) */
    int32_t action_clip;                 /* 1: out-of-Box action components are clipped and NaN components zeroed;
                                            0: passed through as given.  Either way every env whose action fails the
                                            reference's asserts (ctr_reach_env.py:131-132) is counted, ctr_fault_counters */
} ctr_config;

typedef struct ctr_handle ctr_handle;

/* State / output buffers owned by the handle (device memory, SoA).  ctr_buffer() returns the pointer. */
enum ctr_field {
    CTR_F_JOINTS = 0,          /* f32 [6][N]  current joints b0,b1,b2,a0,a1,a2 (Obs.joints)                 */
    CTR_F_DESIRED_GOAL,        /* f64 [3][N]  (env.desired_goal)                                            */
    CTR_F_DESIRED_JOINTS,      /* f32 [6][N]  (env.desired_joints)                                          */
    CTR_F_STARTING_JOINTS,     /* f32 [6][N]  (env.starting_joints)                                         */
    CTR_F_STARTING_POSITION,   /* f64 [3][N]  (env.starting_position)                                       */
    CTR_F_ACHIEVED_GOAL,       /* f64 [3][N]  tip position after the last step / reset                      */
    CTR_F_ORIENTATION,         /* f64 [9][N]  tip rotation matrix, row-major (extension; model.py:162)      */
    CTR_F_T,                   /* i32 [N]     steps in the current episode (env.t)                          */
    CTR_F_SYSTEM,              /* i32 [N]     index into the selected systems (env.system)                  */
    CTR_F_EPISODE,             /* u32 [N]     episode counter (RNG stream position)                         */
    CTR_F_OBS,                 /* f64 [D][N]  'observation' (obs.py:136-164), D = ctr_obs_dim()             */
    CTR_F_REWARD,              /* f64 [N]                                                                   */
    CTR_F_DONE,                /* u8  [N]                                                                   */
    CTR_F_SUCCESS,             /* u8  [N]     info['is_success']                                            */
    CTR_F_ERROR,               /* f64 [N]     info['error']                                                 */
    CTR_F_TERMINAL_OBS,        /* f64 [D][N]  observation of the finished episode (valid where done, auto_reset) */
    CTR_F_TERMINAL_ACHIEVED,   /* f64 [3][N]  achieved goal of the finished episode (valid where done)      */
    CTR_F_TUBE_PARAMS,         /* f64 [18][N] per-env derived constants (only after ctr_set_tube_params / domain_rand) */
    CTR_F_NFEV,                /* i32 [N]     RHS evaluations of the last FK of each env                    */
    CTR_F_STATS,               /* f64 [8]     see ctr_stats()                                               */
    CTR_F_COUNT
};

const char *ctr_last_error(void);
int ctr_abi_version(void);

/* CtrReachEnv.__init__ (ctr_reach_env.py:14-68): build N environments on `device`. */
int ctr_create(const ctr_config *cfg, int64_t n_envs, int32_t device, ctr_handle **out);
int ctr_destroy(ctr_handle *h);
int64_t ctr_num_envs(const ctr_handle *h);
int32_t ctr_obs_dim(const ctr_handle *h);   /* 13, or 14 when n_systems > 1 (obs.py:153-156) */
int ctr_buffer(ctr_handle *h, int32_t field, void **dev_ptr, int64_t *nbytes);

/* Derived per-system constants L,Lc,EI,GJ,Ux,Uy (Tube.py:14-15, Segment.py:57-59) as the device uses them. */
int ctr_system_constants(const ctr_handle *h, int32_t system, double out18[CTR_ENV_PARAMS]);

/* GoalTolerance.get_tol()/update() (goal_tolerance.py:34-55): the schedule runs on the host, the value is set here. */
int ctr_set_tolerance(ctr_handle *h, double tol);
double ctr_get_tolerance(const ctr_handle *h);

/* CtrReachEnv.seed (ctr_reach_env.py:116-122). */
int ctr_seed(ctr_handle *h, uint64_t seed);

/* Model.randomize_parameters (model.py:20-28): explicit per-env constants, f64 [18][N] on the device; NULL
 * returns to the per-system constants. */
int ctr_set_tube_params(ctr_handle *h, const double *params_dev, void *stream);

/* Model.forward_kinematics (model.py:30-70).  joints f64 [6][n]; system i32 [n] or NULL (system 0);
 * params f64 [18][n] or NULL; outputs: tip f64 [3][n], R f64 [9][n] or NULL, info i32 [4][n] or NULL
 * (nfev, accepted steps, rejected steps, sections).  integrator/precision < 0 use the handle's config.
 * Large launches (ctr_fk from 65536 solves, ctr_fk_shape from 32768, ctr_jacobian / ctr_ik_dls from 9363 configurations, the ctr_bvp_*
 * calls from 16384 passes) first build a hand-out
 * order of the configurations in a buffer of the handle (counting sort on section count and deployed length: a lock-step warp costs
 * its longest solve, so similar configurations are solved together).  It changes no result -- every output is addressed by
 * configuration; the one neighbour effect of the env-path kernels exists in any order: a warp that holds a solve with NaN or
 * |angle difference| >= 9e4 rad runs the library sincos for all its 10 solves, 1 ulp from the polynomial one -- but such calls on ONE handle must not overlap on the GPU (one stream, or one handle per stream); the buffer is
 * grown with cudaMalloc when a larger n arrives ($CTR_FK_SORT=0 switches the order off; ctr_jacobian_noalloc never uses it). */
int ctr_fk(ctr_handle *h, const double *joints_dev, const int32_t *system_dev, const double *params_dev, int64_t n,
           int32_t integrator, int32_t precision, double *tip_dev, double *R_dev, int32_t *info_dev, void *stream);
int ctr_fk_host(ctr_handle *h, const double *joints_host, const int32_t *system_host, int64_t n, int32_t integrator,
                int32_t precision, double *tip_host, double *R_host, int32_t *info_host);

/* Obs.set_action x n_substeps (obs.py:166-183, ctr_reach_env.py:134-135) on caller arrays:
 * joints f32 [6][n] in/out, action f32 [6][n], system i32 [n] or NULL. */
int ctr_set_action(ctr_handle *h, float *joints_dev, const float *action_dev, const int32_t *system_dev, int64_t n,
                   int32_t n_substeps, void *stream);

/* Overwrite env state (testing / checkpoint restore): any pointer may be NULL (left unchanged).
 * joints f32 [6][N], desired_goal f64 [3][N], system i32 [N], t i32 [N]; host pointers. */
int ctr_set_state_host(ctr_handle *h, const float *joints, const double *desired_goal, const int32_t *system,
                       const int32_t *t);

/* CtrReachEnv.reset(goal, system) (ctr_reach_env.py:70-114).  mask u8 [N] or NULL (all envs); goal f64 [3][N]
 * or NULL (sample, :98-101); system i32 [N] or NULL (sample uniformly, :95).  Writes joints, desired_*,
 * starting_*, achieved, t = 0 and the observation buffers. */
int ctr_reset(ctr_handle *h, const uint8_t *mask_dev, const double *goal_dev, const int32_t *system_dev, void *stream);

/* CtrReachEnv.step (ctr_reach_env.py:124-158) for all N envs: action f32 [6][N] on the device.  Results land in
 * the CTR_F_OBS / ACHIEVED_GOAL / DESIRED_GOAL / REWARD / DONE / SUCCESS / ERROR buffers. */
int ctr_step(ctr_handle *h, const float *action_dev, void *stream);
/* Same, with an arbitrary action layout: element (joint i, env e) is action_dev[i * stride_field + e * stride_env]
 * (stride_field = N, stride_env = 1 is the SoA layout above; 1 and 6 is a row-major (N, 6) policy output). */
int ctr_step_strided(ctr_handle *h, const float *action_dev, int64_t stride_field, int64_t stride_env, void *stream);

/* The same call with HOST buffers (the end-to-end path): copies action [6][N] host->device, steps, and copies
 * the results back.  Any output pointer may be NULL.  obs f64 [D][N], achieved/desired f64 [3][N],
 * reward f64 [N], done/success u8 [N], error f64 [N]. */
int ctr_step_host(ctr_handle *h, const float *action_host, double *obs_host, double *achieved_host, double *desired_host,
                  double *reward_host, uint8_t *done_host, uint8_t *success_host, double *error_host);
/* Same step, results into TWO host blocks laid out like the device's result block: out_f64 = f64 [(D + 8)][N] with rows
 * achieved (3) | reward (1) | error (1) | obs (D) | desired (3); out_u8 = u8 [2][N] with rows done | success.  Each
 * chunk of the pipeline then needs two DMA transfers instead of seven.  (ctr_step_host pipelines too: the env range
 * is cut into chunks whose H2D / kernel / D2H overlap, and the sequence is replayed as a CUDA graph per pointer set.) */
int ctr_step_host_packed(ctr_handle *h, const float *action_host, double *out_f64_host, uint8_t *out_u8_host);
/* Opt-in DELTA form of the packed contract (float64, same block layout): the caller passes the SAME two blocks every step and
 * keeps their content; the library transfers only what a step can change -- achieved | reward | error | obs rows 0..11 and the
 * flags -- patches desired_goal (and the system row of obs) for the envs that were auto-reset in this step from a short
 * device-built list, and rewrites the tolerance row of obs on the host when the tolerance changed.  After the call the blocks
 * hold exactly what ctr_step_host_packed would have written (tests compare them bit for bit); 138 instead of 170 bytes per
 * env-step cross PCIe.  The first call, a call with different blocks, and the first call after any other call that changes
 * env state (ctr_reset*, ctr_step*, ctr_set_state_host, ctr_set_tube_params, ctr_checkpoint_load) transfer everything. */
int ctr_step_host_delta(ctr_handle *h, const float *action_host, double *out_f64_host, uint8_t *out_u8_host);
/* Opt-in float32 form of the packed contract: out_f32_host is float [D + 8][N] in the same row order, converted on the device
 * from the float64 result block (which stays the handle's state and the default contract: the reference returns float64,
 * obs.py:158-162, although it declares float32 spaces, obs.py:89-132).  Half the device->host bytes per step. */
int ctr_step_host_packed_f32(ctr_handle *h, const float *action_host, float *out_f32_host, uint8_t *out_u8_host);
int ctr_reset_host(ctr_handle *h, double *obs_host, double *achieved_host, double *desired_host);

/* CtrReachEnv.compute_reward (ctr_reach_env.py:160-170) on arbitrary batches (HER relabelling):
 * achieved/desired f64 [3][m] -> reward f64 [m].  tol < 0 uses the handle's tolerance; mode < 0 the config's. */
int ctr_compute_reward(ctr_handle *h, const double *achieved_dev, const double *desired_dev, int64_t m, double tol,
                       int32_t mode, double *reward_dev, void *stream);
/* HER relabel (extension): reward as above and obs[9:12] = new_desired - achieved, obs f64 [D][m]. */
int ctr_relabel(ctr_handle *h, const double *achieved_dev, const double *new_desired_dev, int64_t m, double tol,
                int32_t mode, double *reward_dev, double *obs_dev, void *stream);

/* Episode statistics accumulated on the device since the last ctr_stats_reset: out8 (host) =
 * [episodes, successes, sum final error, sum return, sum episode length, env steps, resets, goal-sampling draws].
 * Multi-GPU callers all-reduce(SUM) these 8 doubles. */
int ctr_stats(ctr_handle *h, double out8_host[8], void *stream);
int ctr_stats_reset(ctr_handle *h, void *stream);

/* FK work since the last ctr_stats_reset: [RHS evaluations, RK step attempts, sections, FK solves] (roofline
 * accounting: algorithmic flops = nfev*120 + attempts*1290 + sections*200). */
int ctr_work_counters(ctr_handle *h, double out4_host[4], void *stream);
/* Where the reference raises per env, a batch counts (since the last ctr_stats_reset): out4 = FK solves that failed
 * (model.py:69 / :152-153), goal samplings stuck after 1001 draws (obs.py:204-205), env-steps whose action was NaN or
 * outside the action Box (ctr_reach_env.py:131-132), 0. */
int ctr_fault_counters(ctr_handle *h, double out4_host[4], void *stream);
/* Checkpoint / restore of the complete handle state: every device buffer the handle owns -- joints, goals, t,
 * system, episode counters (= RNG position), per-env tube parameters, result block, prefetched next episodes, refill queues,
 * slot permutation, statistics -- and the host scalars (tolerance, seed, step index).  Restores into a handle created with the
 * same configuration and env count; the two handles then produce identical bits for identical actions.  Synchronises. */
int64_t ctr_checkpoint_bytes(const ctr_handle *h);
int ctr_checkpoint_save(ctr_handle *h, void *host_buf, int64_t nbytes);
int ctr_checkpoint_load(ctr_handle *h, const void *host_buf, int64_t nbytes);
/* Segment(t1, t2, t3, base) (CTR_Python/Segment.py:6-61) as the env path's solver sees it, for direct comparison with the
 * reference's tables: joints f64 [6][n] (only beta is used) -> integration spans t0 / tf f64 [9][n] of the kept sections
 * (model.py:137-151: [S_k-1, S_k - 1e-6]), pattern i32 [9][n] (bit i: tube i present, bit 3 + i: tube i curved, i.e. EI / GJ
 * resp. U_x / U_y non-zero in the reference's tables), count i32 [n] = kept sections | zero-length spans << 8. */
int ctr_segment_tables(ctr_handle *h, const double *joints_dev, const int32_t *system_dev, const double *params_dev, int64_t n,
                       double *t0_dev, double *tf_dev, int32_t *pattern_dev, int32_t *count_dev, void *stream);
/* Read-only view of the step path's internal scheduling state, for invariant tests (compute-sanitizer is not available on
 * the build pool): slot -> env permutation (i32 [N]), per-window arrival counters (i32), refill-queue counters
 * (i32 [6]: count[3], cursor[3]) and lists (i32 [capacity]), prefetch-slot episode tags (u32 [N]); CTR_DBG_STEP_INDEX
 * returns the number of step launches so far in *count (queue roles: consume = index % 3, produce = (index + 1) % 3). */
enum { CTR_DBG_PERM = 0, CTR_DBG_WINDOW_COUNTERS, CTR_DBG_QUEUE_COUNTERS, CTR_DBG_QUEUE_0, CTR_DBG_QUEUE_1, CTR_DBG_QUEUE_2,
       CTR_DBG_SLOT_TAGS, CTR_DBG_STEP_INDEX };
int ctr_debug_buffer(ctr_handle *h, int32_t which, void **dev_ptr, int64_t *count);

/* Measured FMA-pipe peak of `device` in TFLOP/s (register-resident FMA chains; precision = CTR_PRECISION_*):
 * the roofline denominator of the FK kernels. */
int ctr_measure_fma_peak(int32_t device, int32_t precision, double *tflops_out);

/* Number of kernels launched by this handle so far (bench.py's gpu_launches claim). */
int64_t ctr_launch_count(const ctr_handle *h);

/* ---- backbone shape --------------------------------------------------------------------------
 * Model.r and tip_pos (ctr_reach_envs/envs/model.py:66-68,137-172): 30 samples per section of the backbone position
 * (solve_ivp's t_eval = np.linspace(span[k], span[k+1] - 1e-6, 30), scipy RK45 dense output) and, per tube, the index of
 * the first sample at or beyond L_k + beta_k - 1e-3, which splits r into r1 = r[tip1 : tip0 + 1], r2 = r[tip2 : tip1 + 1],
 * r3 = r[: tip2 + 1].  Replaces the data the matplotlib renderer (envs/ctr_3d_graph.py) and src/utils.py:52-54 consume.
 * joints f64 [6][n] -> shape f64 [n][max_pts][3] (one contiguous polyline per configuration -- the one array of this
 * ABI that is not SoA: the consumers want r[i] as (npts, 3), and the writes of an env land in consecutive sectors; samples
 * past npts are left untouched), npts i32 [n] (30 x sections),
 * tip_idx i32 [3][n], tip f64 [3][n] or NULL.  Always the float64 scipy-RK45 parity integrator.  A section table can hold
 * 9 sections: max_pts = 270 never truncates. */
int ctr_fk_shape(ctr_handle *h, const double *joints_dev, const int32_t *system_dev, const double *params_dev, int64_t n,
                 int32_t max_pts, double *shape_dev, int32_t *npts_dev, int32_t *tip_idx_dev, double *tip_dev, void *stream);

/* ---- finite-difference Jacobian and damped-least-squares inverse kinematics --------------------
 * The algorithm of the reference's baseline controller (src/jacobian_controller.py:19-73 dls_ik_position_only,
 * envs/CTR_Python/CTR_Model.py:251-262 jac): forward differences with step eps (reference 1e-4) on the 6 joints, 7 solves
 * per Jacobian; dq = J^T (J J^T + lam I)^-1 (target - tip), lam 0.25; a configuration stops once the error before its update
 * was below tol (reference 1e-3 m), at most max_iters (reference 500).  Joints are NOT clipped to the joint box (the reference
 * does not).  The reference script runs on the out-of-scope BVP model and cannot be executed: here the kinematics are
 * ctr_fk's, and parity is pinned to oracle/ik_oracle.py only.
 * ctr_jacobian: joints f64 [6][n] -> J f64 [18][n] (3 x 6 row-major: J[r * 6 + c] = d tip_r / d joint_c), tip f64 [3][n] or NULL.
 * ctr_ik_dls: target f64 [3][n], joints f64 [6][n] in / out -> err f64 [n] (the last error norm evaluated), iters i32 [n].
 * Synchronises `stream` (it polls convergence every 4 iterations). */
int ctr_jacobian(ctr_handle *h, const double *joints_dev, const int32_t *system_dev, int64_t n, double eps, double *J_dev,
                 double *tip_dev, void *stream);
int ctr_ik_dls(ctr_handle *h, const double *target_dev, double *joints_dev, const int32_t *system_dev, int64_t n, double lam,
               double eps, double tol, int32_t max_iters, double *err_dev, int32_t *iters_dev, void *stream);

/* ---- ctr_jacobian without any allocation ----------------------------------------------------------------------------
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same arguments, same kernels and same
 * results as ctr_jacobian, but the scratch of the 7n solves (perturbed joints, tips, system ids: 532 B per configuration)
 * is a file-scope static __device__ array of the library instead of the handle's grow-on-demand cudaMalloc, and the call
 * performs no cudaMalloc / cudaFree / allocating library call.
 *  - n > CTR_JACOBIAN_NOALLOC_MAX is rejected (return code 1, ctr_last_error says so); split the batch or use ctr_jacobian.
 *  - every GPU that loads the library holds the array (CTR_JACOBIAN_NOALLOC_MAX * 532 B = 8.3 MiB) until the process exits.
 *  - ALL calls on one device share the array, whatever the handle or stream: calls that overlap on the GPU race.  Calls
 *    on one stream are ordered and safe.  The array keeps the last call's contents; nothing in it has to start at zero
 *    (every value read is written earlier in the same call), so nothing is cleared.
 *  - the caller passes device inputs and outputs as for ctr_jacobian; the launches go to `stream`, nothing synchronises. */
#define CTR_JACOBIAN_NOALLOC_MAX 16384
int ctr_jacobian_noalloc(ctr_handle *h, const double *joints_dev, const int32_t *system_dev, int64_t n, double eps, double *J_dev,
                         double *tip_dev, void *stream);

/* ---- the reference's full model with tip load (CTR_Python/CTR_Model.py) as a second kinematics back-end
 * State of the model: u1x u1y | u1z u2z u3z | twist angles relative to tube 1 | r | R (20 values, ode_eq :56-128), integrated
 * section by section with scipy's RK23 and max_step (ode_solver :130-236).  All arrays are SoA f64 [rows][n].
 *   q     [6][n]  joints = q + q_0 of the reference (beta[3], alpha[3])
 *   f     [3][n]  tip force, NULL = no load;  u [5][n] initial curvatures u_init, NULL = zeros (what jacobian_controller.py passes)
 *   system i32 [n] or NULL (system 0);  params [18][n] per-configuration tube constants or NULL
 * ctr_bvp_pass      = CTR_Model.ode_solver(u_init), method 1: tip [3][n], final R [9][n] / residuals cost [5][n] (:228-235) /
 *                     accepted RK23 steps i32 [n] optional.  A failed pass (step too small) writes NaN and counts in
 *                     ctr_fault_counters()[0].
 * ctr_bvp_jacobian  = CTR_Model.jac (:251-262, eps 1e-4 on q, q restored by subtraction as there) and / or CTR_Model.comp
 *                     (:264-275, 0.05 N on f): J [18][n] (3 x 6 row-major) and / or C [9][n] (3 x 3), tip [3][n] optional.
 * ctr_bvp_solve     = the shooting problem of CTR_Model.minimize (:277-299): starts from the same initial guess and drives the
 *                     same 5 residuals below the same f_tol (max-norm, scipy.optimize.anderson's), by Levenberg-Marquardt steps on a
 *                     forward-difference 5 x 5 Jacobian (6 passes per iteration) instead of Anderson mixing.  u [5][n] out,
 *                     res [n] final residual max-norm, iters i32 [n]; tip / cost of the pass at the solution optional.  Synchronises.
 * ctr_bvp_ik_dls    = src/jacobian_controller.py:19-53 on this model: q [6][n] in/out, target [3][n]; as ctr_ik_dls.  Synchronises.
 * Streams: ctr_bvp_jacobian / _solve / _ik_dls keep their intermediate results in ONE scratch buffer of the handle (grown with
 * cudaMalloc on demand): calls on the same handle must not overlap on the GPU (one stream, or one handle per stream).  The same
 * holds for ctr_bvp_pass from 16384 passes on (its scheduling order -- which configuration a free lane group takes next; it
 * does not change any result -- is built in a buffer of the handle); below that it keeps no state but a pass counter taken from a
 * ring of 64, so up to 64 such launches may be in flight at once. */
int ctr_bvp_pass(ctr_handle *h, const double *q_dev, const double *f_dev, const double *u_dev, const int32_t *system_dev,
                 const double *params_dev, int64_t n, double max_step, double *tip_dev, double *R_dev, double *cost_dev,
                 int32_t *steps_dev, void *stream);
int ctr_bvp_jacobian(ctr_handle *h, const double *q_dev, const double *f_dev, const double *u_dev, const int32_t *system_dev,
                     const double *params_dev, int64_t n, double max_step, double *J_dev, double *C_dev, double *tip_dev,
                     void *stream);
int ctr_bvp_solve(ctr_handle *h, const double *q_dev, const double *f_dev, const int32_t *system_dev, const double *params_dev,
                  int64_t n, double max_step, double f_tol, int32_t max_iters, double *u_dev, double *tip_dev, double *cost_dev,
                  double *res_dev, int32_t *iters_dev, void *stream);
int ctr_bvp_ik_dls(ctr_handle *h, const double *target_dev, double *q_dev, const int32_t *system_dev, int64_t n, double lam,
                   double tol, int32_t max_iters, double max_step, double *err_dev, int32_t *iters_dev, void *stream);

/* ---- hindsight replay on the device --------------------------------------------------------
 * Replaces what the reference delegates to stable_baselines' HER wrapper (third party, not vendored, version not
 * pinned: setup.py:5; call sites src/utils.py:5-7,27-28; settings saved_policies/.../config.yml: goal_selection_strategy
 * future, n_sampled_goal 4, buffer_size 500000, batch_size 256).  The relabelled reward is the reference's own
 * compute_reward (ctr_reach_env.py:160-170).  The ring is CALLER-OWNED device memory, time major and RECORD major:
 * transition (slot, env) is record number slot * n_envs + env; a record is CTR_HER_RECORD doubles (352 bytes = 11 aligned
 * 32-byte sectors), so that sampling -- a gather of random records -- moves whole sectors:
 *   [0, D) obs | [D, 2D) next_obs | 2D + 0..2 achieved | 2D + 3..5 next_achieved | 2D + 6..8 desired | 2D + 9 reward |
 *   2D + 10..12 action (6 x f32 packed in 3 doubles) | padding | 42 done (0.0 / 1.0) | 43 (t_in_ep as the low, ep_len as the
 *   high 32-bit word).
 * The integers the index arithmetic needs before it can fetch anything else (done, t_in_ep, ep_len) live in the separate
 * arrays below -- the authoritative copy, for host-side readers -- AND in the last 16 bytes of the record, which is what the
 * sampler reads: one sector of the record it fetches anyway instead of three scattered 1- / 4-byte gathers. */
#define CTR_HER_RECORD 44
typedef struct {
    int64_t n_envs, capacity;      /* capacity = number of step slots                                            */
    int32_t obs_dim;               /* D = ctr_obs_dim()                                                          */
    int32_t pad_;
    double *rec;                   /* f64 [capacity*n][CTR_HER_RECORD], 32-byte aligned                          */
    uint8_t *done;                 /* u8  [capacity*n]                                                           */
    int32_t *t_in_ep;              /* i32 [capacity*n]  1-based index of the transition inside its episode        */
    int32_t *ep_len;               /* i32 [capacity*n]  length of its episode; 0 while the episode is still open  */
    double *cur_obs;               /* f64 [D][n]  observation the NEXT stored transition starts from              */
    double *cur_achieved, *cur_desired;           /* f64 [3][n]                                                  */
    int32_t *cur_t;                /* i32 [n]     transitions stored so far for the open episode of each env     */
} ctr_her_ring;

typedef struct {                   /* one sampled batch, caller-owned device memory, SoA [rows][m]                */
    int64_t m;
    double *obs, *next_obs;        /* f64 [D][m]                                                                 */
    double *achieved, *next_achieved, *desired;   /* f64 [3][m]; desired is the (possibly relabelled) goal        */
    float *action;                 /* f32 [6][m]                                                                 */
    double *reward;                /* f64 [m]                                                                    */
    uint8_t *done;                 /* u8  [m]   0 for relabelled transitions                                     */
    int64_t *src, *goal_src;       /* i64 [m]   ring column of the transition / of the goal's donor (-1: original) */
} ctr_her_batch;

/* Start recording: cur_* <- the handle's current obs / achieved / desired, cur_t <- 0 (call after ctr_reset). */
int ctr_her_begin(ctr_handle *h, const ctr_her_ring *ring, void *stream);
/* Store the transition of the ctr_step that was just enqueued on `stream` into ring slot `slot`; action as for
 * ctr_step_strided.  Finished episodes (done) take next_obs / next_achieved from the terminal buffers, get their
 * ep_len back-filled, and the env's cur_* continue from the auto-reset state. */
int ctr_her_store(ctr_handle *h, const ctr_her_ring *ring, const float *action_dev, int64_t stride_field,
                  int64_t stride_env, int64_t slot, void *stream);
/* Sample batch->m transitions uniformly from the `n_stored` newest slots (newest = `newest_slot`) with the 'future'
 * strategy: with probability k / (k + 1), k = n_sampled_goal, the desired goal is replaced by the achieved goal the
 * same episode reaches after a uniformly drawn later-or-equal transition (never for the last transition of an episode),
 * reward = compute_reward(next_achieved, goal) with tolerance `tol` (< 0: the handle's), done = 0.  rewrite_error != 0
 * also sets obs[9:12] / next_obs[9:12] = goal - achieved / next_achieved (extension; the reference's wrapper swaps only
 * the desired_goal entry).  Draws are Philox4x32-10 keyed by (seed; sample index, call): reproducible on the host. */
int ctr_her_sample(ctr_handle *h, const ctr_her_ring *ring, int64_t n_stored, int64_t newest_slot,
                   int32_t n_sampled_goal, int32_t rewrite_error, double tol, uint64_t seed, uint64_t call,
                   const ctr_her_batch *batch, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* CTR_B200_H */
