// ctr_kernels.cu -- kernels and C-ABI (include/ctr_b200.h) of libctr_b200.so.  sm_100a only; no CPU fallback.
//
// Kernels (all: 3 lanes per environment, 10 environments per warp, see ctr_device.cuh):
//   k_fk      joints -> tip / R / info                     (Model.forward_kinematics, model.py:30-70)
//   k_step    10 x set_action -> FK -> reward/done/success/error -> observation -> stats -> auto-reset of the envs
//             that finished, by their own warp             (CtrReachEnv.step, ctr_reach_env.py:124-158)
//   k_reset   goal / start sampling (Philox) -> 2 FK -> observation   (CtrReachEnv.reset, ctr_reach_env.py:70-114)
//   k_reward / k_relabel   HER compute_reward              (ctr_reach_env.py:160-170)
//   k_set_action           standalone constraint op        (Obs.set_action, obs.py:166-183)
//   k_mask_to_list         compaction of an explicit reset mask
#include <cuda_runtime.h>
#include <chrono>
#include <mutex>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <string>
#include <vector>

#include "../../include/ctr_b200.h"
#include "ctr_device.cuh"
#include "ctr_her.cuh"
#include "ctr_ik.cuh"
#include "ctr_bvp.cuh"

using namespace ctr;

namespace {

thread_local std::string g_err;

#define CTR_FAIL(msg)                                                           \
    do {                                                                        \
        g_err = std::string(__func__) + ": " + (msg);                           \
        return 1;                                                               \
    } while (0)
#define CTR_CUDA(call)                                                          \
    do {                                                                        \
        cudaError_t e__ = (call);                                               \
        if (e__ != cudaSuccess) {                                               \
            g_err = std::string(__func__) + ": " #call ": " + cudaGetErrorString(e__); \
            return 2;                                                           \
        }                                                                       \
    } while (0)

#ifndef CTR_BLOCK
#define CTR_BLOCK 128
#endif
constexpr int BLOCK = CTR_BLOCK;                 // 4 warps = 40 environments per block
#ifndef CTR_MIN_BLOCKS
#define CTR_MIN_BLOCKS (512 / CTR_BLOCK)
#endif
constexpr int MIN_BLOCKS = CTR_MIN_BLOCKS;       // 16 warps / SM: register cap 65536 / 512 = 128 (measured: 12 and 20 warps / SM are slower)
constexpr int ENVS_PER_BLOCK = (BLOCK / 32) * ENVS_PER_WARP;
// register limit of the FK kernels: by resident blocks (default) or, for occupancy experiments, as an explicit count
#ifdef CTR_MAXNREG
#define CTR_FK_BOUNDS __maxnreg__(CTR_MAXNREG)
#else
#define CTR_FK_BOUNDS __launch_bounds__(BLOCK, MIN_BLOCKS)
#endif
#ifndef CTR_SORT_WINDOW
#define CTR_SORT_WINDOW 640
#endif
#ifndef CTR_SORT_DESC
#define CTR_SORT_DESC 1
#endif
constexpr int SORT_WINDOW = CTR_SORT_WINDOW;                 // slots per sorting window (16 blocks); launches start on window boundaries
constexpr int N_STATS = 16;
enum { ST_EPISODES = 0, ST_SUCCESSES, ST_SUM_ERROR, ST_SUM_RETURN, ST_SUM_LENGTH, ST_STEPS, ST_RESETS, ST_DRAWS,
       ST_FK_FAILED, ST_GOAL_STUCK, ST_NFEV, ST_ATTEMPTS, ST_SECTIONS, ST_FK_COUNT, ST_INVALID_ACTION };

struct DevPtrs {
    float* joints;
    double* desired_goal;
    float* desired_joints;
    float* starting_joints;
    double* starting_position;
    double* achieved;
    double* orientation;
    int32_t* t;
    int32_t* system;
    uint32_t* episode;
    double* obs;
    double* reward;
    uint8_t* done;
    uint8_t* success;
    double* error;
    double* terminal_obs;
    double* terminal_achieved;
    double* tube_params;   // may be null
    int32_t* nfev;
    double* stats;
    int32_t* reset_list;
    int32_t* reset_count;
    // next-episode prefetch (auto-reset off the critical path, see k_step): the state reset() will produce for episode
    // pf_episode[env], computed ahead of time by the refill queue
    float* pf_joints;          // f32 [6][N]  starting joints
    float* pf_desired_joints;  // f32 [6][N]
    double* pf_desired_goal;   // f64 [3][N]
    double* pf_start;          // f64 [3][N]  starting position
    double* pf_orientation;    // f64 [9][N]
    double* pf_tube_params;    // f64 [18][N] (domain randomisation only, else null)
    int32_t* pf_system;        // i32 [N]
    int32_t* pf_nfev;          // i32 [N]
    uint32_t* pf_episode;      // u32 [N]  episode the slot was computed for (0 = invalid)
    int32_t* rq_list[3];       // refill queues, rotated per step: consume / produce / zero
    int32_t* rq_ctr;           // i32 [6]: count[3], cursor[3]
};

struct KArgs {
    DevPtrs p;
    const SysConst* sys;
    const double* raw_tubes;   // [n_systems][3][8] nominal tube table (domain randomisation)
    int64_t n;
    int32_t n_systems, n_substeps, max_steps, rep, reward_mode, auto_reset, resample_joints, rk4_steps, obs_dim;
    int32_t use_params;        // tube_params holds valid per-env constants
    double tol, rtol, atol, domain_rand;
    double domain_rand_length; // +-fraction on L and L_c at every reset (extension, 0 = reference behaviour)
    float act_lim_ext, act_lim_rot;   // action Box half-widths as float32 (ctr_reach_env.py:55-62)
    int32_t action_scale;      // actions arrive in [-1, 1] and are multiplied by the limits (extension)
    int32_t action_clip;       // out-of-box components are clipped, NaN components zeroed (default: passed through, counted)
    uint2 key;
    int64_t env_offset;
    int64_t act_stride_field, act_stride_env;   // action element (i, env) at action[i * stride_field + env * stride_env]
    int32_t prefetch;          // auto-reset through the prefetch slots (auto_reset && resample_joints)
    int32_t rq_consume, rq_produce, rq_zero;    // refill-queue roles of this launch (indices into rq_list / rq_ctr)
    int32_t rq_cap;            // capacity of each refill queue
    int32_t refill_blocks;     // leading blocks of the k_step grid that only drain the refill queue
    int64_t env_begin, env_count;               // env range of this launch (chunked host pipeline), whole handle by default
    int32_t* perm;             // slot -> env, sorted by last step's work inside windows of SORT_WINDOW slots (null: identity)
    int32_t* win_count;        // per-window arrival counters of the blocks (self-resetting)
    int32_t interleave;        // block b serves rank b / W of window b % W (W windows in the launch) instead of rank b % 16 of window b / 16
};

__device__ __forceinline__ void stage_sys(SysConst* s_sys, const SysConst* sys, int n_systems) {
    const int words = n_systems * (int)(sizeof(SysConst) / 4);
    const uint32_t* src = reinterpret_cast<const uint32_t*>(sys);
    uint32_t* dst = reinterpret_cast<uint32_t*>(s_sys);
    for (int i = threadIdx.x; i < words; i += blockDim.x) dst[i] = src[i];
    __syncthreads();
}

// Tube constants of one environment for the FK: LL = L[3], Lc[3] in registers, `der` -> the 15 derived constants in
// shared memory (the system's, or the group's scratch slot when per-env parameters are active).
__device__ __forceinline__ const double* env_consts(const SysConst& sc, const double* params, int64_t n, int64_t env,
                                                    bool use_params, bool writer, GroupScratch& gs, double (&LL)[6]) {
    if (use_params) {
        double C[18];
#pragma unroll
        for (int i = 0; i < 18; ++i) C[i] = __ldg(params + (int64_t)i * n + env);
#pragma unroll
        for (int i = 0; i < 6; ++i) LL[i] = C[i];
        if (writer) store_derived(C, gs);
        return gs.der;
    }
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        LL[i] = sc.L[i];
        LL[3 + i] = sc.Lc[i];
    }
    return sc.der;
}

// queue `env` for a refill of its prefetch slot (consumed at the front of the next k_step launch).  A full queue drops
// the entry: that env then resets in place (correct, only slower).
__device__ __forceinline__ void rq_push(const KArgs& a, int64_t env) {
    const int slot = atomicAdd(a.p.rq_ctr + a.rq_produce, 1);
    if (slot < a.rq_cap) a.p.rq_list[a.rq_produce][slot] = (int32_t)env;
}

constexpr int SLOTS_PER_WARP = ENVS_PER_WARP + 1;   // lanes 30 / 31 form an idle 11th group with its own scratch slot
__device__ __forceinline__ GroupScratch& group_slot(GroupScratch* s_gs, const LaneId& id) {
    return s_gs[(threadIdx.x >> 5) * SLOTS_PER_WARP + id.g];
}

// warp-level sum of a per-environment quantity (counted once per group: only lane j == 0 contributes)
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
    return v;
}

// observation of one env written cooperatively by its 3 lanes (obs.py:136-164, obs_utils.py:41-53,69-80).
// bf / af: the lane's own extension and rotation; bprev / aprev: those of tube j - 1 (0 for j == 0).
// proprioceptive: trig of the float32 joints in float32.  egocentric: prop2ego's np.diff(..., prepend=0) promotes the
// float32 joints to float64, so the first differences and the trig are float64 arithmetic.
__device__ __forceinline__ void write_obs(double* obs, int64_t n, int64_t env, int j, float bf, float af, float bprev,
                                          float aprev, int rep, double des_j, double ach_j, double tol, int system,
                                          int obs_dim) {
    double b, c, s;
    if (rep == CTR_REP_EGOCENTRIC) {
        b = __dsub_rn((double)bf, (double)bprev);
        const double a = __dsub_rn((double)af, (double)aprev);
        sincos(a, &s, &c);
    } else {
        float sf, cf;
        sincosf(af, &sf, &cf);
        b = (double)bf;
        c = (double)cf;
        s = (double)sf;
    }
    obs[(int64_t)(3 * j + 0) * n + env] = c;
    obs[(int64_t)(3 * j + 1) * n + env] = s;
    obs[(int64_t)(3 * j + 2) * n + env] = b;
    obs[(int64_t)(9 + j) * n + env] = __dsub_rn(des_j, ach_j);
    if (j == 0) obs[(int64_t)12 * n + env] = tol;
    if (j == 1 && obs_dim > 13) obs[(int64_t)13 * n + env] = (double)system;
}
__device__ __forceinline__ void write_obs(double* obs, int64_t n, int64_t env, int j, const float (&q)[6], int rep,
                                          double des_j, double ach_j, double tol, int system, int obs_dim) {
    write_obs(obs, n, env, j, pick3(j, q[0], q[1], q[2]), pick3(j, q[3], q[4], q[5]), pick3(j, 0.f, q[0], q[1]),
              pick3(j, 0.f, q[3], q[4]), rep, des_j, ach_j, tol, system, obs_dim);
}

// ------------------------------------------------------------------------------------------------------------
template <typename T, int INTEG>
__global__ void CTR_FK_BOUNDS k_fk(const SysConst* __restrict__ sys, int n_systems,
                                              const double* __restrict__ joints, const int32_t* __restrict__ system,
                                              const double* __restrict__ params, int64_t n, double rtol, double atol,
                                              int rk4_steps, double* __restrict__ tip, double* __restrict__ R,
                                              int32_t* __restrict__ info_out, double* __restrict__ stats,
                                              const int32_t* __restrict__ order, int group) {
    __shared__ SysConst s_sys[CTR_MAX_SYSTEMS];
    __shared__ GroupScratch s_gs[(BLOCK / 32) * SLOTS_PER_WARP];
    stage_sys(s_sys, sys, n_systems);
    const LaneId id = lane_id();
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t slot = warp * ENVS_PER_WARP + id.g;
    const bool valid = (id.g < ENVS_PER_WARP) && (slot < n);
    // work-sorted hand-out order (fk_order): slot -> column; `group` consecutive columns (the 7 solves of a Jacobian) stay together
    const int64_t env = (valid && order) ? (int64_t)order[slot / group] * group + slot % group : slot;
    const int64_t e = valid ? env : 0;
    int sysid = system ? system[e] : 0;
    sysid = min(max(sysid, 0), n_systems - 1);
    GroupScratch& gs = group_slot(s_gs, id);
    double LL[6];
    const double* der = env_consts(s_sys[sysid], params, n, e, params != nullptr, id.j == 0, gs, LL);
    double beta[3], alpha[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        beta[i] = joints[(int64_t)i * n + e];
        alpha[i] = joints[(int64_t)(3 + i) * n + e];
    }
    T y[6];
    FkInfo info;
    fk_group<T, INTEG>(id, valid, gs, LL, LL + 3, der, beta, alpha, rtol, atol, rk4_steps, y, info);
    if (valid) {
        tip[(int64_t)id.j * n + env] = (double)y[2];
        if (R) {
#pragma unroll
            for (int c = 0; c < 3; ++c) R[(int64_t)(3 * id.j + c) * n + env] = (double)y[3 + c];
        }
        if (info_out && id.j == 0) {
            info_out[env] = info.nfev;
            info_out[n + env] = info.nacc;
            info_out[2 * n + env] = info.nrej;
            info_out[3 * n + env] = info.nsec;
        }
        if (info.failed && id.j == 0 && stats) atomicAdd(stats + ST_FK_FAILED, 1.0);
    }
}

// Backbone shape: Model.r and tip_pos of the reference (model.py:66-68,137-172) -- 30 samples per
// section from the integrator's dense output, and for every tube the first sample index whose abscissa reaches
// L_k + beta_k - 1e-3 (the split points of r1 / r2 / r3).  shape f64 [n][max_pts][3], npts i32 [n], tip_idx i32 [3][n].
// UY0: every system has y_curvature == 0 and no per-env parameters are given (the instantiation k_step uses, see rhs()).
template <bool UY0>
__global__ void CTR_FK_BOUNDS k_fk_shape(const SysConst* __restrict__ sys, int n_systems,
                                                    const double* __restrict__ joints, const int32_t* __restrict__ system,
                                                    const double* __restrict__ params, int64_t n, double rtol, double atol,
                                                    int max_pts, double* __restrict__ shape, int32_t* __restrict__ npts,
                                                    int32_t* __restrict__ tip_idx, double* __restrict__ tip,
                                                    double* __restrict__ stats, const int32_t* __restrict__ order) {
    __shared__ SysConst s_sys[CTR_MAX_SYSTEMS];
    __shared__ GroupScratch s_gs[(BLOCK / 32) * SLOTS_PER_WARP];
    stage_sys(s_sys, sys, n_systems);
    const LaneId id = lane_id();
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t slot = warp * ENVS_PER_WARP + id.g;
    const bool valid = (id.g < ENVS_PER_WARP) && (slot < n);
    const int64_t env = (valid && order) ? (int64_t)order[slot] : slot;
    const int64_t e = valid ? env : 0;
    int sysid = system ? system[e] : 0;
    sysid = min(max(sysid, 0), n_systems - 1);
    GroupScratch& gs = group_slot(s_gs, id);
    double LL[6];
    const double* der = env_consts(s_sys[sysid], params, n, e, params != nullptr, id.j == 0, gs, LL);
    double beta[3], alpha[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        beta[i] = joints[(int64_t)i * n + e];
        alpha[i] = joints[(int64_t)(3 + i) * n + e];
    }
    double y[6];
    FkInfo info;
    ShapeOut so;
    so.base = shape + (e * max_pts) * 3 + id.j;      // [n][max_pts][3]: an env's polyline is contiguous, its 3 lanes write 24 B
    so.stride = 3;
    so.max_pts = valid ? max_pts : 0;
    fk_group_shape<UY0>(id, valid, gs, LL, LL + 3, der, beta, alpha, rtol, atol, y, info, so);
    if (valid) {
        // tip_pos[j] = np.argmax(Length >= d_tip[j] - 1e-3) over the concatenated abscissae (0 when none qualifies)
        const double thr = __dsub_rn(__dadd_rn(pick3(id.j, LL[0], LL[1], LL[2]), pick3(id.j, beta[0], beta[1], beta[2])), 1e-3);
        const int nsec = gs.nsec & 0xff;
        int found = 0;
        bool hit = false;
        for (int k = 0; k < nsec && !hit; ++k) {
            const double a = gs.t0[k], b = gs.tf[k];
            const double step = __ddiv_rn(__dsub_rn(b, a), (double)(SHAPE_SAMPLES - 1));
            for (int i = 0; i < SHAPE_SAMPLES; ++i) {
                if (shape_abscissa(i, a, step, b) >= thr) {
                    found = k * SHAPE_SAMPLES + i;
                    hit = true;
                    break;
                }
            }
        }
        tip_idx[(int64_t)id.j * n + env] = found;
        if (tip) tip[(int64_t)id.j * n + env] = y[2];
        if (id.j == 0) {
            npts[env] = info.shape_secs * SHAPE_SAMPLES;
            if (info.failed && stats) atomicAdd(stats + ST_FK_FAILED, 1.0);
        }
    }
}

// The section table of the env path exactly as fk_group sees it (build_sections, Segment.py:6-61 + the spans of model.py:137-151):
// t0 / tf f64 [9][n], pattern i32 [9][n] (bits 0-2 tube present, 3-5 tube curved), count i32 [n] = kept | zero-length << 8.
__global__ void CTR_FK_BOUNDS k_segment_tables(const SysConst* __restrict__ sys, int n_systems, const double* __restrict__ joints,
                                               const int32_t* __restrict__ system, const double* __restrict__ params, int64_t n,
                                               double* __restrict__ t0, double* __restrict__ tf, int32_t* __restrict__ pat,
                                               int32_t* __restrict__ count) {
    __shared__ SysConst s_sys[CTR_MAX_SYSTEMS];
    __shared__ GroupScratch s_gs[(BLOCK / 32) * SLOTS_PER_WARP];
    stage_sys(s_sys, sys, n_systems);
    const LaneId id = lane_id();
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t env = warp * ENVS_PER_WARP + id.g;
    const bool valid = (id.g < ENVS_PER_WARP) && (env < n);
    const int64_t e = valid ? env : 0;
    int sysid = system ? system[e] : 0;
    sysid = min(max(sysid, 0), n_systems - 1);
    GroupScratch& gs = group_slot(s_gs, id);
    double LL[6];
    env_consts(s_sys[sysid], params, n, e, params != nullptr, id.j == 0, gs, LL);
    double beta[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) beta[i] = joints[(int64_t)i * n + e];
    build_sections(LL, LL + 3, beta, gs, id.j == 0);
    if (valid) {
        const int m = gs.nsec & 0xff;
        for (int k = id.j; k < MAXSEC; k += 3) {
            t0[(int64_t)k * n + env] = k < m ? gs.t0[k] : 0.0;
            tf[(int64_t)k * n + env] = k < m ? gs.tf[k] : 0.0;
            pat[(int64_t)k * n + env] = k < m ? gs.pat[k] : 0;
        }
        if (id.j == 0) count[env] = gs.nsec;
    }
}

template <typename T, int INTEG>
__device__ __forceinline__ void reset_envs(const KArgs& a, const LaneId& id, const SysConst* s_sys, GroupScratch* s_gs,
                                           int64_t env, bool valid, const double* __restrict__ goal,
                                           const int32_t* __restrict__ system_in, bool to_slot);

// ------------------------------------------------------------------------------------------------------------
// k_step grid = [refill_blocks leading blocks: only drain the refill queue] + [step blocks: 40 environments each].
//
// Auto-reset off the critical path.  A reset costs two FK solves that depend only on (seed, env, episode), not on the
// step that ends the episode.  Doing them when the env finishes (in the same warp or in a following launch) puts one
// FK latency (~70-100 us) behind the last wave of a 540 us kernel.  Instead every env owns a PREFETCH SLOT holding the
// start state of its next episode; when it finishes, its lanes copy the slot (no FK) and queue the env for a refill,
// and the refill queue of step k is drained at the FRONT of the step k+1 launch (leading blocks, then any warp that
// runs out of step work).  A slot that is not ready (tag != episode + 1: first step of an episode, after seed()/
// explicit resets) falls back to the in-place reset by the env's own warp; both paths give identical bits.
template <typename T, int INTEG>
__global__ void CTR_FK_BOUNDS k_step(const KArgs a, const float* __restrict__ action) {
    __shared__ SysConst s_sys[CTR_MAX_SYSTEMS];
    __shared__ GroupScratch s_gs[(BLOCK / 32) * SLOTS_PER_WARP];
    stage_sys(s_sys, a.sys, a.n_systems);
    const LaneId id = lane_id();
    const int64_t n = a.n;
    const int j = id.j;
    const DevPtrs& p = a.p;
    if (a.prefetch && blockIdx.x == 0 && threadIdx.x == 0) {   // the queue the NEXT launch produces into
        p.rq_ctr[a.rq_zero] = 0;
        p.rq_ctr[3 + a.rq_zero] = 0;
    }
    unsigned pending = 0;          // group-leader lanes whose finished env must be reset in place
    int env32 = 0;                 // env index served by this lane's group
    // Dispatch order: inside a sorting window the slots run from the heaviest env to the lightest.  With `interleave` the grid
    // walks the windows round-robin -- the first W blocks are the heaviest block of every window, the last W the lightest of
    // every window -- so the launch as a whole runs longest-first and drains on its shortest blocks (window by window the
    // last ~600 resident blocks are a random mix of heavy and light ones).
    int64_t lblk = (int64_t)blockIdx.x - a.refill_blocks;          // logical block: 40 consecutive slots
    if (a.interleave && lblk >= 0) {
        constexpr int BPW = SORT_WINDOW / ENVS_PER_BLOCK;
        const int64_t nblk = (a.env_count + ENVS_PER_BLOCK - 1) / ENVS_PER_BLOCK;
        const int64_t W = (nblk + BPW - 1) / BPW;
        lblk = (lblk % W) * BPW + lblk / W;
        if (lblk >= nblk) return;                                   // the last window has fewer blocks (block-uniform exit)
    }
    if ((int)blockIdx.x >= a.refill_blocks) {
        const int64_t warp = (lblk * blockDim.x + threadIdx.x) >> 5;
        const int64_t local = warp * ENVS_PER_WARP + id.g;
        const bool valid = (id.g < ENVS_PER_WARP) && (local < a.env_count);
        // Work-sorted assignment: slot -> env through `perm`, which the previous step sorted by each env's RHS count
        // inside windows of 640 slots.  The 10 environments of a warp then need about the same number of step attempts
        // and start their sections together; a warp costs max-over-its-envs, so this removes ~10 % of the lock-step
        // waste.  Results do not depend on the assignment (every env is computed independently of its neighbours).
        const int64_t env = (valid && a.perm) ? (int64_t)a.perm[a.env_begin + local] : a.env_begin + local;
        env32 = (int)env;
        const int64_t e = valid ? env : a.env_begin;

        const int sysid = p.system[e];
        const SysConst& sc = s_sys[sysid];
        float q[6], act[6];
        bool bad_action = false;       // ctr_reach_env.py:131-132: the reference asserts; a batch cannot raise per env, it counts
#pragma unroll
        for (int i = 0; i < 6; ++i) {
            q[i] = p.joints[(int64_t)i * n + e];
            float v = action[(int64_t)i * a.act_stride_field + e * a.act_stride_env];
            const float lim_i = (i < 3) ? a.act_lim_ext : a.act_lim_rot;
            if (a.action_scale) {                                              // normalised policy output -> joint deltas
                bad_action |= !(fabsf(v) <= 1.0f);
                v = __fmul_rn(fminf(fmaxf(v, -1.0f), 1.0f), lim_i);            // NaN -> -lim (fmaxf drops the NaN); counted above
            } else {
                bad_action |= !(fabsf(v) <= lim_i);                            // action_space.contains(): NaN or outside the Box
            }
            if (a.action_clip) v = (v == v) ? fminf(fmaxf(v, -lim_i), lim_i) : 0.0f;
            act[i] = v;
        }
        JointLimits lim;
        if (a.use_params) {            // per-env tube lengths carry per-env joint limits
            double Le[3];
#pragma unroll
            for (int i = 0; i < 3; ++i) Le[i] = __ldg(p.tube_params + (int64_t)i * n + e);
            limits_of_lengths(sc, Le, lim);
        } else {
            limits_of_system(sc, lim);
        }
        set_action(q, act, lim, a.n_substeps);                                 // ctr_reach_env.py:134-135
        // the new joints are final here: store them now and fetch the lane's own pair back after the solve, so that no
        // joint value occupies a register across the integration loop (it runs at the 128-register limit)
        if (valid) {
            p.joints[(int64_t)j * n + env] = pick3(j, q[0], q[1], q[2]);
            p.joints[(int64_t)(3 + j) * n + env] = pick3(j, q[3], q[4], q[5]);
        }

        GroupScratch& gs = group_slot(s_gs, id);
        T y[6];
        FkInfo info;
        {
            double LL[6];
            const double* der = env_consts(sc, p.tube_params, n, e, a.use_params != 0, j == 0, gs, LL);
            double beta[3] = {(double)q[0], (double)q[1], (double)q[2]};           // model.py:50-51
            double alpha[3] = {(double)q[3], (double)q[4], (double)q[5]};
            fk_group<T, INTEG>(id, valid, gs, LL, LL + 3, der, beta, alpha, a.rtol, a.atol, a.rk4_steps, y, info);
        }
        const float bf = valid ? p.joints[(int64_t)j * n + env] : 0.f, af = valid ? p.joints[(int64_t)(3 + j) * n + env] : 0.f;
        float bprev = __shfl_up_sync(FULL, bf, 1), aprev = __shfl_up_sync(FULL, af, 1);
        if (j == 0) bprev = aprev = 0.f;

        // ---- reward / done / success / error (ctr_reach_env.py:138-140, 155-156, 168-170)
        const double ach_j = (double)y[2];
        const double des_j = p.desired_goal[(int64_t)j * n + e];
        const double diff = __dsub_rn(ach_j, des_j);
        const double sq = __dmul_rn(diff, diff);
        const double s0 = __shfl_sync(FULL, sq, id.l0), s1 = __shfl_sync(FULL, sq, id.l0 + 1), s2 = __shfl_sync(FULL, sq, id.l0 + 2);
        const double dist = __dsqrt_rn(__dadd_rn(__dadd_rn(s0, s1), s2));
        const int t_new = p.t[e] + 1;
        const bool far = !(dist <= a.tol);     // dist > tol (ctr_reach_env.py:169); a NaN distance (failed solve) is far, not done
        const double reward = (a.reward_mode == CTR_REWARD_DENSE) ? -dist : (far ? -1.0 : -0.0);
        const bool done = (!far) || (t_new >= a.max_steps);
        const bool success = dist < a.tol;
        const bool finish = valid && done && a.auto_reset;

        // prefetch slot ready?  (tag read by the group leader, acquire, then shared with the group)
        unsigned episode = 0;
        int ready_i = 0;
        if (finish && a.prefetch && j == 0) {
            episode = p.episode[env];
            const uint32_t tag = *((volatile uint32_t*)(p.pf_episode + env));
            ready_i = (tag == episode + 1u) ? 1 : 0;
            __threadfence();
        }
        const bool ready = __shfl_sync(FULL, ready_i, id.l0) != 0;
        episode = __shfl_sync(FULL, episode, id.l0);

        if (valid) {
            p.achieved[(int64_t)j * n + env] = ach_j;
#pragma unroll
            for (int c = 0; c < 3; ++c) p.orientation[(int64_t)(3 * j + c) * n + env] = (double)y[3 + c];
            write_obs(p.obs, n, env, j, bf, af, bprev, aprev, a.rep, des_j, ach_j, a.tol, sysid, a.obs_dim);
            if (finish) {
                write_obs(p.terminal_obs, n, env, j, bf, af, bprev, aprev, a.rep, des_j, ach_j, a.tol, sysid, a.obs_dim);
                p.terminal_achieved[(int64_t)j * n + env] = ach_j;
            }
            if (j == 0) {
                p.reward[env] = reward;
                p.done[env] = done ? 1 : 0;
                p.success[env] = success ? 1 : 0;
                p.error[env] = dist;
                p.t[env] = t_new;
                p.nfev[env] = info.nfev;
            }
        }
        // ---- consume the prefetched start state of the next episode (CtrReachEnv.reset without its two FK solves)
        if (finish && ready) {
            float qn[6];
#pragma unroll
            for (int i = 0; i < 6; ++i) qn[i] = __ldcg(p.pf_joints + (int64_t)i * n + env);
            const float nb = pick3(j, qn[0], qn[1], qn[2]), na = pick3(j, qn[3], qn[4], qn[5]);
            const double nd = __ldcg(p.pf_desired_goal + (int64_t)j * n + env);
            const double ns = __ldcg(p.pf_start + (int64_t)j * n + env);
            const int nsys = __ldcg(p.pf_system + env);
            p.joints[(int64_t)j * n + env] = nb;
            p.joints[(int64_t)(3 + j) * n + env] = na;
            p.starting_joints[(int64_t)j * n + env] = nb;
            p.starting_joints[(int64_t)(3 + j) * n + env] = na;
            p.desired_joints[(int64_t)j * n + env] = __ldcg(p.pf_desired_joints + (int64_t)j * n + env);
            p.desired_joints[(int64_t)(3 + j) * n + env] = __ldcg(p.pf_desired_joints + (int64_t)(3 + j) * n + env);
            p.desired_goal[(int64_t)j * n + env] = nd;
            p.starting_position[(int64_t)j * n + env] = ns;
            p.achieved[(int64_t)j * n + env] = ns;
#pragma unroll
            for (int c = 0; c < 3; ++c)
                p.orientation[(int64_t)(3 * j + c) * n + env] = __ldcg(p.pf_orientation + (int64_t)(3 * j + c) * n + env);
            if (p.pf_tube_params) {
#pragma unroll
                for (int i = 0; i < 6; ++i)
                    p.tube_params[(int64_t)(3 * i + j) * n + env] = __ldcg(p.pf_tube_params + (int64_t)(3 * i + j) * n + env);
            }
            write_obs(p.obs, n, env, j, qn, a.rep, nd, ns, a.tol, nsys, a.obs_dim);       // ctr_reach_env.py:114
            if (j == 0) {
                p.t[env] = 0;                                                             // :78
                p.system[env] = nsys;
                p.episode[env] = episode + 1u;
                p.nfev[env] = __ldcg(p.pf_nfev + env);
                atomicAdd(p.stats + ST_RESETS, 1.0);
                rq_push(a, env);                                                          // refill for the episode after
            }
        }
        pending = __ballot_sync(FULL, finish && !ready && j == 0);

        // ---- re-sort this window for the next step: the last of its blocks to arrive does a counting sort on nfev
        if (a.perm) {
            __shared__ int s_hist[256];
            __shared__ int s_last;
            const int64_t blk = lblk;
            constexpr int BLOCKS_PER_WINDOW = SORT_WINDOW / ENVS_PER_BLOCK;
            const int64_t win_local = blk / BLOCKS_PER_WINDOW;
            const int64_t slot0 = win_local * SORT_WINDOW;                           // first slot of the window, launch-relative
            const int64_t win = (a.env_begin + slot0) / SORT_WINDOW;
            const int nslots = (int)((a.env_count - slot0 < SORT_WINDOW) ? a.env_count - slot0 : SORT_WINDOW);
            const int nblocks = (nslots + ENVS_PER_BLOCK - 1) / ENVS_PER_BLOCK;
            __threadfence();
            __syncthreads();
            if (threadIdx.x == 0) {
                const int ticket = atomicAdd(a.win_count + win, 1);
                s_last = (ticket == nblocks - 1);
                if (s_last) a.win_count[win] = 0;
            }
            for (int i = threadIdx.x; i < 256; i += BLOCK) s_hist[i] = 0;
            __syncthreads();
            if (s_last) {                                                            // block-uniform
                __threadfence();
                const int64_t gbase = a.env_begin + slot0;                           // window = envs [gbase, gbase + nslots)
                constexpr int PER = (SORT_WINDOW + BLOCK - 1) / BLOCK;
                int bin[PER], rank[PER];
#pragma unroll
                for (int k = 0; k < PER; ++k) {
                    const int i = threadIdx.x + k * BLOCK;
                    bin[k] = -1;
                    if (i < nslots) {
                        bin[k] = min(max(__ldcg(p.nfev + gbase + i), 0), 255);
                        rank[k] = atomicAdd(&s_hist[bin[k]], 1);
                    }
                }
                __syncthreads();
                if (threadIdx.x < 32) {                                              // exclusive scan of the 256 bins
                    int v[8], sum = 0;
#pragma unroll
                    for (int k = 0; k < 8; ++k) {
                        v[k] = s_hist[threadIdx.x * 8 + k];
                        sum += v[k];
                    }
                    int incl = sum;
#pragma unroll
                    for (int o = 1; o < 32; o <<= 1) {
                        const int t = __shfl_up_sync(FULL, incl, o);
                        if ((int)threadIdx.x >= o) incl += t;
                    }
                    int run = incl - sum;
#pragma unroll
                    for (int k = 0; k < 8; ++k) {
                        s_hist[threadIdx.x * 8 + k] = run;
                        run += v[k];
                    }
                }
                __syncthreads();
#pragma unroll
                for (int k = 0; k < PER; ++k) {
                    const int i = threadIdx.x + k * BLOCK;
#if CTR_SORT_DESC
                    // heaviest environments first: the blocks dispatched last in a window (and in the launch) are its lightest
                    if (bin[k] >= 0) a.perm[gbase + (nslots - 1 - (s_hist[bin[k]] + rank[k]))] = (int32_t)(gbase + i);
#else
                    if (bin[k] >= 0) a.perm[gbase + s_hist[bin[k]] + rank[k]] = (int32_t)(gbase + i);
#endif
                }
            }
        }

        // ---- episode statistics: one atomic per warp per counter
        const bool lead = valid && j == 0;
        double v_ep = (lead && done) ? 1.0 : 0.0;
        double v_su = (lead && done && success) ? 1.0 : 0.0;
        double v_er = (lead && done) ? dist : 0.0;
        double v_rt = lead ? reward : 0.0;
        double v_ln = (lead && done) ? (double)t_new : 0.0;
        double v_st = lead ? 1.0 : 0.0;
        double v_fl = (lead && info.failed) ? 1.0 : 0.0;
        const unsigned bad_mask = __ballot_sync(FULL, lead && bad_action);
        double v_nf = lead ? (double)info.nfev : 0.0;
        double v_at = lead ? (double)(info.nacc + info.nrej) : 0.0;
        double v_sc = lead ? (double)info.nsec : 0.0;
        v_nf = warp_sum(v_nf); v_at = warp_sum(v_at); v_sc = warp_sum(v_sc);
        v_ep = warp_sum(v_ep); v_su = warp_sum(v_su); v_er = warp_sum(v_er); v_rt = warp_sum(v_rt);
        v_ln = warp_sum(v_ln); v_st = warp_sum(v_st); v_fl = warp_sum(v_fl);
        if (id.lane == 0) {
            if (v_ep != 0.0) {
                atomicAdd(p.stats + ST_EPISODES, v_ep);
                atomicAdd(p.stats + ST_SUM_ERROR, v_er);
                atomicAdd(p.stats + ST_SUM_LENGTH, v_ln);
                if (v_su != 0.0) atomicAdd(p.stats + ST_SUCCESSES, v_su);
            }
            atomicAdd(p.stats + ST_SUM_RETURN, v_rt);
            atomicAdd(p.stats + ST_STEPS, v_st);
            atomicAdd(p.stats + ST_NFEV, v_nf);
            atomicAdd(p.stats + ST_ATTEMPTS, v_at);
            atomicAdd(p.stats + ST_SECTIONS, v_sc);
            atomicAdd(p.stats + ST_FK_COUNT, v_st);
            if (v_fl != 0.0) atomicAdd(p.stats + ST_FK_FAILED, v_fl);
            if (bad_mask) atomicAdd(p.stats + ST_INVALID_ACTION, (double)__popc(bad_mask));
        }
    }
    // ---- reset work of this warp, 5 environments at a time (group pairs, see reset_envs): first the in-place resets
    //      of its own finished envs whose slot was not ready, then items of the refill queue until it is empty
    if (a.auto_reset) {
        for (;;) {                                                              // every condition is warp-uniform
            const bool from_queue = (pending == 0);
            const int pair = id.g >> 1;
            bool rvalid = false;
            int64_t renv = a.env_begin;
            if (!from_queue) {
                const unsigned src = __fns(pending, 0, pair + 1);               // leader lane of this pair's env
                rvalid = (id.g < ENVS_PER_WARP) && (src != 0xffffffffu);
                const int e32 = __shfl_sync(FULL, env32, src & 31);
                if (rvalid) renv = (int64_t)e32;
#pragma unroll
                for (int k = 0; k < 5; ++k) pending &= pending - 1;
            } else {
                if (!a.prefetch) break;
                int item0 = 0;
                if (id.lane == 0) item0 = atomicAdd(p.rq_ctr + 3 + a.rq_consume, 5);
                item0 = __shfl_sync(FULL, item0, 0);
                const int count = min(__ldcg(p.rq_ctr + a.rq_consume), a.rq_cap);
                if (item0 >= count) break;
                const int item = item0 + pair;
                rvalid = (id.g < ENVS_PER_WARP) && (item < count);
                if (rvalid) renv = (int64_t)__ldcg(p.rq_list[a.rq_consume] + item);
            }
            reset_envs<T, INTEG>(a, id, s_sys, s_gs, renv, rvalid, nullptr, nullptr, from_queue);
        }
    }
}

// ------------------------------------------------------------------------------------------------------------
// reset of up to 5 environments by one warp (CtrReachEnv.reset, ctr_reach_env.py:70-114): group pair p = g >> 1 serves
// `env` (the same value on the 6 lanes of the pair); the even group computes the desired goal, the odd group the
// starting position, so the two FK solves of a reset run side by side.  Shared by k_reset (explicit reset / mask) and by
// the tail of k_step (auto-reset of the environments that finished in this warp).  Results depend only on
// (seed, global env index, episode): who executes the reset does not matter.
template <typename T, int INTEG>
__device__ __forceinline__ void reset_envs(const KArgs& a, const LaneId& id, const SysConst* s_sys, GroupScratch* s_gs,
                                           int64_t env, bool valid, const double* __restrict__ goal,
                                           const int32_t* __restrict__ system_in, bool to_slot) {
    const DevPtrs& p = a.p;
    const int64_t n = a.n;
    const int j = id.j;
    const int which = id.g & 1;        // 0: desired goal, 1: starting position
    const uint64_t genv = (uint64_t)(a.env_offset + env);
    const unsigned env_lo = (unsigned)genv, env_hi = (unsigned)(genv >> 32);
    const unsigned episode = p.episode[env] + 1u;

    int sysid;
    if (system_in) sysid = min(max(system_in[env], 0), a.n_systems - 1);
    else if (a.n_systems > 1) {                                      // ctr_reach_env.py:95 randint(n_systems)
        uint4 r = philox4x32(make_uint4(env_lo, env_hi, episode, 2u << 16), a.key);
        sysid = (int)(((uint64_t)r.x * (uint64_t)a.n_systems) >> 32);
    } else sysid = 0;
    const SysConst& sc = s_sys[sysid];

    GroupScratch& gs = group_slot(s_gs, id);
    double LL[6];
    const double* der;
    __syncwarp();
    if (a.domain_rand > 0.0) {
        // Model.randomize_parameters (model.py:20-28, model_utils.py:5-35): U(v - v r, v + v r) on d_i, d_o, E, G, Ux
        double C[18];
        const double* nom = a.raw_tubes + (int64_t)sysid * 24;
#pragma unroll
        for (int tb = 0; tb < 3; ++tb) {
            uint4 r1 = philox4x32(make_uint4(env_lo, env_hi, episode, (3u << 16) | (unsigned)(2 * tb)), a.key);
            uint4 r2 = philox4x32(make_uint4(env_lo, env_hi, episode, (3u << 16) | (unsigned)(2 * tb + 1)), a.key);
            const double rr = a.domain_rand;
            auto rnd = [rr](double v, unsigned bits) { return (v - v * rr) + ((v + v * rr) - (v - v * rr)) * u01(bits); };
            double di = rnd(nom[tb * 8 + 2], r1.x), d_o = rnd(nom[tb * 8 + 3], r1.y);
            double E = rnd(nom[tb * 8 + 4], r1.z), Gm = rnd(nom[tb * 8 + 5], r1.w), ux = rnd(nom[tb * 8 + 6], r2.x);
            double do2 = d_o * d_o, di2 = di * di;
            double d4 = do2 * do2 - di2 * di2;
            // lengths: untouched in the reference (model_utils.py:12-14); domain_rand_length > 0 is the extension
            // BASELINE.json config 5 names -- the same U(v - v r, v + v r) on L and L_c, joint limits follow (JointLimits)
            const double rl = a.domain_rand_length;
            auto rndl = [rl](double v, unsigned bits) { return (v - v * rl) + ((v + v * rl) - (v - v * rl)) * u01(bits); };
            C[tb] = (rl > 0.0) ? rndl(nom[tb * 8 + 0], r2.y) : nom[tb * 8 + 0];
            C[3 + tb] = (rl > 0.0) ? rndl(nom[tb * 8 + 1], r2.z) : nom[tb * 8 + 1];
            C[6 + tb] = E * ((3.141592653589793 * d4) / 64);           // Tube.py:15
            C[9 + tb] = Gm * ((3.141592653589793 * d4) / 32);          // Tube.py:14
            C[12 + tb] = ux;
            C[15 + tb] = nom[tb * 8 + 7];
        }
#pragma unroll
        for (int i = 0; i < 6; ++i) LL[i] = C[i];
        if (j == 0) store_derived(C, gs);
        der = gs.der;
        double* tp = to_slot ? p.pf_tube_params : p.tube_params;
        if (valid && which == 1 && tp) {
#pragma unroll
            for (int i = 0; i < 6; ++i) tp[(int64_t)(3 * i + j) * n + env] = pick3(j, C[3 * i], C[3 * i + 1], C[3 * i + 2]);
        }
    } else {
        der = env_consts(sc, p.tube_params, n, env, a.use_params != 0, j == 0, gs, LL);
    }

    JointLimits lim;
    if (a.use_params) limits_of_lengths(sc, LL, lim);               // per-env lengths: per-env sampling box and predicates
    else limits_of_system(sc, lim);
    float q[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    bool do_fk = valid;
    int tries = 0;
    bool ok = true;
    if (which == 0) {
        if (goal) do_fk = false;                                     // ctr_reach_env.py:102-103
        else tries = sample_goal(lim, a.key, env_lo, env_hi, episode, 0u, q, ok);   // :100
    } else {
        if (a.resample_joints) tries = sample_goal(lim, a.key, env_lo, env_hi, episode, 1u, q, ok);   // :106
        else {
#pragma unroll
            for (int i = 0; i < 6; ++i) q[i] = p.joints[(int64_t)i * n + env];                      // :111
        }
    }
    if (!do_fk || !ok) {       // stuck sampling (the reference raises, obs.py:204-205): a defined, valid state -- all tubes
#pragma unroll                 // fully extended, alpha = 0 -- and the ST_GOAL_STUCK counter, which ctr_stats turns into rc 3
        for (int i = 0; i < 6; ++i) q[i] = 0.f;
    }
    double beta[3] = {(double)q[0], (double)q[1], (double)q[2]};
    double alpha[3] = {(double)q[3], (double)q[4], (double)q[5]};
    T y[6];
    FkInfo info;
    fk_group<T, INTEG>(id, do_fk, gs, LL, LL + 3, der, beta, alpha, a.rtol, a.atol, a.rk4_steps, y, info);
    double tip_j = (double)y[2];
    if (which == 0 && goal) tip_j = goal[(int64_t)j * n + env];
    const double des_j = __shfl_sync(FULL, tip_j, (id.lane + 32 - 3) & 31);   // the pair's desired-goal group
    const double tol = a.tol;
    if (valid && to_slot) {
        // prefetch slot of the NEXT episode: consumed by k_step when the env finishes (same values as the live path)
        if (which == 0) {
            p.pf_desired_goal[(int64_t)j * n + env] = tip_j;
            p.pf_desired_joints[(int64_t)j * n + env] = pick3(j, q[0], q[1], q[2]);
            p.pf_desired_joints[(int64_t)(3 + j) * n + env] = pick3(j, q[3], q[4], q[5]);
        } else {
            p.pf_joints[(int64_t)j * n + env] = pick3(j, q[0], q[1], q[2]);
            p.pf_joints[(int64_t)(3 + j) * n + env] = pick3(j, q[3], q[4], q[5]);
            p.pf_start[(int64_t)j * n + env] = tip_j;
#pragma unroll
            for (int c = 0; c < 3; ++c) p.pf_orientation[(int64_t)(3 * j + c) * n + env] = (double)y[3 + c];
            if (j == 0) {
                p.pf_system[env] = sysid;
                p.pf_nfev[env] = info.nfev;
            }
        }
    }
    if (to_slot) {
        // publish: data of all 6 lanes of the pair first, then the episode tag (release); a failed solve / stuck
        // sampling leaves the slot invalid, the consumer then takes the in-place path and reports the failure
        __threadfence();
        const unsigned badmask = __ballot_sync(FULL, valid && (info.failed || !ok));
        const unsigned pairmask = 0x3fu << (6 * (id.g >> 1));
        __syncwarp();
        if (valid && which == 1 && j == 0) {
            volatile uint32_t* tag = p.pf_episode + env;
            *tag = (badmask & pairmask) ? 0u : episode;
        }
    }
    if (valid && !to_slot) {
        if (which == 0) {
            p.desired_goal[(int64_t)j * n + env] = tip_j;
            if (!goal) {
                p.desired_joints[(int64_t)j * n + env] = pick3(j, q[0], q[1], q[2]);
                p.desired_joints[(int64_t)(3 + j) * n + env] = pick3(j, q[3], q[4], q[5]);
            }
        } else {
            const float bj = pick3(j, q[0], q[1], q[2]), aj = pick3(j, q[3], q[4], q[5]);
            p.joints[(int64_t)j * n + env] = bj;
            p.joints[(int64_t)(3 + j) * n + env] = aj;
            p.starting_joints[(int64_t)j * n + env] = bj;
            p.starting_joints[(int64_t)(3 + j) * n + env] = aj;
            p.starting_position[(int64_t)j * n + env] = tip_j;
            p.achieved[(int64_t)j * n + env] = tip_j;
#pragma unroll
            for (int c = 0; c < 3; ++c) p.orientation[(int64_t)(3 * j + c) * n + env] = (double)y[3 + c];
            write_obs(p.obs, n, env, j, q, a.rep, des_j, tip_j, tol, sysid, a.obs_dim);   // :114
            if (j == 0) {
                p.t[env] = 0;                                                             // :78
                p.system[env] = sysid;
                p.episode[env] = episode;
                p.nfev[env] = info.nfev;
                if (a.prefetch) rq_push(a, env);           // queue the prefetch of the episode after this one
            }
        }
    }
    if (valid) {
        if (j == 0) {
            if (which == 1 && !to_slot) atomicAdd(p.stats + ST_RESETS, 1.0);   // slot path: counted when consumed
            if (tries) atomicAdd(p.stats + ST_DRAWS, (double)tries);
            if (info.failed) atomicAdd(p.stats + ST_FK_FAILED, 1.0);
            if (!ok) atomicAdd(p.stats + ST_GOAL_STUCK, 1.0);                             // obs.py:204-205
            if (do_fk) {
                atomicAdd(p.stats + ST_NFEV, (double)info.nfev);
                atomicAdd(p.stats + ST_ATTEMPTS, (double)(info.nacc + info.nrej));
                atomicAdd(p.stats + ST_SECTIONS, (double)info.nsec);
                atomicAdd(p.stats + ST_FK_COUNT, 1.0);
            }
        }
    }
}

template <typename T, int INTEG>
__global__ void CTR_FK_BOUNDS k_reset(const KArgs a, const int32_t* __restrict__ list,
                                                 const int32_t* __restrict__ count_ptr, const double* __restrict__ goal,
                                                 const int32_t* __restrict__ system_in) {
    __shared__ SysConst s_sys[CTR_MAX_SYSTEMS];
    __shared__ GroupScratch s_gs[(BLOCK / 32) * SLOTS_PER_WARP];
    stage_sys(s_sys, a.sys, a.n_systems);
    const LaneId id = lane_id();
    const int64_t count = count_ptr ? (int64_t)*count_ptr : a.n;
    const int64_t warps_total = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; warp * 5 < count; warp += warps_total) {
        const int64_t item = warp * 5 + (id.g >> 1);
        const bool valid = (id.g < ENVS_PER_WARP) && (item < count);
        const int64_t env = valid ? (list ? (int64_t)list[item] : item) : 0;
        reset_envs<T, INTEG>(a, id, s_sys, s_gs, env, valid, goal, system_in, false);
    }
}

__global__ void k_mask_to_list(const uint8_t* __restrict__ mask, int64_t n, int32_t* __restrict__ list,
                               int32_t* __restrict__ count) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && mask[i]) list[atomicAdd(count, 1)] = (int32_t)i;
}

__global__ void k_set_action(const SysConst* __restrict__ sys, int n_systems, float* __restrict__ joints,
                             const float* __restrict__ action, const int32_t* __restrict__ system, int64_t n,
                             int n_substeps) {
    __shared__ SysConst s_sys[CTR_MAX_SYSTEMS];
    stage_sys(s_sys, sys, n_systems);
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int s = system ? min(max(system[i], 0), n_systems - 1) : 0;
    float q[6], a[6];
#pragma unroll
    for (int k = 0; k < 6; ++k) {
        q[k] = joints[(int64_t)k * n + i];
        a[k] = action[(int64_t)k * n + i];
    }
    JointLimits lim;
    limits_of_system(s_sys[s], lim);
    set_action(q, a, lim, n_substeps);
#pragma unroll
    for (int k = 0; k < 6; ++k) joints[(int64_t)k * n + i] = q[k];
}

// compute_reward (ctr_reach_env.py:160-170): streaming, 56 B per transition (+24 when the obs error is rewritten)
__global__ void k_reward(const double* __restrict__ ach, const double* __restrict__ des, int64_t m, double tol, int mode,
                         double* __restrict__ reward, double* __restrict__ obs) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < m; i += stride) {
        double a0 = ach[i], a1 = ach[m + i], a2 = ach[2 * m + i];
        double d0 = des[i], d1 = des[m + i], d2 = des[2 * m + i];
        double x = __dsub_rn(a0, d0), y = __dsub_rn(a1, d1), z = __dsub_rn(a2, d2);
        double dist = __dsqrt_rn(__dadd_rn(__dadd_rn(__dmul_rn(x, x), __dmul_rn(y, y)), __dmul_rn(z, z)));
        reward[i] = (mode == CTR_REWARD_DENSE) ? -dist : ((dist > tol) ? -1.0 : -0.0);
        if (obs) {
            obs[9 * m + i] = __dsub_rn(d0, a0);
            obs[10 * m + i] = __dsub_rn(d1, a1);
            obs[11 * m + i] = __dsub_rn(d2, a2);
        }
    }
}

// float32 copy of one chunk of the result block (rows obs | achieved | desired | reward | error, f64 [rows][N]) for the opt-in
// float32 host contract of ctr_step_host_packed_f32: half the bytes over PCIe; the float64 block on the device stays the truth
__global__ void k_pack_f32(const double* __restrict__ blk, float* __restrict__ out, int64_t n, int64_t begin, int64_t count,
                           int rows) {
    const int64_t total = count * rows;
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < total; k += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = k / count, i = begin + (k - r * count);
        out[r * n + i] = (float)blk[r * n + i];
    }
}

// ctr_step_host_delta: the envs of one chunk that were auto-reset in this step, i.e. whose desired goal (and system) changed:
// list[0] = count, then (env, desired[3], system) per entry, at most cap entries (count says when the list overflowed)
__global__ void k_delta_list(const uint8_t* __restrict__ done, const double* __restrict__ desired, const int32_t* __restrict__ system,
                             int64_t n, int64_t begin, int64_t count, int auto_reset, int64_t cap, double* __restrict__ list) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count || !auto_reset) return;
    const int64_t e = begin + i;
    if (!done[e]) return;
    const int64_t k = (int64_t)atomicAdd(list, 1.0);
    if (k >= cap) return;
    double* o = list + 1 + 5 * k;
    o[0] = (double)e;
    o[1] = desired[e];
    o[2] = desired[n + e];
    o[3] = desired[2 * n + e];
    o[4] = (double)system[e];
}

// register-resident FMA chains: the measured FP64 / FP32 FMA-pipe peak that bounds this path (no such figure in
// MEASURED_PEAKS.json)
template <typename T>
__global__ void __launch_bounds__(256) k_fma_peak(T* out, int iters, T b, T c) {
    T a[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) a[k] = (T)(threadIdx.x + k) * (T)1e-3;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int k = 0; k < 8; ++k) a[k] = fma_t(a[k], b, c);   // explicit: the library is built with -fmad=false
    }
    T s = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) s += a[k];
    out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = s;
}

}  // namespace

// ================================================================================================================
#ifndef CTR_COMPUTE_STREAMS
#define CTR_COMPUTE_STREAMS 4
#endif
constexpr int N_COMPUTE_STREAMS = CTR_COMPUTE_STREAMS;   // ctr_step_host: chunk kernels round-robin over this many streams

struct HostGraph {              // one captured ctr_step_host pipeline (keyed by the caller's host pointers)
    const void* key[8];
    int64_t per;
    int packed;                 // 0: separate arrays, 1: one float64 block, 2: one float32 block
    bool uy0;                   // k_step instantiation captured in the kernel nodes (U_y == 0 fast path or generic)
    uint64_t stamp;
    cudaGraph_t graph;
    cudaGraphExec_t exec;
    std::vector<cudaGraphNode_t> knodes;
    std::vector<cudaKernelNodeParams> kparams;
    std::vector<int64_t> begin, count;
    std::vector<int32_t> refill;
};

struct ctr_handle {
    ctr_config cfg;
    int64_t n;
    int32_t device;
    int32_t obs_dim;
    double tol;
    DevPtrs p;
    SysConst* sys_dev;
    double* raw_dev;
    std::vector<SysConst> sys_host;
    void* fields[CTR_F_COUNT];
    int64_t field_bytes[CTR_F_COUNT];
    std::vector<void*> allocs;
    std::vector<size_t> alloc_bytes;   // same order: what a checkpoint dumps
    cudaStream_t stream;       // used by the *_host entry points (= streams[0])
    cudaStream_t streams[N_COMPUTE_STREAMS + 1];   // ctr_step_host: compute streams (chunks round-robin) + one copy-back stream
    int host_flags_late;       // ctr_step_host: done / success bytes in one copy after the last chunk ($CTR_HOST_FLAGS_LATE, default 1)
    int host_chunks;           // chunks per ctr_step_host call (0: one per 16384 envs, at most 16); $CTR_HOST_CHUNKS
    int host_graphs;           // replay the pipeline as a CUDA graph (default; $CTR_HOST_GRAPH=0 submits call by call)
    cudaEvent_t ev_fork, ev_join[N_COMPUTE_STREAMS + 1], ev_chunk[16];
    cudaEvent_t ev_user;       // recorded on the caller's stream by every device-pointer call that changes env state
    int user_pending;          // ... and waited for by the next *_host call, which runs on the handle's private streams
    std::vector<HostGraph> graphs;
    uint64_t graph_clock;
    int use_params;
    int32_t* perm;             // work-sorted slot -> env map of k_step (null: disabled)
    int32_t* win_count;
    int all_uy_zero;           // every tube of every selected system has y_curvature == 0
    int prefetch;              // auto_reset && resample_joints: finished envs take their next start state from the slots
    int64_t step_index;        // k_step launches so far: rotates the three refill queues
    int32_t rq_cap;
    int64_t launches;
    int sm_count;
    int block_interleave;      // k_step grid walks the sorting windows round-robin ($CTR_BLOCK_INTERLEAVE, default 1)
    int refill_blocks;         // leading blocks of a k_step launch that only drain the refill queue ($CTR_REFILL_BLOCKS, default sm_count / 2)
    // lazily allocated staging for host calls
    float* stage_action;
    float* stage_f32;          // device float32 copy of the result block (ctr_step_host_packed_f32)
    double* blk;               // base of the result block: achieved | reward | error | obs | desired, f64 [D + 8][N]
    // ctr_step_host_delta: per-chunk lists of the envs whose desired goal / system changed in the step
    double* delta_dev;         // f64 [16][DELTA_STRIDE]: count, then (env, desired[3], system) x capacity
    double* delta_host;        // pinned mirror
    int64_t delta_cap;         // entries per chunk
    int64_t delta_per;         // envs per chunk of the last delta step
    const void* delta_block;   // host block the caller's persistent copy lives in (a different pointer = full transfer)
    double delta_tol;          // tolerance the host block's tolerance row holds
    int delta_dirty;           // env state changed outside ctr_step_host_delta: the next call transfers everything
    int64_t stage_fk_cap;
    void* stage_fk_dev;
    size_t bvp_cap;            // ctr_bvp_* scratch (bytes)
    void* bvp_dev;
    unsigned long long* bvp_next;   // work counters of k_bvp_pass_g (ring of 64)
    unsigned bvp_launches;
    void* bvp_order_dev;       // scheduling order of k_bvp_pass_g: order i32 [cap] | hist i32 [256] | key u8 [cap]
    int64_t bvp_order_cap;
    int bvp_sort;              // order the configurations before a large k_bvp_pass_g / k_fk / k_fk_shape launch ($CTR_BVP_SORT or $CTR_FK_SORT, default 1)
    int64_t ik_cap;            // ctr_jacobian / ctr_ik_dls scratch: joints f64 [6][7n] | tips f64 [3][7n] | system i32 [7n] | frozen u8 [n] | active i32
    void* ik_dev;
};

namespace {

void derive_system(const double tube[3][8], int constrain_alpha, SysConst& sc) {
    const float PI_F = 3.14159274101257324f;
    for (int i = 0; i < 3; ++i) {
        const double* t = tube[i];
        const double di = t[2], d_o = t[3];
        const double J = (M_PI * (pow(d_o, 4) - pow(di, 4))) / 32;   // Tube.py:14
        const double I = (M_PI * (pow(d_o, 4) - pow(di, 4))) / 64;   // Tube.py:15
        sc.L[i] = t[0];
        sc.Lc[i] = t[1];
        sc.EI[i] = t[4] * I;                                         // Segment.py:57
        sc.GJ[i] = t[5] * J;                                         // Segment.py:59
        sc.Ux[i] = t[6];
        sc.Uy[i] = t[7];
        sc.low[i] = (float)(-t[0] + 1e-3);                           // obs.py:14,59: -L + EXT_TOL as float32
        sc.high[i] = 0.f;
        sc.low[3 + i] = constrain_alpha ? -PI_F : -INFINITY;         // obs.py:64-75
        sc.high[3 + i] = constrain_alpha ? PI_F : INFINITY;
        sc.Lf[i] = (float)t[0];
        const double k = (sc.EI[i] != 0.0) ? sc.EI[i] / sc.GJ[i] : 0.0;   // model.py:101
        sc.der[DER_EI + i] = sc.EI[i];
        sc.der[DER_EIUX + i] = sc.EI[i] * sc.Ux[i];
        sc.der[DER_EIUY + i] = sc.EI[i] * sc.Uy[i];
        sc.der[DER_KX + i] = k * sc.Ux[i];
        sc.der[DER_KY + i] = k * sc.Uy[i];
    }
    sc.d10 = (float)(tube[1][0] - tube[0][0]);                       // obs.py:181 (python-float difference -> float32)
    sc.d21 = (float)(tube[2][0] - tube[1][0]);
}

template <typename Tp>
int dev_alloc(ctr_handle* h, Tp** out, int64_t count) {
    void* ptr = nullptr;
    size_t bytes = (size_t)(count > 0 ? count : 1) * sizeof(Tp);
    CTR_CUDA(cudaMalloc(&ptr, bytes));
    CTR_CUDA(cudaMemset(ptr, 0, bytes));
    h->allocs.push_back(ptr);
    h->alloc_bytes.push_back(bytes);
    *out = (Tp*)ptr;
    return 0;
}

KArgs make_args(const ctr_handle* h) {
    KArgs a;
    a.p = h->p;
    a.sys = h->sys_dev;
    a.raw_tubes = h->raw_dev;
    a.n = h->n;
    a.n_systems = h->cfg.n_systems;
    a.n_substeps = h->cfg.n_substeps;
    a.max_steps = h->cfg.max_steps_per_episode;
    a.rep = h->cfg.joint_representation;
    a.reward_mode = h->cfg.reward_mode;
    a.auto_reset = h->cfg.auto_reset;
    a.resample_joints = h->cfg.resample_joints;
    a.rk4_steps = h->cfg.rk4_steps;
    a.obs_dim = h->obs_dim;
    a.use_params = (h->use_params || h->cfg.domain_rand > 0.0) ? 1 : 0;   // randomised constants live in tube_params
    a.tol = h->tol;
    a.rtol = h->cfg.rtol;
    a.atol = h->cfg.atol;
    a.domain_rand = h->cfg.domain_rand;
    a.domain_rand_length = h->cfg.domain_rand > 0.0 ? h->cfg.domain_rand_length : 0.0;
    a.act_lim_ext = (float)h->cfg.extension_action_limit;                          // np.full(3, limit) in a float32 Box
    a.act_lim_rot = (float)(h->cfg.rotation_action_limit_deg * (M_PI / 180.0));    // np.deg2rad(limit) -> float32
    a.action_scale = h->cfg.action_scale;
    a.action_clip = h->cfg.action_clip;
    a.key = make_uint2((unsigned)h->cfg.seed, (unsigned)(h->cfg.seed >> 32));
    a.env_offset = h->cfg.env_offset;
    a.act_stride_field = h->n;
    a.act_stride_env = 1;
    a.prefetch = h->prefetch;
    a.rq_consume = (int32_t)(h->step_index % 3);
    a.rq_produce = (int32_t)((h->step_index + 1) % 3);
    a.rq_zero = (int32_t)((h->step_index + 2) % 3);
    a.rq_cap = h->rq_cap;
    a.refill_blocks = 0;
    a.env_begin = 0;
    a.env_count = h->n;
    a.perm = h->perm;
    a.interleave = (h->perm && h->block_interleave) ? 1 : 0;
    a.win_count = h->win_count;
    return a;
}

// Stream contract (include/ctr_b200.h): device-pointer calls run on the caller's stream, *_host calls on the handle's private
// streams.  mark_user_work() is called after a device-pointer call that writes env state; order_after_user() makes the private
// stream wait for it, so that `ctr_reset(stream); ctr_step_host(...)` needs no synchronisation by the caller.
int mark_user_work(ctr_handle* h, cudaStream_t st) {
    h->delta_dirty = 1;                    // env state changed outside ctr_step_host_delta: its next call transfers everything
    if (st == h->streams[0]) return 0;
    CTR_CUDA(cudaEventRecord(h->ev_user, st));
    h->user_pending = 1;
    return 0;
}
int order_after_user(ctr_handle* h, cudaStream_t priv) {
    if (!h->user_pending) return 0;
    CTR_CUDA(cudaStreamWaitEvent(priv, h->ev_user, 0));
    h->user_pending = 0;
    return 0;
}

inline int grid_for(int64_t n_envs) { return (int)((n_envs + ENVS_PER_BLOCK - 1) / ENVS_PER_BLOCK); }

// k_step additionally specialises on "every tube has y_curvature == 0" (template INTEG bit 1, see rhs)
#ifdef CTR_DEV_ONLY_D2   // development build: only the headline instantiations are compiled (ptxas -v / SASS turnaround)
#define DISPATCH_STEP(uy0, integ, prec, ...) do { k_step<double, 2> __VA_ARGS__; } while (0)
#define DISPATCH(KERNEL, integ, prec, ...) do { KERNEL<double, 0> __VA_ARGS__; } while (0)
#else
#define DISPATCH_STEP(uy0, integ, prec, ...)                                         \
    do {                                                                             \
        const int sel__ = ((integ) == CTR_INTEGRATOR_DP45 ? 0 : 1) | ((uy0) ? 2 : 0); \
        if ((prec) == CTR_PRECISION_F64) {                                           \
            if (sel__ == 0) k_step<double, 0> __VA_ARGS__;                           \
            else if (sel__ == 1) k_step<double, 1> __VA_ARGS__;                      \
            else if (sel__ == 2) k_step<double, 2> __VA_ARGS__;                      \
            else k_step<double, 3> __VA_ARGS__;                                      \
        } else {                                                                     \
            if (sel__ == 0) k_step<float, 0> __VA_ARGS__;                            \
            else if (sel__ == 1) k_step<float, 1> __VA_ARGS__;                       \
            else if (sel__ == 2) k_step<float, 2> __VA_ARGS__;                       \
            else k_step<float, 3> __VA_ARGS__;                                       \
        }                                                                            \
    } while (0)

#define DISPATCH(KERNEL, integ, prec, ...)                                           \
    do {                                                                             \
        if ((prec) == CTR_PRECISION_F64) {                                           \
            if ((integ) == CTR_INTEGRATOR_DP45) KERNEL<double, 0> __VA_ARGS__;       \
            else KERNEL<double, 1> __VA_ARGS__;                                      \
        } else {                                                                     \
            if ((integ) == CTR_INTEGRATOR_DP45) KERNEL<float, 0> __VA_ARGS__;        \
            else KERNEL<float, 1> __VA_ARGS__;                                       \
        }                                                                            \
    } while (0)
#endif

int launch_reset(ctr_handle* h, const int32_t* list, const int32_t* count_ptr, int64_t max_items, const double* goal,
                 const int32_t* system, cudaStream_t st) {
    if (max_items <= 0) return 0;
    KArgs a = make_args(h);
    a.rq_produce = a.rq_consume;            // explicit resets queue their prefetch for the next k_step launch
    int64_t warps = (max_items + 4) / 5;
    int64_t blocks = (warps + (BLOCK / 32) - 1) / (BLOCK / 32);
    // when the item count lives on the device, cap the grid (grid-stride loop inside) so that the common case of
    // few finished environments does not launch N/20 empty blocks
    if (count_ptr) blocks = blocks < (int64_t)h->sm_count * 8 ? blocks : (int64_t)h->sm_count * 8;
    DISPATCH(k_reset, h->cfg.integrator, h->cfg.precision, <<<(int)blocks, BLOCK, 0, st>>>(a, list, count_ptr, goal, system));
    h->launches += 1;
    CTR_CUDA(cudaGetLastError());
    return 0;
}

}  // namespace

extern "C" {

const char* ctr_last_error(void) { return g_err.c_str(); }
int ctr_abi_version(void) { return CTR_ABI_VERSION; }

static cudaError_t jacobian_scratch_resolve(int device, void** out);   // ctr_jacobian_noalloc, below

int ctr_create(const ctr_config* cfg, int64_t n_envs, int32_t device, ctr_handle** out) {
    if (!cfg || !out) CTR_FAIL("null argument");
    if (cfg->abi_version != CTR_ABI_VERSION) CTR_FAIL("ctr_config.abi_version mismatch");
    if (n_envs < 1 || n_envs > 2000000000LL) CTR_FAIL("n_envs out of range");
    if (cfg->n_systems < 1 || cfg->n_systems > CTR_MAX_SYSTEMS) CTR_FAIL("n_systems out of range");
    if (cfg->joint_representation != CTR_REP_EGOCENTRIC && cfg->joint_representation != CTR_REP_PROPRIOCEPTIVE)
        CTR_FAIL("Incorrect joint representation selected.");
    if (cfg->integrator != CTR_INTEGRATOR_DP45 && cfg->integrator != CTR_INTEGRATOR_RK4) CTR_FAIL("unknown integrator");
    if (cfg->precision != CTR_PRECISION_F64 && cfg->precision != CTR_PRECISION_F32) CTR_FAIL("unknown precision");
    if (cfg->integrator == CTR_INTEGRATOR_RK4 && cfg->rk4_steps < 1) CTR_FAIL("rk4_steps must be >= 1");
    if (cfg->n_substeps < 0 || cfg->max_steps_per_episode < 1) CTR_FAIL("bad n_substeps / max_steps_per_episode");
    if (!(cfg->domain_rand_length >= 0.0 && cfg->domain_rand_length < 0.5)) CTR_FAIL("domain_rand_length must be in [0, 0.5)");
    if (cfg->domain_rand_length > 0.0 && !(cfg->domain_rand > 0.0))
        CTR_FAIL("domain_rand_length needs domain_rand > 0 (lengths are drawn with the other tube parameters at reset)");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) CTR_FAIL("no CUDA device available (this library has no CPU fallback)");
    if (device < 0 || device >= ndev) CTR_FAIL("device index out of range");
    CTR_CUDA(cudaSetDevice(device));
    ctr_handle* h = new ctr_handle();
    h->cfg = *cfg;
    h->n = n_envs;
    h->device = device;
    h->obs_dim = cfg->n_systems > 1 ? 14 : 13;
    h->tol = cfg->initial_tol;
    h->use_params = 0;
    h->launches = 0;
    h->stage_action = nullptr;
    h->stage_f32 = nullptr;
    h->blk = nullptr;
    h->delta_dev = h->delta_host = nullptr;
    h->delta_cap = 0;
    h->delta_per = 0;
    h->delta_block = nullptr;
    h->delta_tol = 0.0;
    h->delta_dirty = 1;
    h->stage_fk_cap = 0;
    h->stage_fk_dev = nullptr;
    h->ik_cap = 0;
    h->ik_dev = nullptr;
    h->bvp_cap = 0;
    h->bvp_dev = nullptr;
    h->bvp_next = nullptr;
    h->bvp_launches = 0;
    h->bvp_order_dev = nullptr;
    h->bvp_order_cap = 0;
    h->bvp_sort = getenv("CTR_BVP_SORT") ? atoi(getenv("CTR_BVP_SORT")) : (getenv("CTR_FK_SORT") ? atoi(getenv("CTR_FK_SORT")) : 1);
    jacobian_scratch_resolve(h->device, nullptr);   // ctr_jacobian_noalloc: device address of its static scratch, once per device (its error, if any, is reported by that call)
    cudaDeviceProp prop;
    CTR_CUDA(cudaGetDeviceProperties(&prop, device));
    h->sm_count = prop.multiProcessorCount;
    h->block_interleave = getenv("CTR_BLOCK_INTERLEAVE") ? atoi(getenv("CTR_BLOCK_INTERLEAVE")) : 1;
    h->refill_blocks = getenv("CTR_REFILL_BLOCKS") ? atoi(getenv("CTR_REFILL_BLOCKS")) : h->sm_count / 2;   // sweep: profiles/r2i_refill_blocks_sweep.txt
    if (h->refill_blocks < 0) h->refill_blocks = 0;
    h->sys_host.resize(CTR_MAX_SYSTEMS);
    memset(h->sys_host.data(), 0, sizeof(SysConst) * CTR_MAX_SYSTEMS);
    for (int s = 0; s < cfg->n_systems; ++s) derive_system(cfg->tube[s], cfg->constrain_alpha, h->sys_host[s]);
    h->all_uy_zero = 1;
    for (int s = 0; s < cfg->n_systems; ++s)
        for (int i = 0; i < 3; ++i)
            if (cfg->tube[s][i][7] != 0.0) h->all_uy_zero = 0;
    if (getenv("CTR_NO_UY0")) h->all_uy_zero = 0;   // A/B switch: force the generic U_y instantiation (read once, here)
    const int64_t N = n_envs, D = h->obs_dim;
    int rc = 0;
    rc |= dev_alloc(h, &h->sys_dev, CTR_MAX_SYSTEMS);
    rc |= dev_alloc(h, &h->raw_dev, CTR_MAX_SYSTEMS * 24);
    // the per-step results live in ONE block, in the order achieved | reward | error | obs | desired (f64 rows) and
    // done | success (u8 rows): a host caller whose arrays have the same order gets two DMA transfers per chunk; what a
    // step can leave unchanged -- the tolerance row of obs (its last or second to last row), the system row, the desired
    // goals -- sits at the END, so that the delta contract (ctr_step_host_delta) moves one contiguous run of 17 rows
    {
        double* blk = nullptr;
        uint8_t* flg = nullptr;
        rc |= dev_alloc(h, &blk, (D + 8) * N);
        rc |= dev_alloc(h, &flg, 2 * N);
        if (!rc) {
            h->blk = blk;                                     // achieved | reward | error | obs | desired
            h->p.achieved = blk;
            h->p.reward = blk + 3 * N;
            h->p.error = blk + 4 * N;
            h->p.obs = blk + 5 * N;
            h->p.desired_goal = blk + (5 + D) * N;
            h->p.done = flg;
            h->p.success = flg + N;
        }
    }
    rc |= dev_alloc(h, &h->p.joints, 6 * N);
    rc |= dev_alloc(h, &h->p.desired_joints, 6 * N);
    rc |= dev_alloc(h, &h->p.starting_joints, 6 * N);
    rc |= dev_alloc(h, &h->p.starting_position, 3 * N);
    rc |= dev_alloc(h, &h->p.orientation, 9 * N);
    rc |= dev_alloc(h, &h->p.t, N);
    rc |= dev_alloc(h, &h->p.system, N);
    rc |= dev_alloc(h, &h->p.episode, N);
    rc |= dev_alloc(h, &h->p.terminal_obs, D * N);
    rc |= dev_alloc(h, &h->p.terminal_achieved, 3 * N);
    rc |= dev_alloc(h, &h->p.tube_params, 18 * N);
    rc |= dev_alloc(h, &h->p.nfev, N);
    rc |= dev_alloc(h, &h->p.stats, N_STATS);
    rc |= dev_alloc(h, &h->p.reset_list, N);
    rc |= dev_alloc(h, &h->p.reset_count, 1);
    h->perm = nullptr;
    h->win_count = nullptr;
    if (cfg->integrator == CTR_INTEGRATOR_DP45 && !getenv("CTR_NO_SORT")) {   // fixed-step RK4 has nothing to balance
        rc |= dev_alloc(h, &h->perm, N);
        rc |= dev_alloc(h, &h->win_count, (N + SORT_WINDOW - 1) / SORT_WINDOW);
        if (!rc) {
            std::vector<int32_t> ident((size_t)N);
            for (int64_t i = 0; i < N; ++i) ident[(size_t)i] = (int32_t)i;
            CTR_CUDA(cudaMemcpy(h->perm, ident.data(), sizeof(int32_t) * N, cudaMemcpyHostToDevice));
        }
    }
    h->prefetch = (cfg->auto_reset && cfg->resample_joints) ? 1 : 0;
    h->step_index = 0;
    h->rq_cap = (int32_t)(2 * N < 2000000000LL ? 2 * N : 2000000000LL);
    h->p.pf_tube_params = nullptr;
    if (h->prefetch) {
        rc |= dev_alloc(h, &h->p.pf_joints, 6 * N);
        rc |= dev_alloc(h, &h->p.pf_desired_joints, 6 * N);
        rc |= dev_alloc(h, &h->p.pf_desired_goal, 3 * N);
        rc |= dev_alloc(h, &h->p.pf_start, 3 * N);
        rc |= dev_alloc(h, &h->p.pf_orientation, 9 * N);
        if (cfg->domain_rand > 0.0) rc |= dev_alloc(h, &h->p.pf_tube_params, 18 * N);
        rc |= dev_alloc(h, &h->p.pf_system, N);
        rc |= dev_alloc(h, &h->p.pf_nfev, N);
        rc |= dev_alloc(h, &h->p.pf_episode, N);
        for (int k = 0; k < 3; ++k) rc |= dev_alloc(h, &h->p.rq_list[k], (int64_t)h->rq_cap);
        rc |= dev_alloc(h, &h->p.rq_ctr, 6);
    }
    if (rc) {
        std::string keep = g_err;
        ctr_destroy(h);
        g_err = keep;
        return 2;
    }
    CTR_CUDA(cudaMemcpy(h->sys_dev, h->sys_host.data(), sizeof(SysConst) * CTR_MAX_SYSTEMS, cudaMemcpyHostToDevice));
    CTR_CUDA(cudaMemcpy(h->raw_dev, cfg->tube, sizeof(double) * CTR_MAX_SYSTEMS * 24, cudaMemcpyHostToDevice));
    CTR_CUDA(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
    h->streams[0] = h->stream;
    for (int k = 1; k <= N_COMPUTE_STREAMS; ++k) {
        CTR_CUDA(cudaStreamCreateWithFlags(&h->streams[k], cudaStreamNonBlocking));
        CTR_CUDA(cudaEventCreateWithFlags(&h->ev_join[k], cudaEventDisableTiming));
    }
    for (int k = 0; k < 16; ++k) CTR_CUDA(cudaEventCreateWithFlags(&h->ev_chunk[k], cudaEventDisableTiming));
    h->host_chunks = getenv("CTR_HOST_CHUNKS") ? atoi(getenv("CTR_HOST_CHUNKS")) : 0;
    h->host_graphs = getenv("CTR_HOST_GRAPH") ? atoi(getenv("CTR_HOST_GRAPH")) : 1;
    h->host_flags_late = getenv("CTR_HOST_FLAGS_LATE") ? atoi(getenv("CTR_HOST_FLAGS_LATE")) : 1;
    h->graph_clock = 0;
    CTR_CUDA(cudaEventCreateWithFlags(&h->ev_fork, cudaEventDisableTiming));
    CTR_CUDA(cudaEventCreateWithFlags(&h->ev_user, cudaEventDisableTiming));
    h->user_pending = 0;
    void* f[CTR_F_COUNT] = {h->p.joints, h->p.desired_goal, h->p.desired_joints, h->p.starting_joints,
                            h->p.starting_position, h->p.achieved, h->p.orientation, h->p.t, h->p.system, h->p.episode,
                            h->p.obs, h->p.reward, h->p.done, h->p.success, h->p.error, h->p.terminal_obs,
                            h->p.terminal_achieved, h->p.tube_params, h->p.nfev, h->p.stats};
    int64_t b[CTR_F_COUNT] = {24 * N, 24 * N, 24 * N, 24 * N, 24 * N, 24 * N, 72 * N, 4 * N, 4 * N, 4 * N,
                              8 * D * N, 8 * N, N, N, 8 * N, 8 * D * N, 24 * N, 144 * N, 4 * N, 8 * 8};
    for (int i = 0; i < CTR_F_COUNT; ++i) {
        h->fields[i] = f[i];
        h->field_bytes[i] = b[i];
    }
    *out = h;
    return 0;
}

int ctr_destroy(ctr_handle* h) {
    if (!h) return 0;
    cudaSetDevice(h->device);
    cudaDeviceSynchronize();
    for (void* ptr : h->allocs) cudaFree(ptr);
    if (h->stage_f32) cudaFree(h->stage_f32);
    if (h->delta_dev) cudaFree(h->delta_dev);
    if (h->delta_host) cudaFreeHost(h->delta_host);
    if (h->stage_action) cudaFree(h->stage_action);      // device staging of ctr_step_host (cudaMalloc, step_host_impl)
    if (h->stage_fk_dev) cudaFree(h->stage_fk_dev);
    if (h->ik_dev) cudaFree(h->ik_dev);
    if (h->bvp_dev) cudaFree(h->bvp_dev);
    if (h->bvp_next) cudaFree(h->bvp_next);
    if (h->bvp_order_dev) cudaFree(h->bvp_order_dev);
    if (h->stream) cudaStreamDestroy(h->stream);
    for (int k = 1; k <= N_COMPUTE_STREAMS; ++k) {
        if (h->streams[k]) cudaStreamDestroy(h->streams[k]);
        if (h->ev_join[k]) cudaEventDestroy(h->ev_join[k]);
    }
    for (int k = 0; k < 16; ++k)
        if (h->ev_chunk[k]) cudaEventDestroy(h->ev_chunk[k]);
    for (auto& g : h->graphs) {
        cudaGraphExecDestroy(g.exec);
        cudaGraphDestroy(g.graph);
    }
    if (h->ev_fork) cudaEventDestroy(h->ev_fork);
    if (h->ev_user) cudaEventDestroy(h->ev_user);
    const cudaError_t pending = cudaGetLastError();      // never hand a stale error to the next handle's first launch check
    delete h;
    if (pending != cudaSuccess) {
        g_err = std::string("ctr_destroy: ") + cudaGetErrorString(pending);
        return 2;
    }
    return 0;
}

int64_t ctr_num_envs(const ctr_handle* h) { return h ? h->n : -1; }
int32_t ctr_obs_dim(const ctr_handle* h) { return h ? h->obs_dim : -1; }
int64_t ctr_launch_count(const ctr_handle* h) { return h ? h->launches : -1; }

int ctr_buffer(ctr_handle* h, int32_t field, void** dev_ptr, int64_t* nbytes) {
    if (!h || field < 0 || field >= CTR_F_COUNT) CTR_FAIL("bad handle / field");
    if (dev_ptr) *dev_ptr = h->fields[field];
    if (nbytes) *nbytes = h->field_bytes[field];
    return 0;
}

int ctr_system_constants(const ctr_handle* h, int32_t system, double out18[CTR_ENV_PARAMS]) {
    if (!h || system < 0 || system >= h->cfg.n_systems) CTR_FAIL("bad handle / system");
    const SysConst& s = h->sys_host[system];
    for (int i = 0; i < 3; ++i) {
        out18[i] = s.L[i];
        out18[3 + i] = s.Lc[i];
        out18[6 + i] = s.EI[i];
        out18[9 + i] = s.GJ[i];
        out18[12 + i] = s.Ux[i];
        out18[15 + i] = s.Uy[i];
    }
    return 0;
}

int ctr_set_tolerance(ctr_handle* h, double tol) {
    if (!h) CTR_FAIL("null handle");
    h->tol = tol;
    return 0;
}
double ctr_get_tolerance(const ctr_handle* h) { return h ? h->tol : -1.0; }

int ctr_seed(ctr_handle* h, uint64_t seed) {
    if (!h) CTR_FAIL("null handle");
    h->cfg.seed = seed;
    if (h->prefetch) {                       // the prefetched episodes were drawn with the old key
        CTR_CUDA(cudaSetDevice(h->device));
        CTR_CUDA(cudaDeviceSynchronize());
        CTR_CUDA(cudaMemset(h->p.pf_episode, 0, sizeof(uint32_t) * h->n));
    }
    return 0;
}

int ctr_set_tube_params(ctr_handle* h, const double* params_dev, void* stream) {
    if (!h) CTR_FAIL("null handle");
    CTR_CUDA(cudaSetDevice(h->device));
    if (h->prefetch)                         // the prefetched FK solves used the old constants
        CTR_CUDA(cudaMemsetAsync(h->p.pf_episode, 0, sizeof(uint32_t) * h->n, (cudaStream_t)stream));
    if (!params_dev) {
        h->use_params = 0;
        return mark_user_work(h, (cudaStream_t)stream);
    }
    CTR_CUDA(cudaMemcpyAsync(h->p.tube_params, params_dev, sizeof(double) * 18 * h->n, cudaMemcpyDeviceToDevice,
                             (cudaStream_t)stream));
    h->use_params = 1;
    return mark_user_work(h, (cudaStream_t)stream);
}

// Work-sorted hand-out order of n configurations (k_bvp_order_* in ctr_bvp.cuh: counting sort on section count and deployed length)
// in the handle's order buffer; *order = null below 4096 configurations or with $CTR_FK_SORT=0 / $CTR_BVP_SORT=0.
static int build_order(ctr_handle* h, const double* q_dev, const int32_t* system_dev, const double* params_dev, int64_t n,
                       int many_first, int64_t min_n, cudaStream_t st, const int32_t** order_out);
// the pre-pass costs ~20 us (a memset and three small launches): it pays from ~65536 env-path solves (k_fk: 3 ns each, 10 % gained)
// and from ~16384 BVP passes (14 ns each, 19 % gained) on
constexpr int64_t ORDER_MIN_FK_SOLVES = 65536, ORDER_MIN_BVP_PASSES = 16384;

// columns = n_cols joint columns; order (or null) permutes groups of `group` consecutive columns
static int fk_launch(ctr_handle* h, const double* joints_dev, const int32_t* system_dev, const double* params_dev, int64_t n,
                     int32_t integrator, int32_t precision, double* tip_dev, double* R_dev, int32_t* info_dev, cudaStream_t st,
                     const int32_t* order, int group) {
    if (integrator < 0) integrator = h->cfg.integrator;
    if (precision < 0) precision = h->cfg.precision;
    int rk4 = h->cfg.rk4_steps > 0 ? h->cfg.rk4_steps : 8;
    DISPATCH(k_fk, integrator, precision,
             <<<grid_for(n), BLOCK, 0, st>>>(h->sys_dev, h->cfg.n_systems, joints_dev, system_dev, params_dev, n,
                                             h->cfg.rtol, h->cfg.atol, rk4, tip_dev, R_dev, info_dev, h->p.stats, order, group));
    h->launches += 1;
    CTR_CUDA(cudaGetLastError());
    return 0;
}

int ctr_fk(ctr_handle* h, const double* joints_dev, const int32_t* system_dev, const double* params_dev, int64_t n,
           int32_t integrator, int32_t precision, double* tip_dev, double* R_dev, int32_t* info_dev, void* stream) {
    if (!h || !joints_dev || !tip_dev) CTR_FAIL("null argument");
    if (n <= 0) return 0;
    CTR_CUDA(cudaSetDevice(h->device));
    cudaStream_t st = (cudaStream_t)stream;
    // a lock-step warp costs its longest solve: hand the configurations out sorted by what their section tables say about their length
    const int32_t* order = nullptr;
    int rc = build_order(h, joints_dev, system_dev, params_dev, n, 1, ORDER_MIN_FK_SOLVES, st, &order);
    if (rc) return rc;
    return fk_launch(h, joints_dev, system_dev, params_dev, n, integrator, precision, tip_dev, R_dev, info_dev, st, order, 1);
}

int ctr_fk_host(ctr_handle* h, const double* joints_host, const int32_t* system_host, int64_t n, int32_t integrator,
                int32_t precision, double* tip_host, double* R_host, int32_t* info_host) {
    if (!h || !joints_host || !tip_host) CTR_FAIL("null argument");
    if (n <= 0) return 0;
    CTR_CUDA(cudaSetDevice(h->device));
    // device scratch: joints 6n f64 | tip 3n f64 | R 9n f64 | system n i32 | info 4n i32
    const size_t bytes = (size_t)n * (8 * 18 + 4 * 5);
    if (h->stage_fk_cap < n) {
        if (h->stage_fk_dev) CTR_CUDA(cudaFree(h->stage_fk_dev));
        h->stage_fk_dev = nullptr;
        CTR_CUDA(cudaMalloc(&h->stage_fk_dev, bytes));
        h->stage_fk_cap = n;
    }
    double* dj = (double*)h->stage_fk_dev;
    double* dt = dj + 6 * n;
    double* dR = dt + 3 * n;
    int32_t* ds = (int32_t*)(dR + 9 * n);
    int32_t* di = ds + n;
    cudaStream_t st = h->stream;
    CTR_CUDA(cudaMemcpyAsync(dj, joints_host, sizeof(double) * 6 * n, cudaMemcpyHostToDevice, st));
    if (system_host) CTR_CUDA(cudaMemcpyAsync(ds, system_host, sizeof(int32_t) * n, cudaMemcpyHostToDevice, st));
    int rc = ctr_fk(h, dj, system_host ? ds : nullptr, nullptr, n, integrator, precision, dt, dR, di, st);
    if (rc) return rc;
    CTR_CUDA(cudaMemcpyAsync(tip_host, dt, sizeof(double) * 3 * n, cudaMemcpyDeviceToHost, st));
    if (R_host) CTR_CUDA(cudaMemcpyAsync(R_host, dR, sizeof(double) * 9 * n, cudaMemcpyDeviceToHost, st));
    if (info_host) CTR_CUDA(cudaMemcpyAsync(info_host, di, sizeof(int32_t) * 4 * n, cudaMemcpyDeviceToHost, st));
    CTR_CUDA(cudaStreamSynchronize(st));
    return 0;
}

int ctr_set_action(ctr_handle* h, float* joints_dev, const float* action_dev, const int32_t* system_dev, int64_t n,
                   int32_t n_substeps, void* stream) {
    if (!h || !joints_dev || !action_dev) CTR_FAIL("null argument");
    if (n <= 0) return 0;
    CTR_CUDA(cudaSetDevice(h->device));
    k_set_action<<<(int)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(h->sys_dev, h->cfg.n_systems, joints_dev,
                                                                          action_dev, system_dev, n, n_substeps);
    h->launches += 1;
    CTR_CUDA(cudaGetLastError());
    return 0;
}

int ctr_set_state_host(ctr_handle* h, const float* joints, const double* desired_goal, const int32_t* system,
                       const int32_t* t) {
    if (!h) CTR_FAIL("null handle");
    CTR_CUDA(cudaSetDevice(h->device));
    CTR_CUDA(cudaDeviceSynchronize());
    if (joints) CTR_CUDA(cudaMemcpy(h->p.joints, joints, sizeof(float) * 6 * h->n, cudaMemcpyHostToDevice));
    if (desired_goal) CTR_CUDA(cudaMemcpy(h->p.desired_goal, desired_goal, sizeof(double) * 3 * h->n, cudaMemcpyHostToDevice));
    if (system) CTR_CUDA(cudaMemcpy(h->p.system, system, sizeof(int32_t) * h->n, cudaMemcpyHostToDevice));
    if (t) CTR_CUDA(cudaMemcpy(h->p.t, t, sizeof(int32_t) * h->n, cudaMemcpyHostToDevice));
    h->delta_dirty = 1;
    return 0;
}

int ctr_reset(ctr_handle* h, const uint8_t* mask_dev, const double* goal_dev, const int32_t* system_dev, void* stream) {
    if (!h) CTR_FAIL("null handle");
    CTR_CUDA(cudaSetDevice(h->device));
    cudaStream_t st = (cudaStream_t)stream;
    if (mask_dev) {
        CTR_CUDA(cudaMemsetAsync(h->p.reset_count, 0, sizeof(int32_t), st));
        k_mask_to_list<<<(int)((h->n + 255) / 256), 256, 0, st>>>(mask_dev, h->n, h->p.reset_list, h->p.reset_count);
        h->launches += 1;
        CTR_CUDA(cudaGetLastError());
        int rc = launch_reset(h, h->p.reset_list, h->p.reset_count, h->n, goal_dev, system_dev, st);
        return rc ? rc : mark_user_work(h, st);
    }
    // full reset: every env re-queues itself, so drop what the queue holds (keeps it within capacity)
    if (h->prefetch) CTR_CUDA(cudaMemsetAsync(h->p.rq_ctr + (h->step_index % 3), 0, sizeof(int32_t), st));
    int rc = launch_reset(h, nullptr, nullptr, h->n, goal_dev, system_dev, st);
    return rc ? rc : mark_user_work(h, st);
}

int ctr_step(ctr_handle* h, const float* action_dev, void* stream) {
    return ctr_step_strided(h, action_dev, h ? h->n : 0, 1, stream);
}

// one k_step launch over the env range [begin, begin + count)
static int launch_step_range(ctr_handle* h, KArgs a, const float* action_dev, int64_t begin, int64_t count,
                             int refill_blocks, cudaStream_t st) {
    a.env_begin = begin;
    a.env_count = count;
    a.refill_blocks = refill_blocks;
    // U_y == 0 fast path: nominal tubes only (explicit per-env constants may carry any U_y)
    const bool uy0 = h->all_uy_zero && !h->use_params;
#ifndef CTR_DEBUG_SKIP_STEP_KERNEL      // measurement-only build: the host pipeline's copies without its kernels
    int grid = grid_for(count);
    if (a.interleave) {                          // whole windows: the blocks past the last window's end leave at once
        constexpr int BPW = SORT_WINDOW / ENVS_PER_BLOCK;
        grid = ((grid + BPW - 1) / BPW) * BPW;
    }
    DISPATCH_STEP(uy0, h->cfg.integrator, h->cfg.precision, <<<grid + refill_blocks, BLOCK, 0, st>>>(a, action_dev));
#else
    (void)uy0;
#endif
    h->launches += 1;
    CTR_CUDA(cudaGetLastError());
    return 0;
}

int ctr_step_strided(ctr_handle* h, const float* action_dev, int64_t stride_field, int64_t stride_env, void* stream) {
    if (!h || !action_dev) CTR_FAIL("null argument");
    CTR_CUDA(cudaSetDevice(h->device));
    KArgs a = make_args(h);
    a.act_stride_field = stride_field;
    a.act_stride_env = stride_env;
    // leading blocks drain the refill queue of the previous step first (see k_step)
    int rc = launch_step_range(h, a, action_dev, 0, h->n, h->prefetch ? h->refill_blocks : 0, (cudaStream_t)stream);
    h->step_index += 1;
    return rc ? rc : mark_user_work(h, (cudaStream_t)stream);
}

static int copy_out(ctr_handle* h, cudaStream_t st, double* obs, double* ach, double* des, double* rew, uint8_t* done,
                    uint8_t* succ, double* err) {
    const int64_t N = h->n;
    if (obs) CTR_CUDA(cudaMemcpyAsync(obs, h->p.obs, sizeof(double) * h->obs_dim * N, cudaMemcpyDeviceToHost, st));
    if (ach) CTR_CUDA(cudaMemcpyAsync(ach, h->p.achieved, sizeof(double) * 3 * N, cudaMemcpyDeviceToHost, st));
    if (des) CTR_CUDA(cudaMemcpyAsync(des, h->p.desired_goal, sizeof(double) * 3 * N, cudaMemcpyDeviceToHost, st));
    if (rew) CTR_CUDA(cudaMemcpyAsync(rew, h->p.reward, sizeof(double) * N, cudaMemcpyDeviceToHost, st));
    if (done) CTR_CUDA(cudaMemcpyAsync(done, h->p.done, N, cudaMemcpyDeviceToHost, st));
    if (succ) CTR_CUDA(cudaMemcpyAsync(succ, h->p.success, N, cudaMemcpyDeviceToHost, st));
    if (err) CTR_CUDA(cudaMemcpyAsync(err, h->p.error, sizeof(double) * N, cudaMemcpyDeviceToHost, st));
    CTR_CUDA(cudaStreamSynchronize(st));
    return 0;
}

// End-to-end step with HOST buffers, pipelined: the env range is cut into chunks; chunk c runs H2D(actions) -> k_step
// -> D2H(results) on stream c & 1, so the 170 B/env of results of one chunk cross PCIe while the next chunks compute
// (a single launch followed by one copy left the GPU idle for the whole 22 MB transfer).  The ~9 API calls per chunk
// cost more CPU time than the chunk takes on the GPU, so the whole sequence is captured ONCE per set of host pointers
// into a CUDA graph; a step is then: refresh the kernel nodes' arguments (tolerance, queue roles), one graph launch.
static int enqueue_host_step(ctr_handle* h, const KArgs& a, int64_t per, const float* action_host, double* obs_host,
                             double* achieved_host, double* desired_host, double* reward_host, uint8_t* done_host,
                             uint8_t* success_host, double* error_host, int packed) {
    const int64_t N = h->n;
    const size_t D = (size_t)h->obs_dim;
    // kernels alternate between two compute streams (the tail of one chunk overlaps the next chunk); all result copies
    // go to a third stream, each waiting only for its own chunk's kernel
    cudaStream_t s0 = h->streams[0], sc = h->streams[N_COMPUTE_STREAMS];
    CTR_CUDA(cudaEventRecord(h->ev_fork, s0));
    for (int k = 1; k <= N_COMPUTE_STREAMS; ++k) CTR_CUDA(cudaStreamWaitEvent(h->streams[k], h->ev_fork, 0));
    int c = 0;
    for (int64_t b = 0; b < N; b += per, ++c) {
        const int64_t cnt = (N - b < per) ? N - b : per;
        cudaStream_t sk = h->streams[c % N_COMPUTE_STREAMS];
        CTR_CUDA(cudaMemcpy2DAsync(h->stage_action + b, sizeof(float) * N, action_host + b, sizeof(float) * N,
                                   sizeof(float) * cnt, 6, cudaMemcpyHostToDevice, sk));
        int rc = launch_step_range(h, a, h->stage_action, b, cnt, (c == 0 && h->prefetch) ? h->refill_blocks : 0, sk);
        if (rc) return rc;
        CTR_CUDA(cudaEventRecord(h->ev_chunk[c], sk));
        CTR_CUDA(cudaStreamWaitEvent(sc, h->ev_chunk[c], 0));
        cudaStream_t st = sc;
        const size_t pitch = sizeof(double) * N, w = sizeof(double) * cnt;
        if (packed == 2) {               // float32 block: convert the chunk's rows on its compute stream, copy half the bytes
            k_pack_f32<<<(int)((cnt * (int64_t)(D + 8) + 1023) / 1024), 256, 0, sk>>>(h->blk, h->stage_f32, N, b, cnt, (int)(D + 8));
            h->launches += 1;
            CTR_CUDA(cudaGetLastError());
            CTR_CUDA(cudaEventRecord(h->ev_chunk[c], sk));
            CTR_CUDA(cudaStreamWaitEvent(sc, h->ev_chunk[c], 0));
            float* dst = reinterpret_cast<float*>(obs_host);
            CTR_CUDA(cudaMemcpy2DAsync(dst + b, sizeof(float) * N, h->stage_f32 + b, sizeof(float) * N, sizeof(float) * cnt, D + 8,
                                       cudaMemcpyDeviceToHost, st));
        } else if (packed == 3) {        // delta contract: the 17 leading rows per chunk; the list of changed goals follows the last chunk
            CTR_CUDA(cudaMemcpy2DAsync(obs_host + b, pitch, h->blk + b, pitch, w, 17, cudaMemcpyDeviceToHost, st));
        } else if (packed) {
            CTR_CUDA(cudaMemcpy2DAsync(obs_host + b, pitch, h->blk + b, pitch, w, D + 8, cudaMemcpyDeviceToHost, st));
        } else {
            if (obs_host) CTR_CUDA(cudaMemcpy2DAsync(obs_host + b, pitch, h->p.obs + b, pitch, w, D, cudaMemcpyDeviceToHost, st));
            if (achieved_host) CTR_CUDA(cudaMemcpy2DAsync(achieved_host + b, pitch, h->p.achieved + b, pitch, w, 3, cudaMemcpyDeviceToHost, st));
            if (desired_host) CTR_CUDA(cudaMemcpy2DAsync(desired_host + b, pitch, h->p.desired_goal + b, pitch, w, 3, cudaMemcpyDeviceToHost, st));
            if (reward_host) CTR_CUDA(cudaMemcpyAsync(reward_host + b, h->p.reward + b, w, cudaMemcpyDeviceToHost, st));
            if (error_host) CTR_CUDA(cudaMemcpyAsync(error_host + b, h->p.error + b, w, cudaMemcpyDeviceToHost, st));
        }
        if (!h->host_flags_late) {
            if (packed) {
                CTR_CUDA(cudaMemcpy2DAsync(done_host + b, (size_t)N, h->p.done + b, (size_t)N, (size_t)cnt, 2, cudaMemcpyDeviceToHost, st));
            } else {
                if (done_host) CTR_CUDA(cudaMemcpyAsync(done_host + b, h->p.done + b, cnt, cudaMemcpyDeviceToHost, st));
                if (success_host) CTR_CUDA(cudaMemcpyAsync(success_host + b, h->p.success + b, cnt, cudaMemcpyDeviceToHost, st));
            }
        }
    }
    if (packed == 3) {
        // one list for the whole env range, built on the copy stream once every chunk's kernel has finished (the stream has
        // waited for all of them by now): per-chunk lists cost a memset, a launch and a 40 KB copy per chunk, more than the
        // rows they save at N = 1
        CTR_CUDA(cudaMemsetAsync(h->delta_dev, 0, sizeof(double), sc));
        k_delta_list<<<(int)((N + 255) / 256), 256, 0, sc>>>(h->p.done, h->p.desired_goal, h->p.system, N, 0, N, h->cfg.auto_reset,
                                                             h->delta_cap, h->delta_dev);
        h->launches += 1;
        CTR_CUDA(cudaGetLastError());
        CTR_CUDA(cudaMemcpyAsync(h->delta_host, h->delta_dev, sizeof(double) * (1 + 5 * h->delta_cap), cudaMemcpyDeviceToHost, sc));
    }
    if (h->host_flags_late) {
        // the done / success bytes of ALL chunks in one copy behind the last chunk's results (the copy stream has waited for
        // every chunk's kernel by now): a 16 KB copy per chunk costs ~4 us of the copy stream, as much as 200 KB of results
        if (packed) {
            CTR_CUDA(cudaMemcpyAsync(done_host, h->p.done, (size_t)(2 * N), cudaMemcpyDeviceToHost, sc));
        } else {
            if (done_host) CTR_CUDA(cudaMemcpyAsync(done_host, h->p.done, (size_t)N, cudaMemcpyDeviceToHost, sc));
            if (success_host) CTR_CUDA(cudaMemcpyAsync(success_host, h->p.success, (size_t)N, cudaMemcpyDeviceToHost, sc));
        }
    }
    for (int k = 1; k <= N_COMPUTE_STREAMS; ++k) {
        CTR_CUDA(cudaEventRecord(h->ev_join[k], h->streams[k]));
        CTR_CUDA(cudaStreamWaitEvent(s0, h->ev_join[k], 0));
    }
    return 0;
}

static int step_host_impl(ctr_handle* h, const float* action_host, double* obs_host, double* achieved_host,
                          double* desired_host, double* reward_host, uint8_t* done_host, uint8_t* success_host,
                          double* error_host, int packed);

int ctr_step_host(ctr_handle* h, const float* action_host, double* obs_host, double* achieved_host, double* desired_host,
                  double* reward_host, uint8_t* done_host, uint8_t* success_host, double* error_host) {
    return step_host_impl(h, action_host, obs_host, achieved_host, desired_host, reward_host, done_host, success_host,
                          error_host, 0);
}

int ctr_step_host_packed(ctr_handle* h, const float* action_host, double* out_f64_host, uint8_t* out_u8_host) {
    if (!h || !out_f64_host || !out_u8_host) CTR_FAIL("null argument");
    const int64_t N = h->n, D = h->obs_dim;
    (void)D;
    h->delta_dirty = 1;                                          // a full transfer into some block: the delta bookkeeping restarts
    return step_host_impl(h, action_host, out_f64_host, nullptr, nullptr, nullptr, out_u8_host, out_u8_host + N, nullptr, 1);
}

int ctr_step_host_delta(ctr_handle* h, const float* action_host, double* out_f64_host, uint8_t* out_u8_host) {
    if (!h || !out_f64_host || !out_u8_host) CTR_FAIL("null argument");
    CTR_CUDA(cudaSetDevice(h->device));
    const int64_t N = h->n, D = h->obs_dim;
    // a different block, or env state changed by another call since the last delta step: everything is transferred once
    if (h->user_pending) h->delta_dirty = 1;
    if (h->delta_dirty || h->delta_block != (const void*)out_f64_host) {
        int rc = step_host_impl(h, action_host, out_f64_host, nullptr, nullptr, nullptr, out_u8_host, out_u8_host + N, nullptr, 1);
        if (rc) return rc;
        h->delta_block = out_f64_host;
        h->delta_tol = h->tol;
        h->delta_dirty = 0;
        return 0;
    }
    if (!h->delta_dev) {
        h->delta_cap = N / 64 > 1024 ? N / 64 : 1024;            // a steady-state step resets ~0.8 % of the envs; more -> the rows are fetched whole
        const size_t bytes = sizeof(double) * (size_t)(1 + 5 * h->delta_cap);
        CTR_CUDA(cudaMalloc((void**)&h->delta_dev, bytes));
        CTR_CUDA(cudaMallocHost((void**)&h->delta_host, bytes));
    }
    int rc = step_host_impl(h, action_host, out_f64_host, nullptr, nullptr, nullptr, out_u8_host, out_u8_host + N, nullptr, 3);
    if (rc) return rc;
    // host side of the contract: the tolerance row when the tolerance changed, the goals (and system row) of the reset envs
    double* obs = out_f64_host + 5 * N;
    if (h->delta_tol != h->tol) {
        double* row = obs + 12 * N;
        for (int64_t i = 0; i < N; ++i) row[i] = h->tol;
        h->delta_tol = h->tol;
    }
    double* des = out_f64_host + (5 + D) * N;
    {
        const double* list = h->delta_host;
        const int64_t cnt = (int64_t)list[0];
        if (cnt > h->delta_cap) {                                // overflow (a wave of resets): the rows are fetched whole
            CTR_CUDA(cudaMemcpy(des, h->p.desired_goal, sizeof(double) * 3 * N, cudaMemcpyDeviceToHost));
            if (D > 13) CTR_CUDA(cudaMemcpy(obs + 13 * N, h->p.obs + 13 * N, sizeof(double) * N, cudaMemcpyDeviceToHost));
        } else {
            for (int64_t k = 0; k < cnt; ++k) {
                const double* e = list + 1 + 5 * k;
                const int64_t env = (int64_t)e[0];
                des[env] = e[1];
                des[N + env] = e[2];
                des[2 * N + env] = e[3];
                if (D > 13) obs[13 * N + env] = e[4];
            }
        }
    }
    return 0;
}

int ctr_step_host_packed_f32(ctr_handle* h, const float* action_host, float* out_f32_host, uint8_t* out_u8_host) {
    if (!h || !out_f32_host || !out_u8_host) CTR_FAIL("null argument");
    CTR_CUDA(cudaSetDevice(h->device));
    const int64_t N = h->n, D = h->obs_dim;
    if (!h->stage_f32) CTR_CUDA(cudaMalloc((void**)&h->stage_f32, sizeof(float) * (size_t)(D + 8) * N));
    double* base = reinterpret_cast<double*>(out_f32_host);      // only the block's base address is used in mode 2
    return step_host_impl(h, action_host, base, nullptr, nullptr, nullptr, out_u8_host, out_u8_host + N, nullptr, 2);
}

static int step_host_impl(ctr_handle* h, const float* action_host, double* obs_host, double* achieved_host,
                          double* desired_host, double* reward_host, uint8_t* done_host, uint8_t* success_host,
                          double* error_host, int packed) {
    if (!h || !action_host) CTR_FAIL("null argument");
    CTR_CUDA(cudaSetDevice(h->device));
    const int64_t N = h->n;
    if (packed != 3) h->delta_dirty = 1;   // (ctr_step_host_delta clears it again after its own full transfer)
    if (!h->stage_action) CTR_CUDA(cudaMalloc((void**)&h->stage_action, sizeof(float) * 6 * N));
    int64_t chunks = h->host_chunks > 0 ? h->host_chunks : N / 16384;
    if (chunks < 1) chunks = 1;
    if (chunks > 16) chunks = 16;
    int64_t per = (N + chunks - 1) / chunks;
    per = ((per + SORT_WINDOW - 1) / SORT_WINDOW) * SORT_WINDOW;      // chunks start on sorting-window boundaries
    if (packed == 3) h->delta_per = per;
    const KArgs a = make_args(h);
    cudaStream_t s0 = h->streams[0];
    {   // env state written by device-pointer calls on the caller's stream (reset, step, tube parameters) comes first
        int rc = order_after_user(h, s0);
        if (rc) return rc;
    }
    if (!h->host_graphs) {
        int rc = enqueue_host_step(h, a, per, action_host, obs_host, achieved_host, desired_host, reward_host, done_host,
                                   success_host, error_host, packed);
        if (rc) return rc;
        h->step_index += 1;
        CTR_CUDA(cudaStreamSynchronize(s0));
        return 0;
    }
    const void* key[8] = {action_host, obs_host, achieved_host, desired_host, reward_host, done_host, success_host, error_host};
    // the captured kernel nodes keep the k_step instantiation chosen at capture time (launch_step_range): a capture made
    // before ctr_set_tube_params() switched per-env constants on (or off) must not be replayed afterwards
    const bool uy0 = h->all_uy_zero && !h->use_params;
    HostGraph* g = nullptr;
    for (auto& c : h->graphs)
        if (c.per == per && c.packed == packed && c.uy0 == uy0 && memcmp(c.key, key, sizeof key) == 0) g = &c;
    if (!g) {
        if (h->graphs.size() >= 8) {                 // evict the least recently used capture
            size_t old = 0;
            for (size_t i = 1; i < h->graphs.size(); ++i)
                if (h->graphs[i].stamp < h->graphs[old].stamp) old = i;
            cudaGraphExecDestroy(h->graphs[old].exec);
            cudaGraphDestroy(h->graphs[old].graph);
            h->graphs.erase(h->graphs.begin() + old);
        }
        HostGraph ng;
        memcpy(ng.key, key, sizeof key);
        ng.per = per;
        ng.packed = packed;
        ng.uy0 = uy0;
        ng.stamp = 0;
        cudaGraph_t graph = nullptr;
        const int64_t launches0 = h->launches;
        CTR_CUDA(cudaStreamBeginCapture(s0, cudaStreamCaptureModeThreadLocal));
        int rc = enqueue_host_step(h, a, per, action_host, obs_host, achieved_host, desired_host, reward_host, done_host,
                                   success_host, error_host, packed);
        cudaError_t ce = cudaStreamEndCapture(s0, &graph);
        h->launches = launches0;
        if (rc) return rc;
        if (ce != cudaSuccess) CTR_FAIL(std::string("stream capture failed: ") + cudaGetErrorString(ce));
        size_t nn = 0;
        CTR_CUDA(cudaGraphGetNodes(graph, nullptr, &nn));
        std::vector<cudaGraphNode_t> nodes(nn);
        CTR_CUDA(cudaGraphGetNodes(graph, nodes.data(), &nn));
        for (cudaGraphNode_t nd : nodes) {
            cudaGraphNodeType ty;
            CTR_CUDA(cudaGraphNodeGetType(nd, &ty));
            if (ty != cudaGraphNodeTypeKernel) continue;
            cudaKernelNodeParams kp;
            CTR_CUDA(cudaGraphKernelNodeGetParams(nd, &kp));
            if (kp.func == (void*)k_pack_f32 || kp.func == (void*)k_delta_list) continue;   // only the k_step nodes carry per-step arguments
            const KArgs* ka = (const KArgs*)kp.kernelParams[0];
            ng.knodes.push_back(nd);
            ng.kparams.push_back(kp);
            ng.begin.push_back(ka->env_begin);
            ng.count.push_back(ka->env_count);
            ng.refill.push_back(ka->refill_blocks);
        }
        CTR_CUDA(cudaGraphInstantiate(&ng.exec, graph, 0));
        ng.graph = graph;                            // kept: the node handles stay valid with it
        h->graphs.push_back(ng);
        g = &h->graphs.back();
    }
    g->stamp = ++h->graph_clock;
#ifdef CTR_HOST_TIMING
    static double t_set = 0, t_launch = 0, t_sync = 0; static int t_n = 0;
    auto T0 = std::chrono::steady_clock::now();
#endif
    // refresh the per-step kernel arguments (goal tolerance, refill-queue roles, seed) of every chunk's k_step node
    const float* act_dev = h->stage_action;
    for (size_t i = 0; i < g->knodes.size(); ++i) {
        KArgs ai = a;
        ai.env_begin = g->begin[i];
        ai.env_count = g->count[i];
        ai.refill_blocks = g->refill[i];
        void* params[2] = {&ai, (void*)&act_dev};
        cudaKernelNodeParams kp = g->kparams[i];
        kp.kernelParams = params;
        kp.extra = nullptr;
        CTR_CUDA(cudaGraphExecKernelNodeSetParams(g->exec, g->knodes[i], &kp));
    }
#ifdef CTR_HOST_TIMING
    auto T1 = std::chrono::steady_clock::now();
#endif
    CTR_CUDA(cudaGraphLaunch(g->exec, s0));
#ifdef CTR_HOST_TIMING
    auto T2 = std::chrono::steady_clock::now();
#endif
    h->launches += (int64_t)g->knodes.size();
    h->step_index += 1;
    CTR_CUDA(cudaStreamSynchronize(s0));
#ifdef CTR_HOST_TIMING
    auto T3 = std::chrono::steady_clock::now();
    t_set += std::chrono::duration<double, std::micro>(T1 - T0).count();
    t_launch += std::chrono::duration<double, std::micro>(T2 - T1).count();
    t_sync += std::chrono::duration<double, std::micro>(T3 - T2).count();
    if (++t_n % 20 == 0) { fprintf(stderr, "[host step] set-params %.1f us, graph launch %.1f us, sync wait %.1f us (mean of 20)\n", t_set / 20, t_launch / 20, t_sync / 20); t_set = t_launch = t_sync = 0; }
#endif
    return 0;
}

int ctr_reset_host(ctr_handle* h, double* obs_host, double* achieved_host, double* desired_host) {
    if (!h) CTR_FAIL("null handle");
    int rc = order_after_user(h, h->stream);
    if (rc) return rc;
    rc = ctr_reset(h, nullptr, nullptr, nullptr, h->stream);
    if (rc) return rc;
    return copy_out(h, h->stream, obs_host, achieved_host, desired_host, nullptr, nullptr, nullptr, nullptr);
}

int ctr_compute_reward(ctr_handle* h, const double* achieved_dev, const double* desired_dev, int64_t m, double tol,
                       int32_t mode, double* reward_dev, void* stream) {
    return ctr_relabel(h, achieved_dev, desired_dev, m, tol, mode, reward_dev, nullptr, stream);
}

int ctr_relabel(ctr_handle* h, const double* achieved_dev, const double* new_desired_dev, int64_t m, double tol,
                int32_t mode, double* reward_dev, double* obs_dev, void* stream) {
    if (!h || !achieved_dev || !new_desired_dev || !reward_dev) CTR_FAIL("null argument");
    if (m <= 0) return 0;
    CTR_CUDA(cudaSetDevice(h->device));
    if (tol < 0) tol = h->tol;
    if (mode < 0) mode = h->cfg.reward_mode;
    int64_t blocks = (m + 255) / 256;
    int64_t cap = (int64_t)h->sm_count * 16;
    if (blocks > cap) blocks = cap;
    k_reward<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(achieved_dev, new_desired_dev, m, tol, mode, reward_dev, obs_dev);
    h->launches += 1;
    CTR_CUDA(cudaGetLastError());
    return 0;
}

int ctr_stats(ctr_handle* h, double out8_host[8], void* stream) {
    if (!h || !out8_host) CTR_FAIL("null argument");
    CTR_CUDA(cudaSetDevice(h->device));
    double tmp[N_STATS];
    CTR_CUDA(cudaMemcpyAsync(tmp, h->p.stats, sizeof(tmp), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    CTR_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    for (int i = 0; i < 8; ++i) out8_host[i] = tmp[i];
    if (tmp[ST_FK_FAILED] != 0.0 || tmp[ST_GOAL_STUCK] != 0.0) {
        char buf[160];
        snprintf(buf, sizeof buf, "%g forward-kinematics solves failed (NaN / step too small, model.py:69), %g goal samplings stuck (obs.py:205)",
                 tmp[ST_FK_FAILED], tmp[ST_GOAL_STUCK]);
        g_err = buf;
        return 3;
    }
    return 0;
}

int ctr_work_counters(ctr_handle* h, double out4_host[4], void* stream) {
    if (!h || !out4_host) CTR_FAIL("null argument");
    CTR_CUDA(cudaSetDevice(h->device));
    double tmp[N_STATS];
    CTR_CUDA(cudaMemcpyAsync(tmp, h->p.stats, sizeof(tmp), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    CTR_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    out4_host[0] = tmp[ST_NFEV];
    out4_host[1] = tmp[ST_ATTEMPTS];
    out4_host[2] = tmp[ST_SECTIONS];
    out4_host[3] = tmp[ST_FK_COUNT];
    return 0;
}

// ---- checkpoint / restore (joints, goals, t, system, tube parameters, RNG position, tolerance -- here: everything)
// A checkpoint is the raw content of every device buffer the handle owns (state, result block, prefetch slots with their tags,
// refill queues, slot permutation, statistics) plus the host-side scalars; it restores into a handle created with the same
// configuration and env count, after which the two handles produce identical bits for identical actions.
struct CkptHeader {
    uint64_t magic;
    int64_t n, step_index, n_allocs;
    double tol;
    uint64_t seed;
    int32_t use_params, obs_dim, abi, pad;
};
static const uint64_t CKPT_MAGIC = 0x43545252424b5032ULL;   // "CTRRBKP2"

int64_t ctr_checkpoint_bytes(const ctr_handle* h) {
    if (!h) return -1;
    size_t total = sizeof(CkptHeader) + sizeof(uint64_t) * h->alloc_bytes.size();
    for (size_t b : h->alloc_bytes) total += b;
    return (int64_t)total;
}

int ctr_checkpoint_save(ctr_handle* h, void* host_buf, int64_t nbytes) {
    if (!h || !host_buf) CTR_FAIL("null argument");
    if (nbytes < ctr_checkpoint_bytes(h)) CTR_FAIL("buffer smaller than ctr_checkpoint_bytes()");
    CTR_CUDA(cudaSetDevice(h->device));
    CTR_CUDA(cudaDeviceSynchronize());
    char* p = (char*)host_buf;
    CkptHeader hd;
    memset(&hd, 0, sizeof hd);
    hd.magic = CKPT_MAGIC;
    hd.n = h->n;
    hd.step_index = h->step_index;
    hd.n_allocs = (int64_t)h->allocs.size();
    hd.tol = h->tol;
    hd.seed = h->cfg.seed;
    hd.use_params = h->use_params;
    hd.obs_dim = h->obs_dim;
    hd.abi = CTR_ABI_VERSION;
    memcpy(p, &hd, sizeof hd);
    p += sizeof hd;
    for (size_t b : h->alloc_bytes) {
        const uint64_t v = b;
        memcpy(p, &v, sizeof v);
        p += sizeof v;
    }
    for (size_t k = 0; k < h->allocs.size(); ++k) {
        CTR_CUDA(cudaMemcpy(p, h->allocs[k], h->alloc_bytes[k], cudaMemcpyDeviceToHost));
        p += h->alloc_bytes[k];
    }
    return 0;
}

int ctr_checkpoint_load(ctr_handle* h, const void* host_buf, int64_t nbytes) {
    if (!h || !host_buf) CTR_FAIL("null argument");
    if (nbytes < (int64_t)sizeof(CkptHeader)) CTR_FAIL("truncated checkpoint");
    const char* p = (const char*)host_buf;
    CkptHeader hd;
    memcpy(&hd, p, sizeof hd);
    p += sizeof hd;
    if (hd.magic != CKPT_MAGIC || hd.abi != CTR_ABI_VERSION) CTR_FAIL("not a checkpoint of this library version");
    if (hd.n != h->n || hd.obs_dim != h->obs_dim || hd.n_allocs != (int64_t)h->allocs.size())
        CTR_FAIL("checkpoint was taken from a handle with a different env count / configuration");
    if (nbytes < ctr_checkpoint_bytes(h)) CTR_FAIL("truncated checkpoint");
    for (size_t k = 0; k < h->allocs.size(); ++k) {
        uint64_t v;
        memcpy(&v, p, sizeof v);
        p += sizeof v;
        if (v != h->alloc_bytes[k]) CTR_FAIL("checkpoint buffer layout differs from this handle's (different configuration)");
    }
    CTR_CUDA(cudaSetDevice(h->device));
    CTR_CUDA(cudaDeviceSynchronize());
    for (size_t k = 0; k < h->allocs.size(); ++k) {
        if (h->allocs[k] == (void*)h->sys_dev || h->allocs[k] == (void*)h->raw_dev) {   // constants of THIS handle's configuration stay
            p += h->alloc_bytes[k];
            continue;
        }
        CTR_CUDA(cudaMemcpy(h->allocs[k], p, h->alloc_bytes[k], cudaMemcpyHostToDevice));
        p += h->alloc_bytes[k];
    }
    h->step_index = hd.step_index;
    h->tol = hd.tol;
    h->cfg.seed = hd.seed;
    h->use_params = hd.use_params;
    h->user_pending = 0;
    h->delta_dirty = 1;
    return 0;
}

int ctr_segment_tables(ctr_handle* h, const double* joints_dev, const int32_t* system_dev, const double* params_dev, int64_t n,
                       double* t0_dev, double* tf_dev, int32_t* pattern_dev, int32_t* count_dev, void* stream) {
    if (!h || !joints_dev || !t0_dev || !tf_dev || !pattern_dev || !count_dev) CTR_FAIL("null argument");
    if (n <= 0) return 0;
    CTR_CUDA(cudaSetDevice(h->device));
    k_segment_tables<<<grid_for(n), BLOCK, 0, (cudaStream_t)stream>>>(h->sys_dev, h->cfg.n_systems, joints_dev, system_dev, params_dev, n,
                                                                      t0_dev, tf_dev, pattern_dev, count_dev);
    h->launches += 1;
    CTR_CUDA(cudaGetLastError());
    return 0;
}

int ctr_debug_buffer(ctr_handle* h, int32_t which, void** dev_ptr, int64_t* count) {
    if (!h || !dev_ptr || !count) CTR_FAIL("null argument");
    *dev_ptr = nullptr;
    *count = 0;
    switch (which) {
        case CTR_DBG_PERM: *dev_ptr = h->perm; *count = h->perm ? h->n : 0; break;
        case CTR_DBG_WINDOW_COUNTERS: *dev_ptr = h->win_count; *count = h->win_count ? (h->n + SORT_WINDOW - 1) / SORT_WINDOW : 0; break;
        case CTR_DBG_QUEUE_COUNTERS: *dev_ptr = h->prefetch ? h->p.rq_ctr : nullptr; *count = h->prefetch ? 6 : 0; break;
        case CTR_DBG_QUEUE_0: case CTR_DBG_QUEUE_1: case CTR_DBG_QUEUE_2:
            *dev_ptr = h->prefetch ? h->p.rq_list[which - CTR_DBG_QUEUE_0] : nullptr; *count = h->prefetch ? h->rq_cap : 0; break;
        case CTR_DBG_SLOT_TAGS: *dev_ptr = h->prefetch ? h->p.pf_episode : nullptr; *count = h->prefetch ? h->n : 0; break;
        case CTR_DBG_STEP_INDEX: *dev_ptr = nullptr; *count = h->step_index; break;
        default: CTR_FAIL("unknown debug buffer");
    }
    return 0;
}

int ctr_fault_counters(ctr_handle* h, double out4_host[4], void* stream) {
    if (!h || !out4_host) CTR_FAIL("null argument");
    CTR_CUDA(cudaSetDevice(h->device));
    double tmp[N_STATS];
    CTR_CUDA(cudaMemcpyAsync(tmp, h->p.stats, sizeof(tmp), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    CTR_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    out4_host[0] = tmp[ST_FK_FAILED];
    out4_host[1] = tmp[ST_GOAL_STUCK];
    out4_host[2] = tmp[ST_INVALID_ACTION];
    out4_host[3] = 0.0;
    return 0;
}

int ctr_measure_fma_peak(int32_t device, int32_t precision, double* tflops_out) {
    if (!tflops_out) CTR_FAIL("null argument");
    CTR_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop;
    CTR_CUDA(cudaGetDeviceProperties(&prop, device));
    const int blocks = prop.multiProcessorCount * 8, threads = 256, iters = 1 << 14;
    void* out = nullptr;
    CTR_CUDA(cudaMalloc(&out, (size_t)blocks * threads * 8));
    cudaEvent_t e0, e1;
    CTR_CUDA(cudaEventCreate(&e0));
    CTR_CUDA(cudaEventCreate(&e1));
    float best = 1e30f;
    for (int rep = 0; rep < 5; ++rep) {
        CTR_CUDA(cudaEventRecord(e0, 0));
        if (precision == CTR_PRECISION_F64) k_fma_peak<double><<<blocks, threads>>>((double*)out, iters, 0.999999, 1e-7);
        else k_fma_peak<float><<<blocks, threads>>>((float*)out, iters, 0.999999f, 1e-7f);
        CTR_CUDA(cudaEventRecord(e1, 0));
        CTR_CUDA(cudaEventSynchronize(e1));
        float ms = 0;
        CTR_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        if (rep > 0 && ms < best) best = ms;
    }
    CTR_CUDA(cudaGetLastError());
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(out);
    *tflops_out = 2.0 * 8.0 * (double)iters * (double)blocks * threads / (best * 1e-3) / 1e12;
    return 0;
}

int ctr_stats_reset(ctr_handle* h, void* stream) {
    if (!h) CTR_FAIL("null handle");
    CTR_CUDA(cudaSetDevice(h->device));
    CTR_CUDA(cudaMemsetAsync(h->p.stats, 0, sizeof(double) * N_STATS, (cudaStream_t)stream));
    return 0;
}

// ---- hindsight replay (kernels in ctr_her.cuh) ------------------------------------------------
static int her_check(const ctr_handle* h, const ctr_her_ring* r) {
    if (!h || !r) CTR_FAIL("null argument");
    if (r->n_envs != h->n) CTR_FAIL("ring.n_envs differs from the handle's environment count");
    if (r->capacity < 1) CTR_FAIL("ring.capacity must be >= 1");
    if (r->obs_dim != h->obs_dim) CTR_FAIL("ring.obs_dim differs from ctr_obs_dim()");
    if (!r->rec || !r->done || !r->t_in_ep || !r->ep_len || !r->cur_obs || !r->cur_achieved || !r->cur_desired || !r->cur_t)
        CTR_FAIL("ring has a null buffer");
    if (((uintptr_t)r->rec & 31) != 0) CTR_FAIL("ring.rec must be 32-byte aligned");
    if (2 * r->obs_dim + 13 > CTR_HER_RECORD) CTR_FAIL("observation does not fit a replay record");
    return 0;
}
static HerSrc her_src(const ctr_handle* h) {
    HerSrc s;
    s.obs = h->p.obs;
    s.achieved = h->p.achieved;
    s.desired = h->p.desired_goal;
    s.terminal_obs = h->p.terminal_obs;
    s.terminal_achieved = h->p.terminal_achieved;
    s.reward = h->p.reward;
    s.done = h->p.done;
    s.auto_reset = h->cfg.auto_reset;
    return s;
}

int ctr_her_begin(ctr_handle* h, const ctr_her_ring* ring, void* stream) {
    int rc = her_check(h, ring);
    if (rc) return rc;
    CTR_CUDA(cudaSetDevice(h->device));
    k_her_begin<<<(int)((h->n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(*ring, her_src(h));
    h->launches += 1;
    CTR_CUDA(cudaGetLastError());
    return 0;
}

int ctr_her_store(ctr_handle* h, const ctr_her_ring* ring, const float* action_dev, int64_t stride_field,
                  int64_t stride_env, int64_t slot, void* stream) {
    int rc = her_check(h, ring);
    if (rc) return rc;
    if (!action_dev) CTR_FAIL("null action");
    if (slot < 0 || slot >= ring->capacity) CTR_FAIL("slot out of range");
    CTR_CUDA(cudaSetDevice(h->device));
    k_her_store<<<(int)((h->n + 32 * HER_WARPS - 1) / (32 * HER_WARPS)), 32 * HER_WARPS, 0, (cudaStream_t)stream>>>(
        *ring, her_src(h), action_dev, stride_field, stride_env, slot);
    h->launches += 1;
    CTR_CUDA(cudaGetLastError());
    return 0;
}

int ctr_her_sample(ctr_handle* h, const ctr_her_ring* ring, int64_t n_stored, int64_t newest_slot, int32_t n_sampled_goal,
                   int32_t rewrite_error, double tol, uint64_t seed, uint64_t call, const ctr_her_batch* b, void* stream) {
    int rc = her_check(h, ring);
    if (rc) return rc;
    if (!b) CTR_FAIL("null batch");
    if (b->m <= 0) return 0;
    if (n_stored < 1 || n_stored > ring->capacity) CTR_FAIL("n_stored must be in [1, capacity]");
    if (newest_slot < 0 || newest_slot >= ring->capacity) CTR_FAIL("newest_slot out of range");
    if (n_sampled_goal < 0) CTR_FAIL("n_sampled_goal must be >= 0");
    if (!b->obs || !b->next_obs || !b->achieved || !b->next_achieved || !b->desired || !b->action || !b->reward ||
        !b->done || !b->src || !b->goal_src)
        CTR_FAIL("batch has a null buffer");
    CTR_CUDA(cudaSetDevice(h->device));
    if (tol < 0) tol = h->tol;
    k_her_sample<<<(int)((b->m + 32 * HER_WARPS - 1) / (32 * HER_WARPS)), 32 * HER_WARPS, 0, (cudaStream_t)stream>>>(
        *ring, *b, n_stored, newest_slot, n_sampled_goal, rewrite_error, tol, h->cfg.reward_mode,
        make_uint2((unsigned)seed, (unsigned)(seed >> 32)), make_uint2((unsigned)call, (unsigned)(call >> 32)));
    h->launches += 1;
    CTR_CUDA(cudaGetLastError());
    return 0;
}

// ---- backbone shape ---------------------------------------------------------------------------
int ctr_fk_shape(ctr_handle* h, const double* joints_dev, const int32_t* system_dev, const double* params_dev, int64_t n,
                 int32_t max_pts, double* shape_dev, int32_t* npts_dev, int32_t* tip_idx_dev, double* tip_dev, void* stream) {
    if (!h || !joints_dev || !shape_dev || !npts_dev || !tip_idx_dev) CTR_FAIL("null argument");
    if (max_pts < 1) CTR_FAIL("max_pts must be >= 1");
    if (n <= 0) return 0;
    CTR_CUDA(cudaSetDevice(h->device));
    const int32_t* order = nullptr;              // work-sorted hand-out order, as for ctr_fk
    {
        int rc = build_order(h, joints_dev, system_dev, params_dev, n, 1, ORDER_MIN_FK_SOLVES / 2, (cudaStream_t)stream, &order);
        if (rc) return rc;
    }
    if (h->all_uy_zero && !params_dev)
        k_fk_shape<true><<<grid_for(n), BLOCK, 0, (cudaStream_t)stream>>>(h->sys_dev, h->cfg.n_systems, joints_dev, system_dev, params_dev, n,
                                                                         h->cfg.rtol, h->cfg.atol, max_pts, shape_dev, npts_dev,
                                                                         tip_idx_dev, tip_dev, h->p.stats, order);
    else
        k_fk_shape<false><<<grid_for(n), BLOCK, 0, (cudaStream_t)stream>>>(h->sys_dev, h->cfg.n_systems, joints_dev, system_dev, params_dev, n,
                                                                          h->cfg.rtol, h->cfg.atol, max_pts, shape_dev, npts_dev,
                                                                          tip_idx_dev, tip_dev, h->p.stats, order);
    h->launches += 1;
    CTR_CUDA(cudaGetLastError());
    return 0;
}

// ---- finite-difference Jacobian / DLS inverse kinematics (kernels in ctr_ik.cuh) ----------------------
struct IkScratch {
    double* block;     // f64 [6][7n]
    double* tips;      // f64 [3][7n]
    int32_t* system;   // i32 [7n]
    int32_t* active;   // i32 [1]
    uint8_t* frozen;   // u8  [n]
};
static int ik_scratch(ctr_handle* h, int64_t n, IkScratch* out) {
    if (h->ik_cap < n) {
        if (h->ik_dev) CTR_CUDA(cudaFree(h->ik_dev));
        h->ik_dev = nullptr;
        h->ik_cap = 0;
        CTR_CUDA(cudaMalloc(&h->ik_dev, (size_t)n * (7 * 6 * 8 + 7 * 3 * 8 + 7 * 4 + 1) + 64));
        h->ik_cap = n;
    }
    const int64_t c = h->ik_cap;
    out->block = (double*)h->ik_dev;
    out->tips = out->block + 42 * c;
    out->system = (int32_t*)(out->tips + 21 * c);
    out->active = out->system + 7 * c;
    out->frozen = (uint8_t*)(out->active + 16);
    return 0;
}
// the 7n kinematics solves of one Jacobian: unperturbed + one column block per joint
static int ik_solve7(ctr_handle* h, const IkScratch& sc, const double* q_dev, const int32_t* system_dev, int64_t n, double eps,
                     cudaStream_t st, bool sorted = true) {
    k_ik_perturb<<<(int)((n + 255) / 256), 256, 0, st>>>(q_dev, system_dev, n, eps, sc.block, system_dev ? sc.system : nullptr);
    h->launches += 1;
    CTR_CUDA(cudaGetLastError());
    const int32_t* order = nullptr;              // configurations in work-sorted order, the 7 columns of each one together
    int rc = sorted ? build_order(h, q_dev, system_dev, nullptr, n, 1, (ORDER_MIN_FK_SOLVES + 6) / 7, st, &order) : 0;   // (the order buffer is a cudaMalloc: not for ctr_jacobian_noalloc)
    if (rc) return rc;
    return fk_launch(h, sc.block, system_dev ? sc.system : nullptr, nullptr, 7 * n, CTR_INTEGRATOR_DP45, CTR_PRECISION_F64, sc.tips,
                     nullptr, nullptr, st, order, 7);
}
static int ik_systems(const IkScratch&, const int32_t*, int64_t, cudaStream_t) {
    return 0;                                    // the system ids of the 7n columns are written by k_ik_perturb (same layout as the joints)
}

int ctr_jacobian(ctr_handle* h, const double* joints_dev, const int32_t* system_dev, int64_t n, double eps, double* J_dev,
                 double* tip_dev, void* stream) {
    if (!h || !joints_dev || !J_dev) CTR_FAIL("null argument");
    if (!(eps > 0.0)) CTR_FAIL("eps must be > 0");
    if (n <= 0) return 0;
    CTR_CUDA(cudaSetDevice(h->device));
    cudaStream_t st = (cudaStream_t)stream;
    IkScratch sc;
    int rc = ik_scratch(h, n, &sc);
    if (!rc) rc = ik_systems(sc, system_dev, n, st);
    if (!rc) rc = ik_solve7(h, sc, joints_dev, system_dev, n, eps, st);
    if (rc) return rc;
    k_ik_jacobian<<<(int)((n + 255) / 256), 256, 0, st>>>(sc.tips, n, eps, J_dev, tip_dev);
    h->launches += 1;
    CTR_CUDA(cudaGetLastError());
    return 0;
}

// ---- ctr_jacobian_noalloc: ctr_jacobian for callers that cannot call cudaMalloc (graph capture, serving loops) ---------------
// No counterpart in the reference.  The 7n perturbed joint columns, their 7n
// tips and the 7n system ids live in ONE file-scope device array instead of the handle's grow-on-demand allocation.  Every
// GPU that loads the library holds the array (CTR_JACOBIAN_NOALLOC_MAX * 532 B = 8.3 MiB) until the process exits.  Every
// call on a device shares it: calls that overlap on the GPU (different streams, different handles of one device) race.
// Nothing in it has to start at zero -- k_ik_perturb writes all 42n joint values, it also writes all 7n system ids, and k_fk all 21n
// tip values (NaN for a failed solve) before k_ik_jacobian reads them -- so the previous call's contents are never seen.
constexpr int64_t NOALLOC_BYTES_PER_CONFIG = 7 * 6 * 8 + 7 * 3 * 8 + 7 * 4;
static __device__ __align__(256) unsigned char g_jacobian_scratch[CTR_JACOBIAN_NOALLOC_MAX * NOALLOC_BYTES_PER_CONFIG];
// In host code `g_jacobian_scratch` is only a stand-in for the symbol: its device address is resolved once per device, with
// that device current (ctr_create does it at set-up; the entry point itself for a device no handle was created on).
constexpr int NOALLOC_MAX_DEVICES = 64;
static void* g_jacobian_scratch_addr[NOALLOC_MAX_DEVICES];
static cudaError_t g_jacobian_scratch_err[NOALLOC_MAX_DEVICES];
static bool g_jacobian_scratch_known[NOALLOC_MAX_DEVICES];
static std::mutex g_jacobian_scratch_mutex;
static cudaError_t jacobian_scratch_resolve(int device, void** out) {      // the caller has made `device` current
    if (device < 0 || device >= NOALLOC_MAX_DEVICES) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(g_jacobian_scratch_mutex);
    if (!g_jacobian_scratch_known[device]) {
        void* ptr = nullptr;
        g_jacobian_scratch_err[device] = cudaGetSymbolAddress((void**)&ptr, g_jacobian_scratch);
        g_jacobian_scratch_addr[device] = g_jacobian_scratch_err[device] == cudaSuccess ? ptr : nullptr;
        g_jacobian_scratch_known[device] = true;
    }
    if (out) *out = g_jacobian_scratch_addr[device];
    return g_jacobian_scratch_err[device];
}

int ctr_jacobian_noalloc(ctr_handle* h, const double* joints_dev, const int32_t* system_dev, int64_t n, double eps, double* J_dev,
                         double* tip_dev, void* stream) {
    if (!h || !joints_dev || !J_dev) CTR_FAIL("null argument");
    if (!(eps > 0.0)) CTR_FAIL("eps must be > 0");
    if (n <= 0) return 0;
    if (n > CTR_JACOBIAN_NOALLOC_MAX)
        CTR_FAIL("ctr_jacobian_noalloc: n exceeds CTR_JACOBIAN_NOALLOC_MAX (16384 configurations, the size of the static device "
                 "scratch array); split the batch or call ctr_jacobian");
    CTR_CUDA(cudaSetDevice(h->device));
    void* base = nullptr;
    const cudaError_t err = jacobian_scratch_resolve(h->device, &base);
    if (err != cudaSuccess || !base) {
        g_err = std::string("ctr_jacobian_noalloc: cudaGetSymbolAddress of the static scratch array failed: ") + cudaGetErrorString(err);
        return 2;
    }
    cudaStream_t st = (cudaStream_t)stream;
    IkScratch sc;                                   // compact layout for this n: joints f64 [6][7n] | tips f64 [3][7n] | system i32 [7n]
    sc.block = (double*)base;
    sc.tips = sc.block + 42 * n;
    sc.system = (int32_t*)(sc.tips + 21 * n);
    sc.active = nullptr;
    sc.frozen = nullptr;
    int rc = ik_systems(sc, system_dev, n, st);
    if (!rc) rc = ik_solve7(h, sc, joints_dev, system_dev, n, eps, st, false);   // given order: no order buffer, no allocation
    if (rc) return rc;
    k_ik_jacobian<<<(int)((n + 255) / 256), 256, 0, st>>>(sc.tips, n, eps, J_dev, tip_dev);
    h->launches += 1;
    CTR_CUDA(cudaGetLastError());
    return 0;
}

int ctr_ik_dls(ctr_handle* h, const double* target_dev, double* joints_dev, const int32_t* system_dev, int64_t n, double lam,
               double eps, double tol, int32_t max_iters, double* err_dev, int32_t* iters_dev, void* stream) {
    if (!h || !target_dev || !joints_dev || !err_dev || !iters_dev) CTR_FAIL("null argument");
    if (!(eps > 0.0) || !(lam >= 0.0) || max_iters < 0) CTR_FAIL("bad eps / lam / max_iters");
    if (n <= 0) return 0;
    CTR_CUDA(cudaSetDevice(h->device));
    cudaStream_t st = (cudaStream_t)stream;
    IkScratch sc;
    int rc = ik_scratch(h, n, &sc);
    if (!rc) rc = ik_systems(sc, system_dev, n, st);
    if (rc) return rc;
    CTR_CUDA(cudaMemsetAsync(sc.frozen, 0, (size_t)n, st));
    CTR_CUDA(cudaMemsetAsync(iters_dev, 0, sizeof(int32_t) * n, st));
    CTR_CUDA(cudaMemsetAsync(err_dev, 0, sizeof(double) * n, st));
    for (int it = 0; it < max_iters; ++it) {
        CTR_CUDA(cudaMemsetAsync(sc.active, 0, sizeof(int32_t), st));
        rc = ik_solve7(h, sc, joints_dev, system_dev, n, eps, st);
        if (rc) return rc;
        k_ik_dls_update<true><<<(int)((n + 255) / 256), 256, 0, st>>>(sc.tips, target_dev, n, eps, lam, tol, joints_dev, err_dev, iters_dev,
                                                               sc.frozen, sc.active);
        h->launches += 1;
        CTR_CUDA(cudaGetLastError());
        if ((it & 3) == 3 || it == max_iters - 1) {          // every 4th iteration: has every configuration converged?
            int32_t active = 0;
            CTR_CUDA(cudaMemcpyAsync(&active, sc.active, sizeof(int32_t), cudaMemcpyDeviceToHost, st));
            CTR_CUDA(cudaStreamSynchronize(st));
            if (active == 0) break;
        }
    }
    CTR_CUDA(cudaStreamSynchronize(st));
    return 0;
}


// ---- BVP back-end (csrc/ctr_bvp.cuh; CTR_Python/CTR_Model.py) --------------------------------------------------------
static int bvp_scratch(ctr_handle* h, size_t bytes, char** out) {
    if (h->bvp_cap < bytes) {
        if (h->bvp_dev) CTR_CUDA(cudaFree(h->bvp_dev));
        h->bvp_dev = nullptr;
        h->bvp_cap = 0;
        CTR_CUDA(cudaMalloc(&h->bvp_dev, bytes));
        h->bvp_cap = bytes;
    }
    *out = (char*)h->bvp_dev;
    return 0;
}
static BvpArgs bvp_args(ctr_handle* h, const double* q, const double* f, const double* u, const int32_t* system,
                        const double* params, int64_t n, int32_t mode, double max_step) {
    BvpArgs a;
    memset(&a, 0, sizeof a);
    a.sys = h->sys_dev;
    a.n_systems = h->cfg.n_systems;
    a.q = q;
    a.f = f;
    a.u = u;
    a.system = system;
    a.params = params;
    a.n = n;
    a.mode = mode;
    a.eps_q = 1e-4;                  // CTR_Model.py:17
    a.eps_f = 0.05;                  // CTR_Model.py:270
    a.eps_u = 1e-4;                  // the reference's own perturbation for its (unused) shooting Jacobian, CTR_Model.py:17,243
    a.max_step = max_step;
    a.rtol = 1e-3;                   // solve_ivp defaults (no tolerances passed, CTR_Model.py:164-167)
    a.atol = 1e-6;
    a.fault = h->p.stats + ST_FK_FAILED;
    return a;
}
static int build_order(ctr_handle* h, const double* q_dev, const int32_t* system_dev, const double* params_dev, int64_t n,
                       int many_first, int64_t min_n, cudaStream_t st, const int32_t** order_out) {
    *order_out = nullptr;
    if (!h->bvp_sort || n < min_n || n >= (int64_t)1 << 31) return 0;
    if (h->bvp_order_cap < n) {
        if (h->bvp_order_dev) CTR_CUDA(cudaFree(h->bvp_order_dev));
        h->bvp_order_dev = nullptr;
        h->bvp_order_cap = 0;
        CTR_CUDA(cudaMalloc(&h->bvp_order_dev, (size_t)n * 5 + 1024 + 64));
        h->bvp_order_cap = n;
    }
    int32_t* order = (int32_t*)h->bvp_order_dev;
    int32_t* hist = order + h->bvp_order_cap;
    uint8_t* key = (uint8_t*)(hist + 256);
    BvpArgs a;
    memset(&a, 0, sizeof a);
    a.sys = h->sys_dev;
    a.n_systems = h->cfg.n_systems;
    a.q = q_dev;
    a.system = system_dev;
    a.params = params_dev;
    a.n = n;
    CTR_CUDA(cudaMemsetAsync(hist, 0, 256 * sizeof(int32_t), st));
    k_bvp_order_key<<<(int)((n + 255) / 256), 256, 0, st>>>(a, many_first, key, hist);
    k_bvp_order_scan<<<1, 256, 0, st>>>(hist);
    k_bvp_order_scatter<<<(int)((n + 255) / 256), 256, 0, st>>>(n, key, hist, order);
    h->launches += 3;
    CTR_CUDA(cudaGetLastError());
    *order_out = order;
    return 0;
}

static int bvp_launch(ctr_handle* h, const BvpArgs& a_in, cudaStream_t st) {
    BvpArgs a = a_in;
    const int64_t total = (int64_t)bvp_variants(a.mode) * a.n;
    static const bool thread_kernel = getenv("CTR_BVP_THREAD_KERNEL") != nullptr;   // A/B switch: one thread per pass (round-2 first form)
    if (thread_kernel) {
        k_bvp_pass<<<(int)((total + CTR_BVP_BLOCK - 1) / CTR_BVP_BLOCK), CTR_BVP_BLOCK, 0, st>>>(a);
    } else {
        // work counters: a ring of 64, one per launch, so that launches the caller runs on different streams do not share one
        if (!h->bvp_next) CTR_CUDA(cudaMalloc(&h->bvp_next, 64 * sizeof(unsigned long long)));
        a.next = h->bvp_next + (h->bvp_launches++ & 63);
        CTR_CUDA(cudaMemsetAsync(a.next, 0, sizeof(unsigned long long), st));
        {   // scheduling order (k_bvp_order_*): similar configurations together, long ones first
            const int32_t* order = nullptr;
            int rc = build_order(h, a.q, a.system, a.params, a.n, 0, (ORDER_MIN_BVP_PASSES + bvp_variants(a.mode) - 1) / bvp_variants(a.mode), st, &order);
            if (rc) return rc;
            a.order = order;
        }
        const int per_block = (CTR_BLOCK / 32) * 10;     // 3 lanes per pass, 10 passes in flight per warp
        int64_t grid = (total + per_block - 1) / per_block;
        const int64_t resident = (int64_t)h->sm_count * CTR_BVP_GMINB;   // persistent: the groups refill themselves from a.next
        if (grid > resident) grid = resident;
        if (a.f || (a.mode & 2)) k_bvp_pass_g<true><<<(int)grid, CTR_BLOCK, 0, st>>>(a);
        else k_bvp_pass_g<false><<<(int)grid, CTR_BLOCK, 0, st>>>(a);
    }
    h->launches += 1;
    CTR_CUDA(cudaGetLastError());
    return 0;
}

int ctr_bvp_pass(ctr_handle* h, const double* q_dev, const double* f_dev, const double* u_dev, const int32_t* system_dev,
                 const double* params_dev, int64_t n, double max_step, double* tip_dev, double* R_dev, double* cost_dev,
                 int32_t* steps_dev, void* stream) {
    if (!h || !q_dev || !tip_dev) CTR_FAIL("null argument");
    if (!(max_step > 0.0)) CTR_FAIL("max_step must be > 0");
    if (n <= 0) return 0;
    CTR_CUDA(cudaSetDevice(h->device));
    BvpArgs a = bvp_args(h, q_dev, f_dev, u_dev, system_dev, params_dev, n, 0, max_step);
    a.tip = tip_dev;
    a.R = R_dev;
    a.cost = cost_dev;
    a.steps = steps_dev;
    return bvp_launch(h, a, (cudaStream_t)stream);
}

int ctr_bvp_jacobian(ctr_handle* h, const double* q_dev, const double* f_dev, const double* u_dev, const int32_t* system_dev,
                     const double* params_dev, int64_t n, double max_step, double* J_dev, double* C_dev, double* tip_dev,
                     void* stream) {
    if (!h || !q_dev || (!J_dev && !C_dev)) CTR_FAIL("null argument");
    if (!(max_step > 0.0)) CTR_FAIL("max_step must be > 0");
    if (n <= 0) return 0;
    CTR_CUDA(cudaSetDevice(h->device));
    cudaStream_t st = (cudaStream_t)stream;
    const int mode = (J_dev ? 1 : 0) | (C_dev ? 2 : 0);
    char* base = nullptr;
    int rc = bvp_scratch(h, sizeof(double) * 3 * (size_t)bvp_variants(mode) * (size_t)n, &base);
    if (rc) return rc;
    BvpArgs a = bvp_args(h, q_dev, f_dev, u_dev, system_dev, params_dev, n, mode, max_step);
    a.tip = (double*)base;
    rc = bvp_launch(h, a, st);
    if (rc) return rc;
    k_bvp_fd<<<(int)((n + 127) / 128), 128, 0, st>>>(a.tip, n, mode, a.eps_q, a.eps_f, J_dev, C_dev, tip_dev);
    h->launches += 1;
    CTR_CUDA(cudaGetLastError());
    return 0;
}

int ctr_bvp_solve(ctr_handle* h, const double* q_dev, const double* f_dev, const int32_t* system_dev, const double* params_dev,
                  int64_t n, double max_step, double f_tol, int32_t max_iters, double* u_dev, double* tip_dev, double* cost_dev,
                  double* res_dev, int32_t* iters_dev, void* stream) {
    if (!h || !q_dev || !u_dev || !res_dev || !iters_dev) CTR_FAIL("null argument");
    if (!(max_step > 0.0) || !(f_tol > 0.0) || max_iters < 0) CTR_FAIL("bad max_step / f_tol / max_iters");
    if (n <= 0) return 0;
    CTR_CUDA(cudaSetDevice(h->device));
    cudaStream_t st = (cudaStream_t)stream;
    // scratch: tips f64 [3][6n] | cost f64 [5][6n] | solver state f64 [37][n] | active i32 [16] | frozen u8 [n]
    char* base = nullptr;
    int rc = bvp_scratch(h, (size_t)n * ((18 + 30 + BVP_LM_STATE) * 8 + 1) + 64, &base);
    if (rc) return rc;
    double* tips = (double*)base;
    double* cost = tips + 18 * n;
    double* nst = cost + 30 * n;
    int32_t* active = (int32_t*)(nst + BVP_LM_STATE * n);
    uint8_t* frozen = (uint8_t*)(active + 16);
    BvpArgs a = bvp_args(h, q_dev, f_dev, u_dev, system_dev, params_dev, n, 4, max_step);
    a.tip = tips;
    a.cost = cost;
    k_bvp_guess<<<(int)((n + 127) / 128), 128, 0, st>>>(a, u_dev);          // CTR_Model.minimize's starting point
    h->launches += 1;
    CTR_CUDA(cudaMemsetAsync(frozen, 0, (size_t)n, st));
    CTR_CUDA(cudaMemsetAsync(iters_dev, 0, sizeof(int32_t) * n, st));
    for (int it = 0; it <= max_iters; ++it) {
        CTR_CUDA(cudaMemsetAsync(active, 0, sizeof(int32_t), st));
        rc = bvp_launch(h, a, st);
        if (rc) return rc;
        k_bvp_lm<<<(int)((n + 127) / 128), 128, 0, st>>>(cost, n, a.eps_u, f_tol, it == 0 ? 1 : 0, it == max_iters ? 1 : 0, u_dev, nst,
                                                          res_dev, iters_dev, frozen, active);
        h->launches += 1;
        CTR_CUDA(cudaGetLastError());
        int32_t act = 0;
        CTR_CUDA(cudaMemcpyAsync(&act, active, sizeof(int32_t), cudaMemcpyDeviceToHost, st));
        CTR_CUDA(cudaStreamSynchronize(st));
        if (act == 0) break;
    }
    if (tip_dev || cost_dev) {                                               // the pass at the solution
        BvpArgs b = bvp_args(h, q_dev, f_dev, u_dev, system_dev, params_dev, n, 0, max_step);
        b.tip = tip_dev ? tip_dev : tips;
        b.cost = cost_dev;
        rc = bvp_launch(h, b, st);
        if (rc) return rc;
    }
    CTR_CUDA(cudaStreamSynchronize(st));
    return 0;
}

int ctr_bvp_ik_dls(ctr_handle* h, const double* target_dev, double* q_dev, const int32_t* system_dev, int64_t n, double lam,
                   double tol, int32_t max_iters, double max_step, double* err_dev, int32_t* iters_dev, void* stream) {
    if (!h || !target_dev || !q_dev || !err_dev || !iters_dev) CTR_FAIL("null argument");
    if (!(lam >= 0.0) || max_iters < 0 || !(max_step > 0.0)) CTR_FAIL("bad lam / max_iters / max_step");
    if (n <= 0) return 0;
    CTR_CUDA(cudaSetDevice(h->device));
    cudaStream_t st = (cudaStream_t)stream;
    // scratch: tips f64 [3][7n] | active i32 [16] | frozen u8 [n]
    char* base = nullptr;
    int rc = bvp_scratch(h, (size_t)n * (21 * 8 + 1) + 64, &base);
    if (rc) return rc;
    double* tips = (double*)base;
    int32_t* active = (int32_t*)(tips + 21 * n);
    uint8_t* frozen = (uint8_t*)(active + 16);
    CTR_CUDA(cudaMemsetAsync(frozen, 0, (size_t)n, st));
    CTR_CUDA(cudaMemsetAsync(iters_dev, 0, sizeof(int32_t) * n, st));
    CTR_CUDA(cudaMemsetAsync(err_dev, 0, sizeof(double) * n, st));
    // src/jacobian_controller.py:19-53: CTR_Model(..., f = 0, q, q_0 = 0, Tol, 1).jac(u_init = 0) every iteration
    BvpArgs a = bvp_args(h, q_dev, nullptr, nullptr, system_dev, nullptr, n, 1, max_step);
    a.tip = tips;
    for (int it = 0; it < max_iters; ++it) {
        CTR_CUDA(cudaMemsetAsync(active, 0, sizeof(int32_t), st));
        rc = bvp_launch(h, a, st);
        if (rc) return rc;
        k_ik_dls_update<false><<<(int)((n + 255) / 256), 256, 0, st>>>(tips, target_dev, n, a.eps_q, lam, tol, q_dev, err_dev, iters_dev,
                                                               frozen, active);
        h->launches += 1;
        CTR_CUDA(cudaGetLastError());
        if ((it & 3) == 3 || it == max_iters - 1) {
            int32_t act = 0;
            CTR_CUDA(cudaMemcpyAsync(&act, active, sizeof(int32_t), cudaMemcpyDeviceToHost, st));
            CTR_CUDA(cudaStreamSynchronize(st));
            if (act == 0) break;
        }
    }
    CTR_CUDA(cudaStreamSynchronize(st));
    return 0;
}

}  // extern "C"
