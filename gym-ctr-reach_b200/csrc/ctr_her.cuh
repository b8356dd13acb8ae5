// ctr_her.cuh -- hindsight-replay kernels, included by ctr_kernels.cu.
//
// The reference hands replay and goal relabelling to stable_baselines' HER wrapper (src/utils.py:5-7,27-28; settings in
// saved_policies/**/config.yml: goal_selection_strategy future, n_sampled_goal 4).  That package is third party, absent
// from /root/reference and from this image, and its version is not pinned by the reference (setup.py:5), so the sampling
// policy below is a restatement of the published 'future' strategy and its parity is UNPINNED; what IS the reference's
// own code on this path -- compute_reward (ctr_reach_env.py:160-170) -- is bit-exact (same arithmetic as k_reward).
//
// Both kernels are HBM-bound byte movers: k_her_store is coalesced (thread = env, rows are contiguous in env),
// k_her_sample is a gather (thread = sample, every row touches one 32-byte sector of a random ring column).
#pragma once
#include "../../include/ctr_b200.h"
#include "ctr_device.cuh"

namespace ctr {

struct HerSrc {                    // the handle's per-step result buffers the store kernel reads
    const double* obs;             // f64 [D][N]  observation after the step (after auto-reset for finished envs)
    const double* achieved;        // f64 [3][N]
    const double* desired;         // f64 [3][N]
    const double* terminal_obs;    // f64 [D][N]  valid where done (auto_reset)
    const double* terminal_achieved;
    const double* reward;
    const uint8_t* done;
    int32_t auto_reset;
};

__global__ void k_her_begin(const ctr_her_ring r, const HerSrc s) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= r.n_envs) return;
    const int64_t n = r.n_envs;
    for (int d = 0; d < r.obs_dim; ++d) r.cur_obs[(int64_t)d * n + e] = s.obs[(int64_t)d * n + e];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        r.cur_achieved[(int64_t)c * n + e] = s.achieved[(int64_t)c * n + e];
        r.cur_desired[(int64_t)c * n + e] = s.desired[(int64_t)c * n + e];
    }
    r.cur_t[e] = 0;
}

__global__ void k_her_store(const ctr_her_ring r, const HerSrc s, const float* __restrict__ action, int64_t stride_field,
                            int64_t stride_env, int64_t slot) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= r.n_envs) return;
    const int64_t n = r.n_envs, CN = r.capacity * n, col = slot * n + e;
    const bool done = s.done[e] != 0;
    const bool term = done && s.auto_reset;       // the live buffers already hold the next episode's start state
    for (int d = 0; d < r.obs_dim; ++d) {
        const double now = s.obs[(int64_t)d * n + e];
        r.obs[(int64_t)d * CN + col] = r.cur_obs[(int64_t)d * n + e];
        r.next_obs[(int64_t)d * CN + col] = term ? s.terminal_obs[(int64_t)d * n + e] : now;
        r.cur_obs[(int64_t)d * n + e] = now;
    }
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const double a_now = s.achieved[(int64_t)c * n + e], d_now = s.desired[(int64_t)c * n + e];
        r.achieved[(int64_t)c * CN + col] = r.cur_achieved[(int64_t)c * n + e];
        r.next_achieved[(int64_t)c * CN + col] = term ? s.terminal_achieved[(int64_t)c * n + e] : a_now;
        r.desired[(int64_t)c * CN + col] = r.cur_desired[(int64_t)c * n + e];
        r.cur_achieved[(int64_t)c * n + e] = a_now;
        r.cur_desired[(int64_t)c * n + e] = d_now;
    }
#pragma unroll
    for (int i = 0; i < 6; ++i) r.action[(int64_t)i * CN + col] = action[(int64_t)i * stride_field + e * stride_env];
    r.reward[col] = s.reward[e];
    r.done[col] = done ? 1 : 0;
    const int t = r.cur_t[e] + 1;
    r.t_in_ep[col] = t;
    if (done) {                                   // close the episode: every transition of it learns its length
        const int back = t < (int)r.capacity ? t : (int)r.capacity;
        int64_t sl = slot;
        for (int k = 0; k < back; ++k) {
            r.ep_len[sl * n + e] = t;
            sl = (sl == 0) ? r.capacity - 1 : sl - 1;
        }
        r.cur_t[e] = 0;
    } else {
        r.ep_len[col] = 0;
        r.cur_t[e] = t;
    }
}

// draws of sample i: Philox4x32-10(counter = (i_lo, i_hi, call_lo, call_hi), key = seed) -> x: slot age, y: env,
// z: relabel coin, w: future offset; integer ranges by multiply-shift ((u32 * range) >> 32).
__global__ void k_her_sample(const ctr_her_ring r, const ctr_her_batch b, int64_t n_stored, int64_t newest_slot,
                             int n_goal, int rewrite_error, double tol, int reward_mode, uint2 key, uint2 call) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= b.m) return;
    const int64_t n = r.n_envs, CN = r.capacity * n, M = b.m;
    const uint4 u = philox4x32(make_uint4((unsigned)i, (unsigned)((uint64_t)i >> 32), call.x, call.y), key);
    const int64_t age = (int64_t)(((uint64_t)u.x * (uint64_t)n_stored) >> 32);
    int64_t slot = newest_slot - age;
    if (slot < 0) slot += r.capacity;
    const int64_t env = (int64_t)(((uint64_t)u.y * (uint64_t)n) >> 32);
    const int64_t col = slot * n + env;
    const int k = r.t_in_ep[col];
    int L = r.ep_len[col];
    if (L == 0) L = r.cur_t[env];                 // open episode: the transitions stored so far
    const bool relabel = (k < L) && ((((uint64_t)u.z * (uint64_t)(n_goal + 1)) >> 32) < (uint64_t)n_goal);
    int64_t gcol = -1;
    if (relabel) {
        const int off = (int)(((uint64_t)u.w * (uint64_t)(L - k)) >> 32);     // donor transition k + off in [k, L - 1]
        int64_t gs = slot + off;
        if (gs >= r.capacity) gs -= r.capacity;
        gcol = gs * n + env;
    }
    double ach[3], nach[3], goal[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        ach[c] = r.achieved[(int64_t)c * CN + col];
        nach[c] = r.next_achieved[(int64_t)c * CN + col];
        goal[c] = relabel ? r.next_achieved[(int64_t)c * CN + gcol] : r.desired[(int64_t)c * CN + col];
        b.achieved[(int64_t)c * M + i] = ach[c];
        b.next_achieved[(int64_t)c * M + i] = nach[c];
        b.desired[(int64_t)c * M + i] = goal[c];
    }
    for (int d = 0; d < r.obs_dim; ++d) {
        double o = r.obs[(int64_t)d * CN + col], o2 = r.next_obs[(int64_t)d * CN + col];
        if (relabel && rewrite_error && d >= 9 && d < 12) {
            o = __dsub_rn(goal[d - 9], ach[d - 9]);          // obs.py:158: desired - achieved
            o2 = __dsub_rn(goal[d - 9], nach[d - 9]);
        }
        b.obs[(int64_t)d * M + i] = o;
        b.next_obs[(int64_t)d * M + i] = o2;
    }
#pragma unroll
    for (int a = 0; a < 6; ++a) b.action[(int64_t)a * M + i] = r.action[(int64_t)a * CN + col];
    double reward = r.reward[col];
    uint8_t done = r.done[col];
    if (relabel) {                                // ctr_reach_env.py:168-170 on (next achieved, new goal)
        const double x = __dsub_rn(nach[0], goal[0]), y = __dsub_rn(nach[1], goal[1]), z = __dsub_rn(nach[2], goal[2]);
        const double dist = __dsqrt_rn(__dadd_rn(__dadd_rn(__dmul_rn(x, x), __dmul_rn(y, y)), __dmul_rn(z, z)));
        reward = (reward_mode == CTR_REWARD_DENSE) ? -dist : ((dist > tol) ? -1.0 : -0.0);
        done = 0;
    }
    b.reward[i] = reward;
    b.done[i] = done;
    b.src[i] = col;
    b.goal_src[i] = gcol;
}

}  // namespace ctr
