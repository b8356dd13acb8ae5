// ctr_her.cuh -- hindsight-replay kernels, included by ctr_kernels.cu.
//
// The reference hands replay and goal relabelling to stable_baselines' HER wrapper (src/utils.py:5-7,27-28; settings in
// saved_policies/**/config.yml: goal_selection_strategy future, n_sampled_goal 4).  That package is third party, absent
// from /root/reference and from this image, and its version is not pinned by the reference (setup.py:5), so the sampling
// policy below is a restatement of the published 'future' strategy and its parity is UNPINNED; what IS the reference's
// own code on this path -- compute_reward (ctr_reach_env.py:160-170) -- is bit-exact (same arithmetic as k_reward).
//
// Both kernels are HBM-bound byte movers.  The ring is record major (include/ctr_b200.h: one 352-byte record per transition):
// k_her_store reads the env's SoA rows coalesced and writes one whole record per thread; k_her_sample gathers whole records,
// a warp per 32 samples, through shared memory.  (The first version kept the ring SoA: a sample then touched one 32-byte
// sector per 8-byte element, 757 GB/s algorithmic = 11.6 % of HBM, profiles/r1i_her_extras.json.)
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

constexpr int HER_REC = CTR_HER_RECORD;      // doubles per record (352 B = 11 sectors)
constexpr int HER_PAD = HER_REC + 1;         // shared-memory row stride (odd: the row-wise reads of phase 3 spread over banks)
#ifndef CTR_HER_MLP
#define CTR_HER_MLP 4
#endif
constexpr int HER_MLP = CTR_HER_MLP;         // records a warp loads per batch in k_her_sample (power of two <= 32)
constexpr int HER_WARPS = 2;
// batch rows are written once and read by another kernel later: streaming stores (evict-first) keep them from displacing ring lines in L2
#ifndef CTR_HER_PLAIN_STORES
#define HER_ST(p, v) __stcs((p), (v))
#else
#define HER_ST(p, v) (*(p) = (v))
#endif                 // warps per sampling block: 2 x 32 x 45 x 8 B = 23 KB of shared memory

// One warp per 32 environments.  Phase 1 (lane = env): the env's SoA rows are read coalesced, the record is assembled in
// shared memory, cur_* move on to the post-step state, the episode counters are updated.  Phase 2: the warp writes the 32
// records one after the other, 352 contiguous bytes each.  (A thread writing its own record 8 bytes at a time made every
// store instruction touch 32 sectors: 87.6 us for 131072 envs against 28.7 us for the SoA ring.)
__global__ void __launch_bounds__(32 * HER_WARPS) k_her_store(const ctr_her_ring r, const HerSrc s,
                                                             const float* __restrict__ action, int64_t stride_field,
                                                             int64_t stride_env, int64_t slot) {
    __shared__ double s_rec[HER_WARPS][32][HER_PAD];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int64_t e0 = ((int64_t)blockIdx.x * HER_WARPS + w) * 32;
    if (e0 >= r.n_envs) return;                                   // warp-uniform
    const int64_t n = r.n_envs, e = e0 + lane;
    const int D = r.obs_dim;
    double* me = s_rec[w][lane];
    if (e < n) {
        const int64_t col = slot * n + e;
        const bool done = s.done[e] != 0;
        const bool term = done && s.auto_reset;   // the live buffers already hold the next episode's start state
        for (int d = 0; d < D; ++d) {
            const double now = s.obs[(int64_t)d * n + e];
            me[d] = r.cur_obs[(int64_t)d * n + e];
            me[D + d] = term ? s.terminal_obs[(int64_t)d * n + e] : now;
            r.cur_obs[(int64_t)d * n + e] = now;
        }
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const double a_now = s.achieved[(int64_t)c * n + e], d_now = s.desired[(int64_t)c * n + e];
            me[2 * D + c] = r.cur_achieved[(int64_t)c * n + e];
            me[2 * D + 3 + c] = term ? s.terminal_achieved[(int64_t)c * n + e] : a_now;
            me[2 * D + 6 + c] = r.cur_desired[(int64_t)c * n + e];
            r.cur_achieved[(int64_t)c * n + e] = a_now;
            r.cur_desired[(int64_t)c * n + e] = d_now;
        }
        me[2 * D + 9] = s.reward[e];
        float* act = reinterpret_cast<float*>(me + 2 * D + 10);
#pragma unroll
        for (int i = 0; i < 6; ++i) act[i] = action[(int64_t)i * stride_field + e * stride_env];
        for (int d = 2 * D + 13; d < HER_REC - 2; ++d) me[d] = 0.0;   // padding: defined bytes
        r.done[col] = done ? 1 : 0;
        const int t = r.cur_t[e] + 1;
        r.t_in_ep[col] = t;
        // the sampler's index arithmetic needs (t_in_ep, ep_len, done) of a transition BEFORE it can fetch anything else: a copy
        // in the record's last 16 bytes makes that ONE sector of the record it reads anyway instead of three 4- / 1-byte
        // gathers from the side arrays (which stay authoritative for host-side readers)
        me[HER_REC - 2] = done ? 1.0 : 0.0;
        me[HER_REC - 1] = __hiloint2double(done ? t : 0, t);       // high word: ep_len (0 = episode still open), low word: t_in_ep
        if (done) {                               // close the episode: every transition of it learns its length
            const int back = t < (int)r.capacity ? t : (int)r.capacity;
            int64_t sl = slot;
            for (int k = 0; k < back; ++k) {
                r.ep_len[sl * n + e] = t;
                if (k > 0) reinterpret_cast<int32_t*>(r.rec + (sl * n + e) * HER_REC + (HER_REC - 1))[1] = t;   // the record's copy
                sl = (sl == 0) ? r.capacity - 1 : sl - 1;
            }
            r.cur_t[e] = 0;
        } else {
            r.ep_len[col] = 0;
            r.cur_t[e] = t;
        }
    }
    __syncwarp();
    const int cnt = (int)((n - e0 < 32) ? n - e0 : 32);
    double* out = r.rec + (slot * n + e0) * HER_REC;              // the 32 records of this warp are contiguous in the ring
    for (int k = lane; k < cnt * HER_REC; k += 32) out[k] = s_rec[w][k / HER_REC][k % HER_REC];
}

// One warp per 32 samples.  Phase 1: lane l draws sample l (Philox4x32-10, counter = (i_lo, i_hi, call_lo, call_hi),
// key = seed -> x: slot age, y: env, z: relabel coin, w: future offset; integer ranges by multiply-shift) and reads the three
// integers the index arithmetic needs.  Phase 2: the warp copies the 32 records into shared memory, one record per iteration,
// 352 coalesced bytes each (+ the donor's next_achieved for relabelled samples).  Phase 3: every output row is written by the
// 32 lanes at once (256 contiguous bytes).  A sample costs 13 sectors of reads (11 of its record, whose last one is also what
// phase 1 needs, the env's open-episode counter, the donor's goal) instead of one sector per 8-byte element.
__global__ void __launch_bounds__(32 * HER_WARPS) k_her_sample(const ctr_her_ring r, const ctr_her_batch b, int64_t n_stored,
                                                              int64_t newest_slot, int n_goal, int rewrite_error, double tol,
                                                              int reward_mode, uint2 key, uint2 call) {
    __shared__ double s_rec[HER_WARPS][32][HER_PAD];
    __shared__ double s_goal[HER_WARPS][32][3];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int64_t i0 = ((int64_t)blockIdx.x * HER_WARPS + w) * 32;
    if (i0 >= b.m) return;                                        // warp-uniform
    const int64_t i = i0 + lane, M = b.m, n = r.n_envs;
    const bool live = i < M;
    const int D = r.obs_dim;
    // ---- phase 1
    int64_t col = 0, gcol = -1;
    bool relabel = false, done_rec = false;
    if (live) {
        const uint4 u = philox4x32(make_uint4((unsigned)i, (unsigned)((uint64_t)i >> 32), call.x, call.y), key);
        const int64_t age = (int64_t)(((uint64_t)u.x * (uint64_t)n_stored) >> 32);
        int64_t slot = newest_slot - age;
        if (slot < 0) slot += r.capacity;
        const int64_t env = (int64_t)(((uint64_t)u.y * (uint64_t)n) >> 32);
        col = slot * n + env;
        const double2 tail = __ldg(reinterpret_cast<const double2*>(r.rec + col * HER_REC + (HER_REC - 2)));
        done_rec = tail.x != 0.0;
        const int k = __double2loint(tail.y);
        int L = __double2hiint(tail.y);
        if (L == 0) L = r.cur_t[env];                             // open episode: the transitions stored so far
        relabel = (k < L) && ((((uint64_t)u.z * (uint64_t)(n_goal + 1)) >> 32) < (uint64_t)n_goal);
        if (relabel) {
            const int off = (int)(((uint64_t)u.w * (uint64_t)(L - k)) >> 32);   // donor transition k + off in [k, L - 1]
            int64_t gs = slot + off;
            if (gs >= r.capacity) gs -= r.capacity;
            gcol = gs * n + env;
        }
    }
    // ---- phase 2
    // HER_MLP records per batch: all their loads are issued before the first shared-memory store, so a warp keeps
    // HER_MLP x 376 bytes in flight instead of one record (the kernel is latency bound: random 352-byte reads)
#pragma unroll 1
    for (int s0 = 0; s0 < 32; s0 += HER_MLP) {
        double v0[HER_MLP], v1[HER_MLP], vg[HER_MLP];
#pragma unroll
        for (int u = 0; u < HER_MLP; ++u) {
            const int64_t c = __shfl_sync(FULL, col, s0 + u), g = __shfl_sync(FULL, gcol, s0 + u);
            const double* rec = r.rec + c * HER_REC;
            v0[u] = __ldg(rec + lane);
            v1[u] = (lane < HER_REC - 32) ? __ldg(rec + 32 + lane) : 0.0;
            vg[u] = (g >= 0 && lane < 3) ? __ldg(r.rec + g * HER_REC + 2 * D + 3 + lane) : 0.0;
        }
#pragma unroll
        for (int u = 0; u < HER_MLP; ++u) {
            s_rec[w][s0 + u][lane] = v0[u];
            if (lane < HER_REC - 32) s_rec[w][s0 + u][32 + lane] = v1[u];
            if (lane < 3) s_goal[w][s0 + u][lane] = vg[u];
        }
    }
    __syncwarp();
    // ---- phase 3 (lane = sample)
    const double* me = s_rec[w][lane];
    double ach[3], nach[3], goal[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        ach[c] = me[2 * D + c];
        nach[c] = me[2 * D + 3 + c];
        goal[c] = relabel ? s_goal[w][lane][c] : me[2 * D + 6 + c];
    }
    double reward = me[2 * D + 9];
    uint8_t done = (live && done_rec) ? 1 : 0;
    if (relabel) {                                                // ctr_reach_env.py:168-170 on (next achieved, new goal)
        const double x = __dsub_rn(nach[0], goal[0]), y = __dsub_rn(nach[1], goal[1]), z = __dsub_rn(nach[2], goal[2]);
        const double dist = __dsqrt_rn(__dadd_rn(__dadd_rn(__dmul_rn(x, x), __dmul_rn(y, y)), __dmul_rn(z, z)));
        reward = (reward_mode == CTR_REWARD_DENSE) ? -dist : ((dist > tol) ? -1.0 : -0.0);
        done = 0;
    }
    if (!live) return;
    for (int d = 0; d < D; ++d) {
        double o = me[d], o2 = me[D + d];
        if (relabel && rewrite_error && d >= 9 && d < 12) {
            o = __dsub_rn(goal[d - 9], ach[d - 9]);               // obs.py:158: desired - achieved
            o2 = __dsub_rn(goal[d - 9], nach[d - 9]);
        }
        HER_ST(b.obs + (int64_t)d * M + i, o);
        HER_ST(b.next_obs + (int64_t)d * M + i, o2);
    }
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        HER_ST(b.achieved + (int64_t)c * M + i, ach[c]);
        HER_ST(b.next_achieved + (int64_t)c * M + i, nach[c]);
        HER_ST(b.desired + (int64_t)c * M + i, goal[c]);
    }
    const float* act = reinterpret_cast<const float*>(me + 2 * D + 10);
#pragma unroll
    for (int a = 0; a < 6; ++a) HER_ST(b.action + (int64_t)a * M + i, act[a]);
    HER_ST(b.reward + i, reward);
    b.done[i] = done;
    b.src[i] = col;
    b.goal_src[i] = gcol;
}

}  // namespace ctr
