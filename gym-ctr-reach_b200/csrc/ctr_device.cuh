// ctr_device.cuh -- device-side building blocks of the B200 CTR-Reach step path (sm_100a).
//
// Execution model: one environment is owned by a GROUP of 3 adjacent lanes (one lane per tube).  Lane j holds
//   y_j = { uz_j, alpha_j, r_j, R[j][0], R[j][1], R[j][2] }
// i.e. the torsion pair of tube j, component j of the position and ROW j of the rotation matrix.  The 18-dim ODE
// of the reference (model.py:72-117) couples these slices only through
//   * the three angle differences  (one sincos per lane, the other one arrives by shuffle), and
//   * (ux_0, uy_0, uz_0) of tube 0  (broadcast from lane 0 of the group).
// so the whole Dormand-Prince state (y, 7 stage derivatives) is 6 x 8 values per lane and stays in registers
// even in float64.  A warp carries 10 environments (lanes 30/31 idle).  All shuffles are executed at
// warp-uniform program points with the full mask; per-environment control flow is expressed by predication.
//
// Reference citations are relative to /root/reference/ctr_reach_envs/ ; "rk.py"/"common.py" are
// scipy/integrate/_ivp/* (the third-party integrator the reference calls, model.py:149-151).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace ctr {

constexpr unsigned FULL = 0xffffffffu;
constexpr int GROUP = 3;
constexpr int ENVS_PER_WARP = 10;
constexpr int MAX_ATTEMPTS = 4000;  // hard cap on RK step attempts per FK (termination guard for NaN inputs)

// ---- per-system constants staged once per block in shared memory --------------------------------------------
struct SysConst {
    double L[3], Lc[3], EI[3], GJ[3], Ux[3], Uy[3];  // derived tube constants (Tube.py:14-15, Segment.py:57-59)
    float low[6], high[6];                           // joint Box bounds, float32 (obs.py:59-75)
    float d10, d21;                                  // f32(L1 - L0), f32(L2 - L1) (obs.py:181, NEP-50 weak python float)
    float Lf[3];                                     // f32(L_i) for the goal-sampling predicate (obs.py:198-200)
    float slow[3];                                   // sample-space low for beta (obs.py:59-63)
};

struct LaneId {
    int lane, g, j;            // lane in warp, group in warp, tube index (lane in group)
    int l0, lnext, lprev;      // absolute lane ids: group lane 0, tube (j+1)%3, tube (j+2)%3
};

__device__ __forceinline__ LaneId lane_id() {
    LaneId id;
    id.lane = threadIdx.x & 31;
    id.g = id.lane / GROUP;
    id.j = id.lane - id.g * GROUP;
    id.l0 = id.g * GROUP;
    id.lnext = id.l0 + (id.j + 1) % GROUP;
    id.lprev = id.l0 + (id.j + 2) % GROUP;
    return id;
}

template <typename T>
__device__ __forceinline__ T pick3(int j, T a0, T a1, T a2) {
    return j == 0 ? a0 : (j == 1 ? a1 : a2);
}

// (p0 + p1) + p2 of the three lanes' partials, identical bits on all three lanes.
template <typename T>
__device__ __forceinline__ T group_sum(const LaneId& id, T p) {
    T p0 = __shfl_sync(FULL, p, id.l0);
    T p1 = __shfl_sync(FULL, p, id.l0 + 1);
    T p2 = __shfl_sync(FULL, p, id.l0 + 2);
    return (p0 + p1) + p2;
}

// ---- math helpers, overloaded on the state type --------------------------------------------------------------
__device__ __forceinline__ void sincos_t(double x, double* s, double* c) { sincos(x, s, c); }
__device__ __forceinline__ void sincos_t(float x, float* s, float* c) { sincosf(x, s, c); }
__device__ __forceinline__ double fma_t(double a, double b, double c) { return fma(a, b, c); }
__device__ __forceinline__ float fma_t(float a, float b, float c) { return fmaf(a, b, c); }
__device__ __forceinline__ double abs_t(double a) { return fabs(a); }
__device__ __forceinline__ float abs_t(float a) { return fabsf(a); }
__device__ __forceinline__ double max_t(double a, double b) { return fmax(a, b); }
__device__ __forceinline__ float max_t(float a, float b) { return fmaxf(a, b); }
__device__ __forceinline__ double rcp_t(double a) { return 1.0 / a; }
__device__ __forceinline__ float rcp_t(float a) { return 1.0f / a; }

// ---- section geometry (Segment.py:6-61) -- IEEE float64, explicit _rn intrinsics so nothing is contracted ----
struct Geometry {
    double pts[10];      // sorted transition points (dynamic index -> local memory; touched once per section)
    double beta[3], dtip[3], dc[3];
    double bmin;
};

__device__ __forceinline__ void cswap(double& a, double& b) {
    double lo = fmin(a, b), hi = fmax(a, b);
    a = lo;
    b = hi;
}

__device__ __forceinline__ void build_geometry(const double* L, const double* Lc, const double* beta, Geometry& G) {
    double p[10];
    p[0] = 0.0;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        G.beta[i] = beta[i];
        G.dtip[i] = __dadd_rn(L[i], beta[i]);          // Segment.py:12
        G.dc[i] = __dsub_rn(G.dtip[i], Lc[i]);         // Segment.py:13
        p[1 + i] = beta[i];
        p[4 + i] = G.dc[i];
        p[7 + i] = G.dtip[i];
    }
    G.bmin = fmin(fmin(beta[0], beta[1]), beta[2]);    // Segment.py:53 min(base)
    // 29-comparator sorting network for 10 keys (Segment.py:15-16 np.sort; tie order is irrelevant)
#define CS(a, b) cswap(p[a], p[b])
    CS(0, 8); CS(1, 9); CS(2, 7); CS(3, 5); CS(4, 6);
    CS(0, 2); CS(1, 4); CS(5, 8); CS(7, 9);
    CS(0, 3); CS(2, 4); CS(5, 7); CS(6, 9);
    CS(0, 1); CS(3, 6); CS(8, 9);
    CS(1, 5); CS(2, 3); CS(4, 8); CS(6, 7);
    CS(1, 2); CS(3, 5); CS(4, 6); CS(7, 8);
    CS(2, 3); CS(4, 5); CS(6, 7);
    CS(3, 4); CS(5, 6);
#undef CS
#pragma unroll
    for (int i = 0; i < 10; ++i) G.pts[i] = p[i];
}

// ---- per-section constants in the lane-rotated form -----------------------------------------------------------
template <typename T>
struct Section {
    T Ao, Bo;   // A_j, B_j of the lane's own tube:   A_m = EI_m Ux_m / sum(EI),  B_m = EI_m Uy_m / sum(EI)
    T An, Bn;   // tube (j+1)%3
    T Ap, Bp;   // tube (j+2)%3
    T kx, ky;   // (EI_j / GJ_j) * Ux_j , (EI_j / GJ_j) * Uy_j   (0 when tube j is absent: model.py:97-99)
    T present;  // 1 when tube j exists on this section (alpha_j' = uz_j), else 0
};

template <typename T>
__device__ __forceinline__ void make_section(const Geometry& G, const double* EI, const double* GJ, const double* Ux,
                                             const double* Uy, double lo, double hi, int j, Section<T>& sc) {
    double ei[3], a[3], b[3];
    bool curved[3];
#pragma unroll
    for (int m = 0; m < 3; ++m) {
        bool present = (G.beta[m] <= lo) && (hi <= G.dtip[m]);   // Segment.py:23-41 (value form)
        curved[m] = present && (G.dc[m] <= lo);                  // Segment.py:42-43
        ei[m] = present ? EI[m] : 0.0;
    }
    double inv = 1.0 / (ei[0] + ei[1] + ei[2]);                  // model.py:87
#pragma unroll
    for (int m = 0; m < 3; ++m) {
        a[m] = curved[m] ? ei[m] * Ux[m] * inv : 0.0;
        b[m] = curved[m] ? ei[m] * Uy[m] * inv : 0.0;
    }
    sc.Ao = (T)pick3(j, a[0], a[1], a[2]);
    sc.Bo = (T)pick3(j, b[0], b[1], b[2]);
    sc.An = (T)pick3(j, a[1], a[2], a[0]);
    sc.Bn = (T)pick3(j, b[1], b[2], b[0]);
    sc.Ap = (T)pick3(j, a[2], a[0], a[1]);
    sc.Bp = (T)pick3(j, b[2], b[0], b[1]);
    double eij = pick3(j, ei[0], ei[1], ei[2]);
    double gjj = pick3(j, GJ[0], GJ[1], GJ[2]);
    bool cj = pick3(j, curved[0], curved[1], curved[2]);
    double k = (eij != 0.0) ? eij / gjj : 0.0;                   // model.py:101
    sc.kx = (T)(cj ? k * pick3(j, Ux[0], Ux[1], Ux[2]) : 0.0);
    sc.ky = (T)(cj ? k * pick3(j, Uy[0], Uy[1], Uy[2]) : 0.0);
    sc.present = (eij != 0.0) ? (T)1 : (T)0;
}

// ---- ODE right-hand side (model.py:72-117), one lane's slice ------------------------------------------------
// Y = {uz_j, alpha_j, r_j, R_j0, R_j1, R_j2} -> D = d/ds of the same.  Executed by the full warp.
template <typename T>
__device__ __forceinline__ void rhs(const LaneId& id, const Section<T>& sc, const T (&Y)[6], T (&D)[6]) {
    T a_next = __shfl_sync(FULL, Y[1], id.lnext);
    T s_o, c_o;
    sincos_t(Y[1] - a_next, &s_o, &c_o);                 // angle alpha_j - alpha_{j+1}
    T c_p = __shfl_sync(FULL, c_o, id.lprev);            // cos(alpha_{j-1} - alpha_j)
    T s_p = __shfl_sync(FULL, s_o, id.lprev);            // sin(alpha_{j-1} - alpha_j) = -sin(alpha_j - alpha_{j-1})
    // model.py:87-94 with cos(0) = 1, sin(0) = 0 folded
    T ux = fma_t(sc.An, c_o, sc.Ao);
    ux = fma_t(sc.Bn, s_o, ux);
    ux = fma_t(sc.Ap, c_p, ux);
    ux = fma_t(-sc.Bp, s_p, ux);
    T uy = fma_t(-sc.An, s_o, sc.Bo);
    uy = fma_t(sc.Bn, c_o, uy);
    uy = fma_t(sc.Ap, s_p, uy);
    uy = fma_t(sc.Bp, c_p, uy);
    D[0] = fma_t(ux, sc.ky, -(uy * sc.kx));              // model.py:101  (0 when absent: kx = ky = 0)
    D[1] = sc.present * Y[0];                            // model.py:99,102
    T ux0 = __shfl_sync(FULL, ux, id.l0);
    T uy0 = __shfl_sync(FULL, uy, id.l0);
    T uz0 = __shfl_sync(FULL, Y[0], id.l0);
    D[2] = Y[5];                                         // r' = R e3          (model.py:108)
    D[3] = fma_t(Y[4], uz0, -(Y[5] * uy0));              // R' = R u_hat, row j (model.py:107,109)
    D[4] = fma_t(Y[5], ux0, -(Y[3] * uz0));
    D[5] = fma_t(Y[3], uy0, -(Y[4] * ux0));
}

// Dormand-Prince 5(4) tableau (rk.py: class RK45)
#define DP_A21 (1.0 / 5)
#define DP_A31 (3.0 / 40)
#define DP_A32 (9.0 / 40)
#define DP_A41 (44.0 / 45)
#define DP_A42 (-56.0 / 15)
#define DP_A43 (32.0 / 9)
#define DP_A51 (19372.0 / 6561)
#define DP_A52 (-25360.0 / 2187)
#define DP_A53 (64448.0 / 6561)
#define DP_A54 (-212.0 / 729)
#define DP_A61 (9017.0 / 3168)
#define DP_A62 (-355.0 / 33)
#define DP_A63 (46732.0 / 5247)
#define DP_A64 (49.0 / 176)
#define DP_A65 (-5103.0 / 18656)
#define DP_B1 (35.0 / 384)
#define DP_B3 (500.0 / 1113)
#define DP_B4 (125.0 / 192)
#define DP_B5 (-2187.0 / 6784)
#define DP_B6 (11.0 / 84)
#define DP_E1 (-71.0 / 57600)
#define DP_E3 (71.0 / 16695)
#define DP_E4 (-71.0 / 1920)
#define DP_E5 (17253.0 / 339200)
#define DP_E6 (-22.0 / 525)
#define DP_E7 (1.0 / 40)

__device__ __forceinline__ double ulp_up(double t) {  // |nextafter(t, +inf) - t|   (rk.py:116)
    long long b = __double_as_longlong(t);
    double n = (t >= 0.0) ? __longlong_as_double(b + 1) : __longlong_as_double(b - 1);
    if (t == 0.0) n = __longlong_as_double(1);
    return fabs(n - t);
}

struct FkInfo {
    int nfev, nacc, nrej, nsec, failed;
};

// ---- forward kinematics (model.py:30-70, 119-174) for the lane group ----------------------------------------
// INTEG 0: scipy RK45 restated (rk.py rk_step/_step_impl, common.py select_initial_step), per-env adaptive.
// INTEG 1: classical RK4, rk4_steps per section.
// beta/alpha: float64 joint values (float32 joints promoted exactly, model.py:50-61).
// On return lane j holds y = {uz_j, alpha_j, r_j, R_j0..2} at the tip.
template <typename T, int INTEG>
__device__ __forceinline__ void fk_group(const LaneId& id, bool valid, const double* L, const double* Lc,
                                         const double* EI, const double* GJ, const double* Ux, const double* Uy,
                                         const double* beta, const double* alpha, double rtol_d, double atol_d,
                                         int rk4_steps, T (&y)[6], FkInfo& info) {
    Geometry G;
    build_geometry(L, Lc, beta, G);
    const int j = id.j;
    {   // initial state (model.py:56-64): uz = 0, alpha = joint[3:], r = 0, R = Rz(alpha_0)
        double s0, c0;
        sincos(alpha[0], &s0, &c0);
        y[0] = (T)0;
        y[1] = (T)pick3(j, alpha[0], alpha[1], alpha[2]);
        y[2] = (T)0;
        y[3] = (T)pick3(j, c0, s0, 0.0);
        y[4] = (T)pick3(j, -s0, c0, 0.0);
        y[5] = (T)pick3(j, 0.0, 0.0, 1.0);
    }
    info.nfev = info.nacc = info.nrej = info.nsec = info.failed = 0;
    const T rtol = (T)rtol_d, atol = (T)atol_d;
    Section<T> sc;
    sc.Ao = sc.Bo = sc.An = sc.Bn = sc.Ap = sc.Bp = sc.kx = sc.ky = sc.present = (T)0;
    T f[6] = {0, 0, 0, 0, 0, 0};   // derivative at (t, y)  (FSAL: rk.py:170)
    int k = -1;
    double cum = 0.0, prevS = 0.0;
    double t = 0.0, tf = 0.0, h_abs = 0.0;
    bool done = !valid, need_sec = true, rejected = false;
    int steps_left = 0;
    int attempts = 0;

    while (true) {
        // ---- advance to the next kept section (Segment.py:46-53, model.py:137-151); scalar, no shuffles
        bool init = false;
        while (!done && need_sec) {
            bool found = false;
            double lo = 0.0, hi = 0.0, S = 0.0;
            while (++k < 9) {
                lo = G.pts[k];
                hi = G.pts[k + 1];
                double len = __dmul_rn(1e-5, floor(__dmul_rn(1e5, __dsub_rn(hi, lo))));   // Segment.py:16
                if (len == 0.0) continue;                                                  // Segment.py:46
                cum = __dadd_rn(cum, len);                                                 // Segment.py:52
                S = __dadd_rn(cum, G.bmin);                                                // Segment.py:53
                if (!(S > 0.0)) continue;
                found = true;
                break;
            }
            if (!found) {
                done = true;
                break;
            }
            make_section<T>(G, EI, GJ, Ux, Uy, lo, hi, j, sc);
            double t0 = prevS;
            tf = __dsub_rn(S, 1e-6);                      // model.py:141
            if (tf < t0) {                                // model.py:145-151: sorted span, integrate (min, max)
                double z = t0;
                t0 = tf;
                tf = z;
            }
            prevS = S;
            info.nsec += 1;
            t = t0;
            if (tf == t0) {                               // zero-length span: solver returns y0 (base.py step())
                info.nfev += 1;
                continue;
            }
            need_sec = false;
            init = true;
            rejected = false;
            if (INTEG == 1) {
                h_abs = (tf - t0) / (double)rk4_steps;
                steps_left = rk4_steps;
            }
        }
        if (!__any_sync(FULL, !done)) break;

        if (INTEG == 0) {
            // ---- RungeKutta.__init__: f0 and select_initial_step (rk.py:93-101, common.py:68-134)
            if (__any_sync(FULL, init)) {
                T f0[6], f1[6], y1[6], rs[6];
                rhs<T>(id, sc, y, f0);
                T p0 = 0, p1 = 0;
#pragma unroll
                for (int i = 0; i < 6; ++i) {
                    rs[i] = rcp_t(fma_t(abs_t(y[i]), rtol, atol));     // 1 / scale
                    T a = y[i] * rs[i], b = f0[i] * rs[i];
                    p0 = fma_t(a, a, p0);
                    p1 = fma_t(b, b, p1);
                }
                double d0 = sqrt((double)group_sum(id, p0) * (1.0 / 18.0));   // common.py:63-65 RMS norm
                double d1 = sqrt((double)group_sum(id, p1) * (1.0 / 18.0));
                double interval = fabs(tf - t);
                double h0 = (d0 < 1e-5 || d1 < 1e-5) ? 1e-6 : 0.01 * d0 / d1;
                h0 = fmin(h0, interval);
                T h0t = (T)h0;
#pragma unroll
                for (int i = 0; i < 6; ++i) y1[i] = fma_t(h0t, f0[i], y[i]);
                rhs<T>(id, sc, y1, f1);
                T p2 = 0;
#pragma unroll
                for (int i = 0; i < 6; ++i) {
                    T c = (f1[i] - f0[i]) * rs[i];
                    p2 = fma_t(c, c, p2);
                }
                double d2 = sqrt((double)group_sum(id, p2) * (1.0 / 18.0)) / h0;
                double h1;
                if (d1 <= 1e-15 && d2 <= 1e-15) h1 = fmax(1e-6, h0 * 1e-3);
                else h1 = pow(0.01 / fmax(d1, d2), 0.2);
                if (init) {
                    h_abs = fmin(fmin(100.0 * h0, h1), interval);
                    info.nfev += 2;
#pragma unroll
                    for (int i = 0; i < 6; ++i) f[i] = f0[i];
                }
            }
        }

        const bool active = !done;   // every lane that is not done has a live section here
        // ---- one step attempt (rk.py:108-178)
        double min_step = 10.0 * ulp_up(t);
        if (INTEG == 0) {
            if (!rejected && h_abs < min_step) h_abs = min_step;
        }
        double t_new = t + h_abs;
        if (t_new - tf > 0.0) t_new = tf;
        if (INTEG == 1 && steps_left == 1) t_new = tf;
        const double h_d = t_new - t;
        const T h = (T)h_d;
        if (INTEG == 0) h_abs = fabs(h_d);   // rk.py:136-137: the factor below scales the step actually taken

        T ynew[6];
        bool accept = true;
        double factor = 1.0;
        if (INTEG == 0) {
            T k2[6], k3[6], k4[6], k5[6], k6[6], k7[6], Y[6];
#pragma unroll
            for (int i = 0; i < 6; ++i) Y[i] = fma_t(h * (T)DP_A21, f[i], y[i]);
            rhs<T>(id, sc, Y, k2);
#pragma unroll
            for (int i = 0; i < 6; ++i) Y[i] = fma_t(h, fma_t((T)DP_A32, k2[i], (T)DP_A31 * f[i]), y[i]);
            rhs<T>(id, sc, Y, k3);
#pragma unroll
            for (int i = 0; i < 6; ++i)
                Y[i] = fma_t(h, fma_t((T)DP_A43, k3[i], fma_t((T)DP_A42, k2[i], (T)DP_A41 * f[i])), y[i]);
            rhs<T>(id, sc, Y, k4);
#pragma unroll
            for (int i = 0; i < 6; ++i)
                Y[i] = fma_t(h, fma_t((T)DP_A54, k4[i], fma_t((T)DP_A53, k3[i], fma_t((T)DP_A52, k2[i], (T)DP_A51 * f[i]))), y[i]);
            rhs<T>(id, sc, Y, k5);
#pragma unroll
            for (int i = 0; i < 6; ++i)
                Y[i] = fma_t(h, fma_t((T)DP_A65, k5[i], fma_t((T)DP_A64, k4[i], fma_t((T)DP_A63, k3[i],
                             fma_t((T)DP_A62, k2[i], (T)DP_A61 * f[i])))), y[i]);
            rhs<T>(id, sc, Y, k6);
#pragma unroll
            for (int i = 0; i < 6; ++i)
                ynew[i] = fma_t(h, fma_t((T)DP_B6, k6[i], fma_t((T)DP_B5, k5[i], fma_t((T)DP_B4, k4[i],
                                fma_t((T)DP_B3, k3[i], (T)DP_B1 * f[i])))), y[i]);
            rhs<T>(id, sc, ynew, k7);
            T pe = 0;
#pragma unroll
            for (int i = 0; i < 6; ++i) {
                T e = fma_t((T)DP_E7, k7[i], fma_t((T)DP_E6, k6[i], fma_t((T)DP_E5, k5[i], fma_t((T)DP_E4, k4[i],
                            fma_t((T)DP_E3, k3[i], (T)DP_E1 * f[i])))));
                T sc_i = fma_t(max_t(abs_t(y[i]), abs_t(ynew[i])), rtol, atol);     // rk.py:141
                T q = (e * h) * rcp_t(sc_i);
                pe = fma_t(q, q, pe);
            }
            double err = sqrt((double)group_sum(id, pe) * (1.0 / 18.0));            // rk.py:142
            if (err < 1.0) {                                                        // rk.py:144-155
                factor = (err == 0.0) ? 10.0 : fmin(10.0, 0.9 * pow(err, -0.2));
                if (rejected) factor = fmin(1.0, factor);
            } else {                                                                // rk.py:157-159
                accept = false;
                factor = fmax(0.2, 0.9 * pow(err, -0.2));
            }
            if (active) {
                info.nfev += 6;
                if (accept) {
#pragma unroll
                    for (int i = 0; i < 6; ++i) f[i] = k7[i];
                }
            }
        } else {
            // classical RK4 (fixed step)
            T k1[6], k2[6], k3[6], k4[6], Y[6];
            rhs<T>(id, sc, y, k1);
#pragma unroll
            for (int i = 0; i < 6; ++i) Y[i] = fma_t(h * (T)0.5, k1[i], y[i]);
            rhs<T>(id, sc, Y, k2);
#pragma unroll
            for (int i = 0; i < 6; ++i) Y[i] = fma_t(h * (T)0.5, k2[i], y[i]);
            rhs<T>(id, sc, Y, k3);
#pragma unroll
            for (int i = 0; i < 6; ++i) Y[i] = fma_t(h, k3[i], y[i]);
            rhs<T>(id, sc, Y, k4);
#pragma unroll
            for (int i = 0; i < 6; ++i)
                ynew[i] = fma_t(h * (T)(1.0 / 6.0), (k1[i] + k4[i]) + (T)2 * (k2[i] + k3[i]), y[i]);
            if (active) info.nfev += 4;
        }

        if (active) {
            attempts += 1;
            if (accept) {
#pragma unroll
                for (int i = 0; i < 6; ++i) y[i] = ynew[i];
                t = t_new;
                info.nacc += 1;
                rejected = false;
                if (INTEG == 0) h_abs *= factor;
                if (INTEG == 1) steps_left -= 1;
                if (t == tf) need_sec = true;
            } else {
                h_abs *= factor;
                rejected = true;
                info.nrej += 1;
                if (h_abs < min_step) {   // rk.py:129-130 TOO_SMALL_STEP
                    info.failed = 1;
                    done = true;
                }
            }
            if (attempts >= MAX_ATTEMPTS) {
                info.failed = 1;
                done = true;
            }
        }
    }
}

// ---- Obs.set_action x n_substeps (obs.py:166-183, ctr_reach_env.py:134-135), float32 arithmetic -------------
__device__ __forceinline__ void set_action(float (&q)[6], const float (&a)[6], const SysConst& sc, int n_substeps) {
    for (int s = 0; s < n_substeps; ++s) {
#pragma unroll
        for (int i = 0; i < 6; ++i) q[i] = fminf(fmaxf(__fadd_rn(q[i], a[i]), sc.low[i]), sc.high[i]);   // obs.py:172
        q[0] = fminf(q[0], q[1]);                                   // obs.py:180 (i = 1)
        q[0] = fmaxf(q[0], __fadd_rn(sc.d10, q[1]));                // obs.py:181
        q[1] = fminf(q[1], q[2]);                                   // i = 2
        q[1] = fmaxf(q[1], __fadd_rn(sc.d21, q[2]));
    }
}

// ---- Philox4x32-10 counter-based RNG (results depend only on (seed, global env, episode, stream, draw)) -----
__device__ __forceinline__ uint4 philox4x32(uint4 ctr, uint2 key) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        unsigned hi0 = __umulhi(0xD2511F53u, ctr.x), lo0 = 0xD2511F53u * ctr.x;
        unsigned hi1 = __umulhi(0xCD9E8D57u, ctr.z), lo1 = 0xCD9E8D57u * ctr.z;
        ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
        key.x += 0x9E3779B9u;
        key.y += 0xBB67AE85u;
    }
    return ctr;
}

__device__ __forceinline__ double u01(unsigned x) { return (double)x * (1.0 / 4294967296.0); }   // [0, 1)

// Obs.sample_goal (obs.py:185-207): uniform float32 box beta in [-L+1e-3, 0], alpha in [-pi, pi]; accept iff
// b0 <= b1 <= b2 and b_{i-1} + L_{i-1} >= L_i + b_i (float32 arithmetic).  Returns the number of draws.
__device__ __forceinline__ int sample_goal(const SysConst& sc, uint2 key, unsigned env_lo, unsigned env_hi,
                                           unsigned episode, unsigned stream, float (&q)[6], bool& ok) {
    int tries = 0;
    ok = false;
    const float PI_F = 3.14159274101257324f;   // float32(np.pi): the Box bounds (obs.py:60-63)
    while (tries < 1001) {                     // obs.py:204: counter > 1000 -> ValueError
        uint4 r = philox4x32(make_uint4(env_lo, env_hi, episode, (stream << 16) | (unsigned)tries), key);
        unsigned rr[3] = {r.x, r.y, r.z};
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            double lo = (double)sc.slow[i];
            q[i] = (float)(lo + (0.0 - lo) * u01(rr[i]));   // uniform(low, high).astype(float32)
        }
        tries += 1;
        bool v = (q[0] <= q[1]) && (__fadd_rn(q[0], sc.Lf[0]) >= __fadd_rn(sc.Lf[1], q[1])) &&
                 (q[1] <= q[2]) && (__fadd_rn(q[1], sc.Lf[1]) >= __fadd_rn(sc.Lf[2], q[2]));
        if (v) {
            unsigned ra[3];
            ra[0] = r.w;
            uint4 r2 = philox4x32(make_uint4(env_lo, env_hi, episode, (stream << 16) | 0xFFFFu), key);
            ra[1] = r2.x;
            ra[2] = r2.y;
#pragma unroll
            for (int i = 0; i < 3; ++i) {
                double lo = -(double)PI_F, hi = (double)PI_F;
                q[3 + i] = (float)(lo + (hi - lo) * u01(ra[i]));
            }
            ok = true;
            break;
        }
    }
    return tries;
}

}  // namespace ctr
