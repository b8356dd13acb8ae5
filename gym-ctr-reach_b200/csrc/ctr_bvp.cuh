// ctr_bvp.cuh -- the reference's full concentric-tube model with tip load as a second kinematics back-end (SURVEY.md 8(f)
// row 4), included by ctr_kernels.cu.
//
// Reference: envs/CTR_Python/CTR_Model.py -- ode_eq :56-128 (20 states: u1x u1y | u1z u2z u3z | alpha_1..3 relative to tube 1 |
// r | R), ode_solver :130-236 (one shooting pass: per section solve_ivp(method='RK23', max_step=Tol), moment continuity at
// the section boundaries :186-226, the 5 residuals :228-235), jac :251-262, comp :264-275, minimize :277-299.  The
// integrator is scipy's RK23 (Bogacki-Shampine 3(2), rk.py / common.py) restated here like RK45 is in ctr_device.cuh.
//
// One thread per pass: this is the analysis / controller path (a Jacobian is 7 passes, a compliance matrix 4, a shooting
// iteration 6), not the env step; its 20-dim state with four stage vectors lives in local memory.
// Entries the reference leaves unassigned (np.empty, :63: twist derivatives of a tube absent on the section) are 0 here;
// that is also what the golden vectors pin (oracle/ref_loader.py: _BvpLoader, oracle/make_golden_bvp.py).
#pragma once
#include <cuda_runtime.h>
#include <math_constants.h>
#include <stdint.h>

namespace ctr {

constexpr int NB = 20;

struct BvpSection {
    double ei[3], gj[3], ux[3], uy[3];
};

// Segment.py:6-61 for one thread: end abscissae S[m] and the presence / curvature pattern of every kept section
// (same arithmetic as build_sections: IEEE float64 in source order, no contraction)
__device__ inline int bvp_sections(const double* L, const double* Lc, const double* beta, double* S, int* pat) {
    double p[10], dtip[3], dc[3];
    p[0] = 0.0;
    for (int i = 0; i < 3; ++i) {
        dtip[i] = __dadd_rn(L[i], beta[i]);
        dc[i] = __dsub_rn(dtip[i], Lc[i]);
        p[1 + i] = beta[i];
        p[4 + i] = dc[i];
        p[7 + i] = dtip[i];
    }
    for (int i = 1; i < 10; ++i) {                      // insertion sort (np.sort; tie order is irrelevant)
        const double v = p[i];
        int k = i - 1;
        while (k >= 0 && p[k] > v) {
            p[k + 1] = p[k];
            --k;
        }
        p[k + 1] = v;
    }
    const double bmin = fmin(fmin(beta[0], beta[1]), beta[2]);
    double cum = 0.0;
    int m = 0;
    for (int k = 0; k < 9; ++k) {
        const double lo = p[k], hi = p[k + 1];
        const double len = __dmul_rn(1e-5, floor(__dmul_rn(1e5, __dsub_rn(hi, lo))));   // Segment.py:16
        if (len == 0.0) continue;
        cum = __dadd_rn(cum, len);
        const double s = __dadd_rn(cum, bmin);
        if (!(s > 0.0)) continue;
        int pt = 0;
        for (int i = 0; i < 3; ++i) {
            const bool present = (beta[i] <= lo) && (hi <= dtip[i]);
            const bool curved = present && (dc[i] <= lo);
            pt |= (present ? 1 : 0) << i;
            pt |= (curved ? 8 : 0) << i;
        }
        S[m] = s;
        pat[m] = pt;
        m += 1;
    }
    return m;
}

__device__ inline void bvp_section_consts(const double* C /* L Lc EI GJ Ux Uy x3 */, int pat, BvpSection& c) {
    for (int i = 0; i < 3; ++i) {
        const bool present = (pat >> i) & 1, curved = (pat >> (3 + i)) & 1;
        c.ei[i] = present ? C[6 + i] : 0.0;
        c.gj[i] = present ? C[9 + i] : 0.0;
        c.ux[i] = curved ? C[12 + i] : 0.0;
        c.uy[i] = curved ? C[15 + i] : 0.0;
    }
}

// CTR_Model.py:56-128
__device__ inline void bvp_rhs(const BvpSection& c, const double* f, const double* y, double* d) {
    double u[9], ca[3], sa[3];
    ca[0] = 1.0;                                        // alpha_1 (relative to tube 1) stays 0: its derivative is u1z - u1z
    sa[0] = 0.0;
    sincos(y[6], &sa[1], &ca[1]);
    sincos(y[7], &sa[2], &ca[2]);
    if (y[5] != 0.0) sincos(y[5], &sa[0], &ca[0]);
    u[0] = y[0];
    u[1] = y[1];
    u[2] = y[2];
    for (int i = 1; i < 3; ++i) {                       // u_i = Rz(alpha_i)^T u_1 in xy, own twist (:64-73)
        u[3 * i + 0] = ca[i] * y[0] + sa[i] * y[1];
        u[3 * i + 1] = -sa[i] * y[0] + ca[i] * y[1];
        u[3 * i + 2] = y[2 + i];
    }
    for (int i = 0; i < 3; ++i) {                       // :76-82
        if (c.gj[i] != 0.0) {
            d[2 + i] = (c.ei[i] / c.gj[i]) * (u[3 * i] * c.uy[i] - u[3 * i + 1] * c.ux[i]);
            d[5 + i] = y[2 + i] - y[2];
        } else {
            d[2 + i] = 0.0;
            d[5 + i] = 0.0;
        }
    }
    double du0 = 0.0, du1 = 0.0, sei = 0.0;
    for (int i = 0; i < 3; ++i) {                       // :84-112
        const double ei = c.ei[i], gj = c.gj[i], uz = u[3 * i + 2], ux = u[3 * i], uyy = u[3 * i + 1];
        sei += ei;
        du0 += d[5 + i] * ei * y[1] + c.uy[i] * ei * uz * ca[i] + c.ux[i] * ei * uz * sa[i] - ei * uyy * uz * ca[i] +
               gj * uyy * uz * ca[i] - ei * ux * uz * sa[i] + gj * ux * uz * sa[i];
        du1 += -d[5 + i] * ei * y[0] + c.uy[i] * ei * uz * sa[i] - c.ux[i] * ei * uz * ca[i] + ei * ux * uz * ca[i] -
               gj * ux * uz * ca[i] - ei * uyy * uz * sa[i] + gj * uyy * uz * sa[i];
    }
    const double* R = y + 11;
    const double g0 = R[0] * f[0] + R[3] * f[1] + R[6] * f[2];      // R^T f
    const double g1 = R[1] * f[0] + R[4] * f[1] + R[7] * f[2];
    d[0] = -du0 / sei + g1 / sei;                                    // :114-118
    d[1] = -du1 / sei - g0 / sei;
    d[8] = R[2];
    d[9] = R[5];
    d[10] = R[8];
    for (int r = 0; r < 3; ++r) {                                    // R' = R u_hat(u_1)
        const double a = R[3 * r], b = R[3 * r + 1], cc = R[3 * r + 2];
        d[11 + 3 * r + 0] = b * y[2] - cc * y[1];
        d[11 + 3 * r + 1] = cc * y[0] - a * y[2];
        d[11 + 3 * r + 2] = a * y[1] - b * y[0];
    }
}

__device__ inline double bvp_rms(const double* x, const double* scale) {
    double s = 0.0;
    for (int i = 0; i < NB; ++i) {
        const double q = x[i] / scale[i];
        s += q * q;
    }
    return sqrt(s / NB);
}

struct BvpTipTwist {    // u_z of tubes 2, 3 at the stored point nearest their tips (:231-235, first minimum)
    double best[2], uz[2];
};
__device__ inline void bvp_tip_note(BvpTipTwist& tz, const double* dtip, double t, const double* y) {
    for (int i = 1; i < 3; ++i) {
        const double dist = fabs(t - dtip[i]);
        if (dist < tz.best[i - 1]) {
            tz.best[i - 1] = dist;
            tz.uz[i - 1] = y[2 + i];
        }
    }
}

// solve_ivp(fun, [t0, tf], y, method='RK23', max_step): y advanced in place; accepted steps, or -1 (step too small / NaN)
__device__ inline int bvp_rk23_section(const BvpSection& c, const double* f3, double t0, double tf, double* y, double rtol,
                                       double atol, double max_step, BvpTipTwist& tz, const double* dtip) {
    const double A21 = 0.5, A32 = 0.75, B1 = 2.0 / 9, B2 = 1.0 / 3, B3 = 4.0 / 9, E1 = 5.0 / 72, E2 = -1.0 / 12, E3 = -1.0 / 9,
                 E4 = 1.0 / 8;
    double f[NB], k2[NB], k3[NB], k4[NB], yn[NB], tmp[NB];
    double t = t0;
    int steps = 0;
    bvp_rhs(c, f3, y, f);
    bvp_tip_note(tz, dtip, t, y);
    double h_abs;
    {   // common.py select_initial_step, error-estimator order 2
        const double interval = fabs(tf - t0);
        if (interval == 0.0) return 0;
        for (int i = 0; i < NB; ++i) tmp[i] = atol + fabs(y[i]) * rtol;          // scale
        const double d0 = bvp_rms(y, tmp), d1 = bvp_rms(f, tmp);
        double h0 = (d0 < 1e-5 || d1 < 1e-5) ? 1e-6 : 0.01 * d0 / d1;
        h0 = fmin(h0, interval);
        for (int i = 0; i < NB; ++i) yn[i] = y[i] + h0 * f[i];
        bvp_rhs(c, f3, yn, k2);
        for (int i = 0; i < NB; ++i) k2[i] = k2[i] - f[i];
        const double d2 = bvp_rms(k2, tmp) / h0;
        const double h1 = (d1 <= 1e-15 && d2 <= 1e-15) ? fmax(1e-6, h0 * 1e-3) : pow(0.01 / fmax(d1, d2), 1.0 / 3.0);
        h_abs = fmin(fmin(100 * h0, h1), fmin(interval, max_step));
    }
    while (t != tf) {                                                 // rk.py _step_impl
        const double min_step = 10 * fabs(nextafter(t, CUDART_INF) - t);
        if (h_abs > max_step) h_abs = max_step;
        else if (h_abs < min_step) h_abs = min_step;
        bool rejected = false;
        for (;;) {
            if (h_abs < min_step) return -1;
            double t_new = t + h_abs;
            if (t_new - tf > 0) t_new = tf;
            const double h = t_new - t;
            h_abs = fabs(h);
            for (int i = 0; i < NB; ++i) tmp[i] = y[i] + (f[i] * A21) * h;
            bvp_rhs(c, f3, tmp, k2);
            for (int i = 0; i < NB; ++i) tmp[i] = y[i] + (f[i] * 0.0 + k2[i] * A32) * h;
            bvp_rhs(c, f3, tmp, k3);
            for (int i = 0; i < NB; ++i) yn[i] = y[i] + h * (f[i] * B1 + k2[i] * B2 + k3[i] * B3);
            bvp_rhs(c, f3, yn, k4);
            double s = 0.0;
            for (int i = 0; i < NB; ++i) {
                const double sc = atol + fmax(fabs(y[i]), fabs(yn[i])) * rtol;
                const double e = (f[i] * E1 + k2[i] * E2 + k3[i] * E3 + k4[i] * E4) * h / sc;
                s += e * e;
            }
            const double err = sqrt(s / NB);
            if (err < 1) {
                double factor = (err == 0) ? 10.0 : fmin(10.0, 0.9 * pow(err, -1.0 / 3.0));
                if (rejected) factor = fmin(1.0, factor);
                h_abs *= factor;
                t = t_new;
                for (int i = 0; i < NB; ++i) {
                    y[i] = yn[i];
                    f[i] = k4[i];
                }
                steps += 1;
                bvp_tip_note(tz, dtip, t, y);
                break;
            }
            if (!(err == err)) return -1;
            h_abs *= fmax(0.2, 0.9 * pow(err, -1.0 / 3.0));
            rejected = true;
        }
        if (steps > 100000) return -1;
    }
    return steps;
}

// CTR_Model.ode_solver(u_init), method 1.  C: the 18 tube constants; q: beta[3], alpha[3] (= q + q_0); f: tip force; u: u_init.
// y_out[20]: final state; cost[5]; returns accepted steps or -1.
__device__ inline int bvp_pass(const double* C, const double* q, const double* f, const double* u, double max_step, double rtol,
                               double atol, double* y, double* cost) {
    double S[9];
    int pat[9];
    const int m = bvp_sections(C, C + 3, q, S, pat);
    for (int i = 0; i < NB; ++i) y[i] = 0.0;
    for (int i = 0; i < 5; ++i) y[i] = u[i];
    const double a1 = q[3];
    y[6] = q[4] - a1;                                                // :31: twist angles relative to tube 1
    y[7] = q[5] - a1;
    double s1, c1;
    sincos(a1, &s1, &c1);
    y[11] = c1;
    y[12] = -s1;
    y[14] = s1;
    y[15] = c1;
    y[19] = 1.0;
    double dtip[3];
    for (int i = 0; i < 3; ++i) dtip[i] = C[i] + q[i];
    BvpTipTwist tz;
    tz.best[0] = tz.best[1] = CUDART_INF;
    tz.uz[0] = tz.uz[1] = 0.0;
    int total = 0;
    double prev = 0.0;
    BvpSection c, cn;
    if (m > 0) bvp_section_consts(C, pat[0], c);
    for (int seg = 0; seg < m; ++seg) {
        const int st = bvp_rk23_section(c, f, prev, S[seg], y, rtol, atol, max_step, tz, dtip);
        if (st < 0) return -1;
        total += st;
        prev = S[seg];
        if (seg < m - 1) {                                           // :186-226: moment continuity
            bvp_section_consts(C, pat[seg + 1], cn);
            double sx = 0.0, sy = 0.0, sein = 0.0;
            for (int i = 0; i < 3; ++i) {
                double ct = 1.0, st_ = 0.0;
                if (i > 0) sincos(y[5 + i], &st_, &ct);
                const double uix = ct * y[0] + st_ * y[1], uiy = -st_ * y[0] + ct * y[1];
                const double vx = c.ei[i] * (uix - c.ux[i]) + cn.ei[i] * cn.ux[i];
                const double vy = c.ei[i] * (uiy - c.uy[i]) + cn.ei[i] * cn.uy[i];
                sx += ct * vx - st_ * vy;
                sy += st_ * vx + ct * vy;
                sein += cn.ei[i];
            }
            y[0] = sx / sein;
            y[1] = sy / sein;
            c = cn;
        }
    }
    cost[0] = y[0] - c.ux[0];                                        // :228-235
    cost[1] = y[1] - c.uy[0];
    cost[2] = y[2];
    cost[3] = tz.uz[0];
    cost[4] = tz.uz[1];
    return total;
}

// Launch shape: problem p = variant * n + i.  variant 0: the configuration as given; variants 1..6 (mode bit 0): joint
// variant - 1 advanced by eps_q exactly as CTR_Model.jac's loop leaves q (:257-261: earlier joints went + eps - eps);
// the next 3 variants (mode bit 1): tip-force component advanced by eps_f as comp does (:270-274); the next 5 (mode
// bit 2): u_init component advanced by eps_u (shooting Jacobian).  Outputs are [rows][variants * n].
struct BvpArgs {
    const SysConst* sys;
    int n_systems;
    const double* q;        // f64 [6][n]
    const double* f;        // f64 [3][n] or null (no tip load)
    const double* u;        // f64 [5][n] or null (zeros)
    const int32_t* system;  // i32 [n] or null
    const double* params;   // f64 [18][n] per-config tube constants or null
    int64_t n;
    int32_t mode;
    double eps_q, eps_f, eps_u, max_step, rtol, atol;
    double* tip;            // f64 [3][V n]
    double* R;              // f64 [9][V n] or null
    double* cost;           // f64 [5][V n] or null
    int32_t* steps;         // i32 [V n] or null
    double* fault;          // stats slot for failed passes or null
    unsigned long long* next;  // work counter of the persistent lane-group kernel (zeroed before the launch)
    const int32_t* order;      // configuration served by scheduling rank r (k_bvp_order_*), or null: rank r is configuration r
};

__device__ __host__ inline int bvp_variants(int mode) { return 1 + ((mode & 1) ? 6 : 0) + ((mode & 2) ? 3 : 0) + ((mode & 4) ? 5 : 0); }

#ifndef CTR_BVP_BLOCK
#define CTR_BVP_BLOCK 64
#endif
#ifndef CTR_BVP_MINB
#define CTR_BVP_MINB 1
#endif
__global__ void __launch_bounds__(CTR_BVP_BLOCK, CTR_BVP_MINB) k_bvp_pass(const BvpArgs a) {
    const int V = bvp_variants(a.mode);
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= (int64_t)V * a.n) return;
    const int v = (int)(p / a.n);
    const int64_t i = p - (int64_t)v * a.n;
    double C[18], q[6], f[3] = {0.0, 0.0, 0.0}, u[5] = {0.0, 0.0, 0.0, 0.0, 0.0};
    if (a.params) {
        for (int k = 0; k < 18; ++k) C[k] = a.params[(int64_t)k * a.n + i];
    } else {
        int s = a.system ? a.system[i] : 0;
        s = min(max(s, 0), a.n_systems - 1);
        const SysConst& sc = a.sys[s];
        for (int k = 0; k < 3; ++k) {
            C[k] = sc.L[k];
            C[3 + k] = sc.Lc[k];
            C[6 + k] = sc.EI[k];
            C[9 + k] = sc.GJ[k];
            C[12 + k] = sc.Ux[k];
            C[15 + k] = sc.Uy[k];
        }
    }
    for (int k = 0; k < 6; ++k) q[k] = a.q[(int64_t)k * a.n + i];
    if (a.f)
        for (int k = 0; k < 3; ++k) f[k] = a.f[(int64_t)k * a.n + i];
    if (a.u)
        for (int k = 0; k < 5; ++k) u[k] = a.u[(int64_t)k * a.n + i];
    int w = v - 1;                                       // which perturbation this variant is
    if (v > 0 && (a.mode & 1)) {
        if (w < 6) {
            for (int k = 0; k < w; ++k) q[k] = __dsub_rn(__dadd_rn(q[k], a.eps_q), a.eps_q);
            q[w] = __dadd_rn(q[w], a.eps_q);
            w = -1;
        } else {
            w -= 6;
        }
    }
    if (w >= 0 && (a.mode & 2)) {
        if (w < 3) {
            for (int k = 0; k < w; ++k) f[k] = __dsub_rn(__dadd_rn(f[k], a.eps_f), a.eps_f);
            f[w] = __dadd_rn(f[w], a.eps_f);
            w = -1;
        } else {
            w -= 3;
        }
    }
    if (w >= 0 && (a.mode & 4) && w < 5) u[w] = __dadd_rn(u[w], a.eps_u);
    double y[NB], cost[5];
    bool finite = true;
    for (int k = 0; k < 6; ++k) finite = finite && fabs(q[k]) < CUDART_INF;
    const int st = finite ? bvp_pass(C, q, f, u, a.max_step, a.rtol, a.atol, y, cost) : -1;
    const int64_t M = (int64_t)V * a.n;
    const double bad = __longlong_as_double(0x7ff8000000000000LL);
    for (int k = 0; k < 3; ++k) a.tip[(int64_t)k * M + p] = (st < 0) ? bad : y[8 + k];
    if (a.R)
        for (int k = 0; k < 9; ++k) a.R[(int64_t)k * M + p] = (st < 0) ? bad : y[11 + k];
    if (a.cost)
        for (int k = 0; k < 5; ++k) a.cost[(int64_t)k * M + p] = (st < 0) ? bad : cost[k];
    if (a.steps) a.steps[p] = st;
    if (st < 0 && a.fault) atomicAdd(a.fault, 1.0);
}

// ------------------------------------------------------------------------------------------------------------------------
// Lane-group form of the pass (the default): like the env path, a pass is owned by 3 adjacent lanes, one per tube, 10 passes
// per warp, state in registers.  Lane j holds {u1x, u1y (replicated, identical bits on the three lanes), u_jz, alpha_j, r_j,
// R row j}: 8 values, so the six vectors of RK23 (y, f, k2, k3, y_new, k4) are 96 registers.  Per RHS the lanes exchange
// u_1z (broadcast from lane 0) and two 3-lane sums (the curvature sums of CTR_Model.py:84-112; R^T f when there is a load).
// Control flow is warp-synchronous as in fk_dp45: per-pass decisions are predicates, exchanges sit at warp-uniform points.
// The thread-per-pass kernel above is kept (CTR_BVP_THREAD_KERNEL) as the straight-line statement of the same arithmetic.
struct BvpLaneSec {
    double ei, dge, ux, uy;  // own tube on this section: EI, GJ - EI, precurvature (0 when absent / straight)
    double k;                // EI / GJ (0 when absent)
    double inv_sei;          // 1 / sum of the present tubes' EI
};

// x^(-1/3) for the step-size controller (rk.py:147 with error_exponent = -1/3, common.py:127): float MUFU seed and two Newton
// steps on y^-3 = x (y <- y (1 + (1 - x y^3) / 3), error -> 2 e^2: 1e-6 -> 2e-12 -> 8e-24), like pow_m01 of the env path.
__device__ __forceinline__ double pow_m13(double x) {
    const float xf = fminf(fmaxf((float)x, 1e-30f), 1e30f);
    double y = (double)exp2f(-0.333333333f * __log2f(xf));
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        const double e = fma(-x, y * y * y, 1.0);
        y = fma(y * (1.0 / 3.0), e, y);
    }
    return y;
}

// sin / cos of a twist angle: the env path's polynomial form below 9e4 rad, the library beyond
__device__ __forceinline__ void bvp_sincos(double x, double* s, double* c) {
    if (fabs(x) < 9.0e4) sincos_t<false>(x, s, c);
    else sincos_slow(x, s, c);
}

struct BvpLaneTube {       // the lane's own tube and what it needs of the others, in registers (no indexed constant array)
    double ei, gj, ux, uy, k;   // k = ei / gj
    double EI[3];
};

__device__ __forceinline__ void bvp_lane_consts(const BvpLaneTube& T, int pat, int j, BvpLaneSec& c, double& sei) {
    const bool present = (pat >> j) & 1, curved = (pat >> (3 + j)) & 1;
    c.ei = present ? T.ei : 0.0;
    c.dge = present ? T.gj - T.ei : 0.0;
    c.ux = curved ? T.ux : 0.0;
    c.uy = curved ? T.uy : 0.0;
    c.k = present ? T.k : 0.0;
    sei = ((pat & 1) ? T.EI[0] : 0.0) + ((pat & 2) ? T.EI[1] : 0.0) + ((pat & 4) ? T.EI[2] : 0.0);   // the order of the loop :84-112
    c.inv_sei = 1.0 / sei;
}

// S = {u1x, u1y, u_jz, alpha_j, r_j, R_j0, R_j1, R_j2} -> D; fj = the lane's component of the tip force.
// The curvature sums of CTR_Model.py:84-112 in factored form.  With u_j = Rz(alpha_j)^T u_1 in xy (:64-73), the seven terms of
// tube j in du0 collapse (u_jy cos + u_jx sin = u_1y, u_jx cos - u_jy sin = u_1x) to
//   EI_j (alpha_j' u_1y + u_jz A_j) + (GJ_j - EI_j) u_jz u_1y,   A_j = U_jy cos alpha_j + U_jx sin alpha_j
// and those in du1 to
//   EI_j (-alpha_j' u_1x + u_jz B_j) - (GJ_j - EI_j) u_jz u_1x,  B_j = U_jy sin alpha_j - U_jx cos alpha_j,
// and the twist equation :79 to (EI_j / GJ_j)(u_1x A_j + u_1y B_j): 16 FP64 operations instead of 62.  Same value to ~1e-16
// relative; the thread-per-pass kernel and oracle/bvp_oracle.c keep the reference's literal expression.
template <bool LOAD>
__device__ __forceinline__ void bvp_lane_rhs(const LaneId& id, const BvpLaneSec& c, double fj, const double (&S)[8], double (&D)[8]) {
    __syncwarp();
    xch_put2(id, 0, S[2], 0.0);                                // u_1z comes from lane 0
    __syncwarp();
    const double u1z = xch_get2a<double>(id, 0, id.l0);
    double sa = 0.0, ca = 1.0;
    if (id.j != 0) bvp_sincos(S[3], &sa, &ca);                 // alpha_1 (relative to tube 1) is identically 0
    const double uz = S[2];
    const double A = fma(ca, c.uy, sa * c.ux), B = fma(sa, c.uy, -(ca * c.ux));
    const bool on = c.ei != 0.0;
    D[2] = on ? c.k * fma(S[0], A, S[1] * B) : 0.0;            // :76-82
    D[3] = on ? S[2] - u1z : 0.0;
    const double w = c.dge * uz;
    double p0 = fma(w, S[1], c.ei * fma(uz, A, D[3] * S[1]));
    double p1 = fma(-w, S[0], c.ei * fma(uz, B, -(D[3] * S[0])));
    group_sum2(id, p0, p1);
    double g0 = 0.0, g1 = 0.0;
    if (LOAD) {                                                // R^T f: rows are spread over the lanes
        g0 = S[5] * fj;
        g1 = S[6] * fj;
        group_sum2(id, g0, g1);
    }
    D[0] = (g1 - p0) * c.inv_sei;                              // :114-118
    D[1] = -(p1 + g0) * c.inv_sei;
    D[4] = S[7];                                               // r' = R e3
    D[5] = fma(S[6], u1z, -(S[7] * S[1]));                     // R' = R u_hat(u_1), row j
    D[6] = fma(S[7], S[0], -(S[5] * u1z));
    D[7] = fma(S[5], S[1], -(S[6] * S[0]));
}

// sum over the 20 states of (x / scale)^2: lane 0 counts its 8 values, lanes 1 / 2 their 6 own ones (u1x, u1y are replicated)
__device__ __forceinline__ double bvp_lane_sq(int j, const double (&x)[8], const double (&rs)[8]) {
    double s = 0.0;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const double q = x[i] * rs[i];
        if (i >= 2 || j == 0) s = fma(q, q, s);
    }
    return s;
}

// What a lane group carries between attempts of its current pass
struct BvpLaneState {
    double y[8], fv[8];
    BvpLaneSec c;
    BvpLaneTube T;
    double fj, dtip, best, uz_tip, t, tf, h_abs;
    int64_t p;                 // output column (variant * n + configuration) of the pass being integrated
    int ksec, nsec, steps;
    bool done, have, exhausted, need_sec, rejected, failed;
};

// Loads pass p and puts the group at s = 0 of it (CTR_Model.py:22-36, :130-160); executed by the lanes of the fetching groups
// only (no exchange inside; the caller synchronises the warp before the section table is read).
template <bool LOAD>
__device__ __forceinline__ void bvp_fetch(const BvpArgs& a, int j, int64_t p, GroupScratch& gs, BvpLaneState& s) {
    // scheduling order: the variants of one configuration are consecutive, so they run in the same warp at the same time -- they
    // take (nearly) the same steps and start their sections together; the outputs keep the [rows][variant * n + i] layout
    const int V = bvp_variants(a.mode);
    const int64_t r = p / V;
    const int v = (int)(p - r * V);
    const int64_t i = a.order ? (int64_t)a.order[r] : r;
    double L[3], Lc[3], q[6], f[3] = {0.0, 0.0, 0.0}, u[5] = {0.0, 0.0, 0.0, 0.0, 0.0};
    if (a.params) {
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            L[k] = a.params[(int64_t)k * a.n + i];
            Lc[k] = a.params[(int64_t)(3 + k) * a.n + i];
            s.T.EI[k] = a.params[(int64_t)(6 + k) * a.n + i];
        }
        s.T.gj = a.params[(int64_t)(9 + j) * a.n + i];
        s.T.ux = a.params[(int64_t)(12 + j) * a.n + i];
        s.T.uy = a.params[(int64_t)(15 + j) * a.n + i];
    } else {
        int sy = a.system ? a.system[i] : 0;
        sy = min(max(sy, 0), a.n_systems - 1);
        const SysConst& sc = a.sys[sy];
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            L[k] = sc.L[k];
            Lc[k] = sc.Lc[k];
            s.T.EI[k] = sc.EI[k];
        }
        s.T.gj = sc.GJ[j];
        s.T.ux = sc.Ux[j];
        s.T.uy = sc.Uy[j];
    }
    s.T.ei = pick3(j, s.T.EI[0], s.T.EI[1], s.T.EI[2]);
    s.T.k = s.T.ei / s.T.gj;
#pragma unroll
    for (int k = 0; k < 6; ++k) q[k] = a.q[(int64_t)k * a.n + i];
    if (LOAD && a.f) {
#pragma unroll
        for (int k = 0; k < 3; ++k) f[k] = a.f[(int64_t)k * a.n + i];
    }
    if (a.u) {
#pragma unroll
        for (int k = 0; k < 5; ++k) u[k] = a.u[(int64_t)k * a.n + i];
    }
    int w = v - 1;                                       // which perturbation this variant is (see k_bvp_pass)
    if (v > 0 && (a.mode & 1)) {
        if (w < 6) {
#pragma unroll
            for (int k = 0; k < 6; ++k) {
                if (k < w) q[k] = __dsub_rn(__dadd_rn(q[k], a.eps_q), a.eps_q);
                if (k == w) q[k] = __dadd_rn(q[k], a.eps_q);
            }
            w = -1;
        } else {
            w -= 6;
        }
    }
    if (w >= 0 && (a.mode & 2)) {
        if (w < 3) {
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                if (k < w) f[k] = __dsub_rn(__dadd_rn(f[k], a.eps_f), a.eps_f);
                if (k == w) f[k] = __dadd_rn(f[k], a.eps_f);
            }
            w = -1;
        } else {
            w -= 3;
        }
    }
    if (w >= 0 && (a.mode & 4)) {
#pragma unroll
        for (int k = 0; k < 5; ++k)
            if (k == w) u[k] = __dadd_rn(u[k], a.eps_u);
    }
    build_sections<false, false>(L, Lc, q, gs, j == 0);
    s.fj = pick3(j, f[0], f[1], f[2]);
    const double a1 = q[3];
    double s1, c1;
    sincos(a1, &s1, &c1);
    s.y[0] = u[0];
    s.y[1] = u[1];
    s.y[2] = pick3(j, u[2], u[3], u[4]);
    s.y[3] = pick3(j, 0.0, q[4] - a1, q[5] - a1);              // :31: twist angles relative to tube 1
    s.y[4] = 0.0;
    s.y[5] = pick3(j, c1, s1, 0.0);
    s.y[6] = pick3(j, -s1, c1, 0.0);
    s.y[7] = pick3(j, 0.0, 0.0, 1.0);
    s.dtip = pick3(j, L[0] + q[0], L[1] + q[1], L[2] + q[2]);
    s.best = CUDART_INF;                                       // :231-235 for the lane's own tube (lanes 1, 2)
    s.uz_tip = 0.0;
    s.c.ei = s.c.dge = s.c.ux = s.c.uy = s.c.k = s.c.inv_sei = 0.0;
    s.ksec = 0;
    s.steps = 0;
    s.t = s.tf = s.h_abs = 0.0;
    s.p = (int64_t)v * a.n + i;                                 // where the results go
    // a NaN / infinite joint has no section table (every comparison of Segment.py is false): a failed pass, not an empty one
    const bool finite = fabs(q[0]) < CUDART_INF && fabs(q[1]) < CUDART_INF && fabs(q[2]) < CUDART_INF && fabs(q[3]) < CUDART_INF &&
                        fabs(q[4]) < CUDART_INF && fabs(q[5]) < CUDART_INF;
    s.done = !finite;
    s.have = true;
    s.need_sec = true;
    s.rejected = false;
    s.failed = !finite;
}

// residuals (:228-235) and the tip frame of a finished pass: lane 0 owns cost 0..2, lanes 1 / 2 the twist of their tube at its tip
__device__ __forceinline__ void bvp_store(const BvpArgs& a, int j, int64_t M, const BvpLaneState& s) {
    const double bad = __longlong_as_double(0x7ff8000000000000LL);
    const bool f = s.failed;
    const int64_t p = s.p;
    a.tip[(int64_t)j * M + p] = f ? bad : s.y[4];
    if (a.R) {
#pragma unroll
        for (int k = 0; k < 3; ++k) a.R[(int64_t)(3 * j + k) * M + p] = f ? bad : s.y[5 + k];
    }
    if (a.cost) {
        if (j == 0) {
            a.cost[p] = f ? bad : s.y[0] - s.c.ux;
            a.cost[M + p] = f ? bad : s.y[1] - s.c.uy;
            a.cost[2 * M + p] = f ? bad : s.y[2];
        } else {
            a.cost[(int64_t)(2 + j) * M + p] = f ? bad : s.uz_tip;
        }
    }
    if (j == 0) {
        if (a.steps) a.steps[p] = f ? -1 : s.steps;
        if (f && a.fault) atomicAdd(a.fault, 1.0);
    }
}

// Persistent form: every lane group takes its next pass from a global counter the moment its current one ends, so the 10 groups
// of a warp stay busy although the passes differ in length (accepted steps 50 on average, 207 at most over the reset
// distribution: waiting for the longest of 10 would cost 2.3x, measured).  Outputs are addressed by pass, so which group
// integrates which pass does not show in the results.
#ifndef CTR_BVP_GMINB
#define CTR_BVP_GMINB 3
#endif
template <bool LOAD>
__global__ void __launch_bounds__(CTR_BLOCK, CTR_BVP_GMINB) k_bvp_pass_g(const BvpArgs a) {
    __shared__ GroupScratch s_gs[(CTR_BLOCK / 32) * 11];
    const LaneId id = lane_id();
    GroupScratch& gs = s_gs[id.w * 11 + id.g];
    const int j = id.j;
    const int64_t M = (int64_t)bvp_variants(a.mode) * a.n;
    const double rtol = a.rtol, atol = a.atol, max_step = a.max_step;
    BvpLaneState s;
#pragma unroll
    for (int i = 0; i < 8; ++i) s.y[i] = s.fv[i] = 0.0;
    s.c.ei = s.c.dge = s.c.ux = s.c.uy = s.c.k = s.c.inv_sei = 0.0;
    s.T.ei = s.T.gj = s.T.ux = s.T.uy = s.T.k = s.T.EI[0] = s.T.EI[1] = s.T.EI[2] = 0.0;
    s.fj = s.dtip = s.best = s.uz_tip = s.t = s.tf = s.h_abs = 0.0;
    s.p = 0;
    s.ksec = s.nsec = s.steps = 0;
    s.done = true;
    s.have = false;
    s.exhausted = id.g >= 10;                                  // lanes 30 / 31 never own a pass
    s.need_sec = s.rejected = s.failed = false;
    for (;;) {
        if (__any_sync(FULL, s.done && !s.exhausted)) {
            // ---- finished groups store their pass and take the next one
            if (s.done && s.have) {
                bvp_store(a, j, M, s);
                s.have = false;
            }
            const bool want = s.done && !s.exhausted;
            const unsigned m = __ballot_sync(FULL, want && j == 0);
            unsigned long long base = 0;
            if (id.lane == 0) base = atomicAdd(a.next, (unsigned long long)__popc(m));
            base = __shfl_sync(FULL, base, 0);
            const int64_t pn = (int64_t)base + __popc(m & ((1u << id.l0) - 1u));
            if (want) {
                if (pn < M) bvp_fetch<LOAD>(a, j, pn, gs, s);
                else s.exhausted = true;
            }
            __syncwarp();
            if (want && !s.exhausted) s.nsec = gs.nsec & 0xff;
        }
        if (!__any_sync(FULL, !s.done)) {
            if (__any_sync(FULL, s.have)) continue;              // passes that failed at fetch are stored at the top of the loop
            break;
        }
        bool init = false;
        if (__any_sync(FULL, s.need_sec && !s.done)) {
            // ---- next section: moment continuity gives u_1xy of the new section (:186-226), executed by the whole warp
            const bool adv = s.need_sec && !s.done && s.ksec < s.nsec;
            if (s.need_sec && !s.done && s.ksec >= s.nsec) s.done = true;
            double sein = 1.0;
            BvpLaneSec cn = s.c;
            if (adv) bvp_lane_consts(s.T, gs.pat[s.ksec], j, cn, sein);
            double sa = 0.0, ca = 1.0;
            if (j != 0) bvp_sincos(s.y[3], &sa, &ca);
            const double uix = ca * s.y[0] + sa * s.y[1], uiy = -sa * s.y[0] + ca * s.y[1];
            const double vx = s.c.ei * (uix - s.c.ux) + cn.ei * cn.ux, vy = s.c.ei * (uiy - s.c.uy) + cn.ei * cn.uy;
            double sx = ca * vx - sa * vy, sy = sa * vx + ca * vy;
            group_sum2(id, sx, sy);
            if (adv) {
                if (s.ksec > 0) {
                    s.y[0] = sx / sein;
                    s.y[1] = sy / sein;
                }
                s.c = cn;
                s.t = gs.t0[s.ksec];
                s.tf = gs.tf[s.ksec];
                s.ksec += 1;
                s.need_sec = false;
                s.rejected = false;
                init = true;
            }
        }
        if (__any_sync(FULL, init)) {
            // ---- RungeKutta.__init__: f0, select_initial_step with max_step (rk.py:93-101, common.py)
            double D[8], rs[8], Y[8], D1[8];
            bvp_lane_rhs<LOAD>(id, s.c, s.fj, s.y, D);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                if (init) s.fv[i] = D[i];
                rs[i] = rcp_fast(fma(fabs(s.y[i]), rtol, atol));
            }
            if (init) {                                        // s.t[0] of the section is a stored point too
                const double dist = fabs(s.t - s.dtip);
                if (dist < s.best) {
                    s.best = dist;
                    s.uz_tip = s.y[2];
                }
            }
            double p0 = bvp_lane_sq(j, s.y, rs), p1 = bvp_lane_sq(j, D, rs);
            group_sum2(id, p0, p1);
            const double m0 = p0 * (1.0 / NB), m1 = p1 * (1.0 / NB);                   // d0^2, d1^2 of common.py:105-108
            const double interval = fabs(s.tf - s.t);
            double h0 = (m0 < 1e-10 || m1 < 1e-10) ? 1e-6 : 0.01 * sqrt_fast(m0 * rcp_fast(m1));
            h0 = dmin(h0, interval);
#pragma unroll
            for (int i = 0; i < 8; ++i) Y[i] = fma(h0, D[i], s.y[i]);
            bvp_lane_rhs<LOAD>(id, s.c, s.fj, Y, D1);
#pragma unroll
            for (int i = 0; i < 8; ++i) D1[i] = D1[i] - D[i];
            double p2 = bvp_lane_sq(j, D1, rs), dummy = 0.0;
            group_sum2(id, p2, dummy);
            const double m2 = p2 * (1.0 / NB) * rcp_fast(h0 * h0);                        // d2^2
            const double mx = dmax(m1, m2);
            const double h1 = (mx <= 1e-30) ? dmax(1e-6, h0 * 1e-3) : pow_m13(100.0 * sqrt_fast(mx));   // (0.01 / max(d1, d2))^(1/3)
            if (init) s.h_abs = dmin(dmin(100 * h0, h1), dmin(interval, max_step));
        }
        // ---- one step attempt (rk.py _step_impl, RK23 tableau)
        const double min_step = 10 * fabs(nextafter(s.t, CUDART_INF) - s.t);
        if (!s.rejected) {
            if (s.h_abs > max_step) s.h_abs = max_step;
            else if (s.h_abs < min_step) s.h_abs = min_step;
        }
        const bool too_small = s.rejected && s.h_abs < min_step;   // rk.py: "if h_abs < min_step: return False, TOO_SMALL_STEP" in the retry loop
        double t_new = s.t + s.h_abs;
        if (t_new - s.tf > 0) t_new = s.tf;
        const double h = t_new - s.t;
        const double h_try = fabs(h);
        double k2[8], k3[8], k4[8], yn[8], Y[8], ep[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) Y[i] = fma(s.fv[i] * 0.5, h, s.y[i]);
        bvp_lane_rhs<LOAD>(id, s.c, s.fj, Y, k2);
#pragma unroll
        for (int i = 0; i < 8; ++i) Y[i] = fma(k2[i] * 0.75, h, s.y[i]);
        bvp_lane_rhs<LOAD>(id, s.c, s.fj, Y, k3);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            yn[i] = fma(h, fma(k3[i], 4.0 / 9, fma(k2[i], 1.0 / 3, s.fv[i] * (2.0 / 9))), s.y[i]);
            ep[i] = fma(k3[i], -1.0 / 9, fma(k2[i], -1.0 / 12, s.fv[i] * (5.0 / 72)));   // k2, k3 end here: 16 registers fewer across the last stage
        }
        bvp_lane_rhs<LOAD>(id, s.c, s.fj, yn, k4);
        double pe = 0.0, dummy = 0.0;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const double sc = fma(fabs(absmax_arg(s.y[i], yn[i])), rtol, atol);
            const double e = fma(k4[i], 1.0 / 8, ep[i]) * h * rcp_fast(sc);   // a zero numerator sends '/' down its slow path
            if (i >= 2 || j == 0) pe = fma(e, e, pe);
        }
        group_sum2(id, pe, dummy);
        const double me = pe * (1.0 / NB);
        const double err = !(me <= 1e-200) ? sqrt_fast(me) : 0.0;   // below: the controller saturates at factor 10 long before; a NaN stays a NaN
        const double fac = 0.9 * pow_m13(err);
        if (!s.done) {
            s.h_abs = h_try;
            if (too_small) {
                s.failed = true;
                s.done = true;
            } else if (err < 1.0) {
                double factor = (err == 0.0) ? 10.0 : dmin(10.0, fac);
                if (s.rejected) factor = dmin(1.0, factor);
                s.h_abs *= factor;
                s.t = t_new;
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    s.y[i] = yn[i];
                    s.fv[i] = k4[i];
                }
                s.steps += 1;
                s.rejected = false;
                const double dist = fabs(s.t - s.dtip);
                if (dist < s.best) {
                    s.best = dist;
                    s.uz_tip = s.y[2];
                }
                if (s.t == s.tf) s.need_sec = true;
            } else {
                if (!(err == err)) {
                    s.failed = true;
                    s.done = true;
                }
                s.h_abs *= dmax(0.2, fac);
                s.rejected = true;
            }
            if (s.steps > 100000) {
                s.failed = true;
                s.done = true;
            }
        }
    }
}

// ---- scheduling order of the configurations --------------------------------------------------------------------------------
// The groups of a warp run their passes independently, but a warp whose 10 passes look alike wastes less: its sections start in the
// same iterations (the initial-step block runs for several groups at once) and its groups run dry together.  Passes that are
// handed out in an order that puts similar configurations next to each other, the long ones first, finish sooner: measured on
// 131072 reset configurations 2.17 ms in the given order, 1.59 ms sorted by the (unknowable) step count itself, 1.74 ms by the key
// used here -- few sections first (passes with few, long sections take the most steps), longer robots first among those.  Any order
// gives the same results (outputs are addressed by configuration).  Counting sort: key + histogram, scan of the 256 bins, scatter.
// many_first: the env path's solves (k_fk) cost a fixed price per section, so there the configurations with MANY sections are the long ones
__global__ void k_bvp_order_key(const BvpArgs a, int many_first, uint8_t* __restrict__ key, int32_t* __restrict__ hist) {
    __shared__ int s_hist[256];
    for (int k = threadIdx.x; k < 256; k += blockDim.x) s_hist[k] = 0;
    __syncthreads();
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < a.n) {
        double L[3], Lc[3], beta[3];
        if (a.params) {
            for (int k = 0; k < 3; ++k) {
                L[k] = a.params[(int64_t)k * a.n + i];
                Lc[k] = a.params[(int64_t)(3 + k) * a.n + i];
            }
        } else {
            int sy = a.system ? a.system[i] : 0;
            sy = min(max(sy, 0), a.n_systems - 1);
            for (int k = 0; k < 3; ++k) {
                L[k] = a.sys[sy].L[k];
                Lc[k] = a.sys[sy].Lc[k];
            }
        }
        for (int k = 0; k < 3; ++k) beta[k] = a.q[(int64_t)k * a.n + i];
        // kept sections and deployed length (Segment.py:6-61) in registers: the 29-comparator network of build_sections, no table
        double p[10], dtip[3];
        p[0] = 0.0;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            dtip[k] = __dadd_rn(L[k], beta[k]);
            p[1 + k] = beta[k];
            p[4 + k] = __dsub_rn(dtip[k], Lc[k]);
            p[7 + k] = dtip[k];
        }
        const double bmin = dmin(dmin(beta[0], beta[1]), beta[2]);
#define CS(x, y) cswap(p[x], p[y])
        CS(0, 8); CS(1, 9); CS(2, 7); CS(3, 5); CS(4, 6);
        CS(0, 2); CS(1, 4); CS(5, 8); CS(7, 9);
        CS(0, 3); CS(2, 4); CS(5, 7); CS(6, 9);
        CS(0, 1); CS(3, 6); CS(8, 9);
        CS(1, 5); CS(2, 3); CS(4, 8); CS(6, 7);
        CS(1, 2); CS(3, 5); CS(4, 6); CS(7, 8);
        CS(2, 3); CS(4, 5); CS(6, 7);
        CS(3, 4); CS(5, 6);
#undef CS
        double cum = 0.0, len = 0.0;
        int m = 0;
#pragma unroll
        for (int k = 0; k < 9; ++k) {
            const double ln = __dmul_rn(1e-5, floor(__dmul_rn(1e5, __dsub_rn(p[k + 1], p[k]))));
            cum = __dadd_rn(cum, ln);
            const double S = __dadd_rn(cum, bmin);
            if (ln != 0.0 && S > 0.0) {
                m += 1;
                len = S;
            }
        }
        if (!(bmin == bmin)) len = bmin;                                 // NaN joints: the key below sends them last
        const double lmax = fmax(fmax(L[0], L[1]), L[2]);
        int lb = 15 - (int)(16.0 * len / lmax);                         // longer first
        lb = min(max(lb, 0), 15);
        m = min(max(m, 0), 15);
        if (many_first) m = 15 - m;
        const int kk = (len == len) ? m * 16 + lb : 255;                // NaN joints last
        key[i] = (uint8_t)kk;
        atomicAdd(&s_hist[kk], 1);
    }
    __syncthreads();
    for (int k = threadIdx.x; k < 256; k += blockDim.x)
        if (s_hist[k]) atomicAdd(&hist[k], s_hist[k]);
}
__global__ void k_bvp_order_scan(int32_t* __restrict__ hist) {          // one block of 256 threads: exclusive scan in place
    __shared__ int s[256];
    const int t = threadIdx.x;
    s[t] = hist[t];
    __syncthreads();
    for (int o = 1; o < 256; o <<= 1) {
        const int v = (t >= o) ? s[t - o] : 0;
        __syncthreads();
        s[t] += v;
        __syncthreads();
    }
    hist[t] = s[t] - hist[t];
}
// Scatter without hot global atomics (most configurations fall into a handful of bins): a block ranks its keys in shared memory, reserves one
// range per bin with a single global atomic, and its threads write into the reserved ranges.
__global__ void k_bvp_order_scatter(int64_t n, const uint8_t* __restrict__ key, int32_t* __restrict__ offs, int32_t* __restrict__ order) {
    __shared__ int s_cnt[256];
    __shared__ int s_base[256];
    for (int k = threadIdx.x; k < 256; k += blockDim.x) s_cnt[k] = 0;
    __syncthreads();
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int kk = 0, rank = 0;
    if (i < n) {
        kk = key[i];
        rank = atomicAdd(&s_cnt[kk], 1);
    }
    __syncthreads();
    for (int k = threadIdx.x; k < 256; k += blockDim.x)
        if (s_cnt[k]) s_base[k] = atomicAdd(&offs[k], s_cnt[k]);
    __syncthreads();
    if (i < n) order[s_base[kk] + rank] = (int32_t)i;
}

// Forward differences over the variant block: J f64 [18][n] (3 x 6 row-major), Cm f64 [9][n] (3 x 3), tip0 f64 [3][n]
__global__ void k_bvp_fd(const double* __restrict__ tips, int64_t n, int mode, double eps_q, double eps_f, double* __restrict__ J,
                         double* __restrict__ Cm, double* __restrict__ tip0) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int V = bvp_variants(mode);
    const int64_t M = (int64_t)V * n;
    double P[3];
    for (int r = 0; r < 3; ++r) {
        P[r] = tips[(int64_t)r * M + i];
        if (tip0) tip0[(int64_t)r * n + i] = P[r];
    }
    int base = 1;
    if (mode & 1) {
        if (J)
            for (int c = 0; c < 6; ++c)
                for (int r = 0; r < 3; ++r)
                    J[(int64_t)(r * 6 + c) * n + i] = __ddiv_rn(__dsub_rn(tips[(int64_t)r * M + (int64_t)(base + c) * n + i], P[r]), eps_q);
        base += 6;
    }
    if ((mode & 2) && Cm)
        for (int c = 0; c < 3; ++c)
            for (int r = 0; r < 3; ++r)
                Cm[(int64_t)(r * 3 + c) * n + i] = __ddiv_rn(__dsub_rn(tips[(int64_t)r * M + (int64_t)(base + c) * n + i], P[r]), eps_f);
}

// One iteration of the shooting problem cost(u) = 0 on the 6-variant block (trial point + 5 perturbed u): Levenberg-Marquardt.
// The trial point is accepted when its sum of squares is below that of the last accepted point; then the forward-difference
// Jacobian of the trial point becomes the model and the damping shrinks (mu / 5).  Otherwise the damping grows (4 mu) and the
// step is recomputed from the stored model: du = -(J^T J + mu diag(J^T J))^-1 J^T c.  Convergence is judged on the max-norm
// (scipy.optimize.anderson's default, what the reference's f_tol refers to, CTR_Model.py:296).  Plain Newton with step
// halving stalls on the high-twist configurations (measured on the reference's own rows: 16 of 18 against 17 of 18 here; the
// reference's Anderson iteration itself gives up on 2 of 20).
// st: u_base f64 [5][n] | c_base [5][n] | J [25][n] | mu [n] | merit [n]
constexpr int BVP_LM_STATE = 37;
__global__ void k_bvp_lm(const double* __restrict__ cost, int64_t n, double eps_u, double f_tol, int first, int last,
                         double* __restrict__ u, double* __restrict__ st, double* __restrict__ res, int32_t* __restrict__ iters,
                         uint8_t* __restrict__ frozen, int32_t* __restrict__ active) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n || frozen[i]) return;
    const int64_t M = 6 * n;
    double* ub = st;
    double* cb = st + 5 * n;
    double* Jb = st + 10 * n;
    double* mup = st + 35 * n;
    double* mer = st + 36 * n;
    double c[5], J[5][5];
    double nrm = 0.0, ssq = 0.0;
    bool bad = false;
    for (int r = 0; r < 5; ++r) {
        c[r] = cost[(int64_t)r * M + i];
        bad |= !(c[r] == c[r]);
        nrm = fmax(nrm, fabs(c[r]));
        ssq += c[r] * c[r];
    }
    double mu;
    if (!bad && (first || ssq < mer[i])) {                   // accept the trial point: new model
        bool jbad = false;
        for (int r = 0; r < 5; ++r)
            for (int k = 0; k < 5; ++k) {
                const double cp = cost[(int64_t)r * M + (int64_t)(1 + k) * n + i];
                jbad |= !(cp == cp);
                J[r][k] = (cp - c[r]) / eps_u;
            }
        mer[i] = ssq;
        res[i] = nrm;
        for (int k = 0; k < 5; ++k) {
            ub[(int64_t)k * n + i] = u[(int64_t)k * n + i];
            cb[(int64_t)k * n + i] = c[k];
        }
        if (nrm < f_tol || last || jbad) {
            frozen[i] = 1;
            return;
        }
        for (int r = 0; r < 5; ++r)
            for (int k = 0; k < 5; ++k) Jb[(int64_t)(r * 5 + k) * n + i] = J[r][k];
        mu = first ? 1e-2 : fmax(mup[i] * 0.2, 1e-9);
        iters[i] += 1;
    } else {                                                   // reject: more damping on the stored model
        if (first) {                                           // the pass at the initial guess failed: nothing to iterate from
            res[i] = CUDART_NAN;
            frozen[i] = 1;
            return;
        }
        mu = mup[i] * 4.0;
        if (mu > 1e12 || last) {
            for (int k = 0; k < 5; ++k) u[(int64_t)k * n + i] = ub[(int64_t)k * n + i];
            frozen[i] = 1;
            return;
        }
        for (int r = 0; r < 5; ++r) {
            c[r] = cb[(int64_t)r * n + i];
            for (int k = 0; k < 5; ++k) J[r][k] = Jb[(int64_t)(r * 5 + k) * n + i];
        }
    }
    mup[i] = mu;
    double A[5][6];
    for (int a = 0; a < 5; ++a) {
        for (int b = 0; b < 5; ++b) {
            double s = 0.0;
            for (int r = 0; r < 5; ++r) s += J[r][a] * J[r][b];
            A[a][b] = s;
        }
        double g = 0.0;
        for (int r = 0; r < 5; ++r) g += J[r][a] * c[r];
        A[a][5] = -g;
    }
    for (int a = 0; a < 5; ++a) A[a][a] += mu * (A[a][a] + 1e-12);
    bool singular = false;
    for (int k = 0; k < 5; ++k) {
        int piv = k;
        for (int r = k + 1; r < 5; ++r)
            if (fabs(A[r][k]) > fabs(A[piv][k])) piv = r;
        for (int cc = 0; cc < 6; ++cc) {
            const double t = A[k][cc];
            A[k][cc] = A[piv][cc];
            A[piv][cc] = t;
        }
        const double d = A[k][k];
        if (!(fabs(d) > 1e-300)) {
            singular = true;
            break;
        }
        for (int r = k + 1; r < 5; ++r) {
            const double m = A[r][k] / d;
            for (int cc = k; cc < 6; ++cc) A[r][cc] -= m * A[k][cc];
        }
    }
    if (singular) {
        for (int k = 0; k < 5; ++k) u[(int64_t)k * n + i] = ub[(int64_t)k * n + i];
        frozen[i] = 1;
        return;
    }
    double du[5];
    for (int k = 4; k >= 0; --k) {
        double s = A[k][5];
        for (int cc = k + 1; cc < 5; ++cc) s -= A[k][cc] * du[cc];
        du[k] = s / A[k][k];
    }
    for (int k = 0; k < 5; ++k) u[(int64_t)k * n + i] = ub[(int64_t)k * n + i] + du[k];
    atomicAdd(active, 1);
}

// initial guess of CTR_Model.minimize (:279-294): u1_xy from the precurvatures of the first section, zero twist curvature
__global__ void k_bvp_guess(const BvpArgs a, double* __restrict__ u) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.n) return;
    double C[18], q[6];
    if (a.params) {
        for (int k = 0; k < 18; ++k) C[k] = a.params[(int64_t)k * a.n + i];
    } else {
        int s = a.system ? a.system[i] : 0;
        s = min(max(s, 0), a.n_systems - 1);
        const SysConst& sc = a.sys[s];
        for (int k = 0; k < 3; ++k) {
            C[k] = sc.L[k];
            C[3 + k] = sc.Lc[k];
            C[6 + k] = sc.EI[k];
            C[9 + k] = sc.GJ[k];
            C[12 + k] = sc.Ux[k];
            C[15 + k] = sc.Uy[k];
        }
    }
    for (int k = 0; k < 6; ++k) q[k] = a.q[(int64_t)k * a.n + i];
    double S[9];
    int pat[9];
    const int m = bvp_sections(C, C + 3, q, S, pat);
    BvpSection c;
    bvp_section_consts(C, m > 0 ? pat[0] : 0, c);
    double sx = 0.0, sy = 0.0, sei = 0.0;
    for (int k = 0; k < 3; ++k) {
        const double al = (k == 0) ? 0.0 : q[3 + k] - q[3];
        const double ct = cos(-al), st = sin(-al);
        sx += c.ei[k] * c.ux[k] * ct + c.ei[k] * c.uy[k] * st;
        sy += -c.ei[k] * c.ux[k] * st + c.ei[k] * c.uy[k] * ct;
        sei += c.ei[k];
    }
    u[i] = sx / sei;
    u[a.n + i] = sy / sei;
    for (int k = 2; k < 5; ++k) u[(int64_t)k * a.n + i] = 0.0;
}

}  // namespace ctr
