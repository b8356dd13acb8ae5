// ctr_device.cuh -- device-side building blocks of the B200 CTR-Reach step path (sm_100a).
//
// Execution model: one environment is owned by a GROUP of 3 adjacent lanes (one lane per tube).  Lane j holds
//   y_j = { uz_j, alpha_j, r_j, R[j][0], R[j][1], R[j][2] }
// i.e. the torsion pair of tube j, component j of the position and ROW j of the rotation matrix.  The 18-dim ODE
// of the reference (model.py:72-117) couples these slices only through
//   * the three angle differences  (one sincos per lane, the other one arrives from the neighbouring lane), and
//   * (ux_0, uy_0, uz_0) of tube 0  (broadcast from lane 0 of the group).
// so the whole Dormand-Prince state (y, 7 stage derivatives) is 6 x 8 values per lane and stays in registers
// even in float64.  A warp carries 10 environments (lanes 30/31 idle).  The lane exchanges go through a per-warp
// shared-memory mailbox (CTR_XCH_SMEM, below) at warp-uniform program points; per-environment control flow is
// expressed by predication.
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
    double der[15];                                  // EI, EI Ux, EI Uy, (EI/GJ) Ux, (EI/GJ) Uy per tube (see DER_*)
    float low[6], high[6];                           // joint Box bounds, float32 (obs.py:59-75)
    float d10, d21;                                  // f32(L1 - L0), f32(L2 - L1) (obs.py:181, NEP-50 weak python float)
    float Lf[3];                                     // f32(L_i) for the goal-sampling predicate (obs.py:198-200)
};

struct LaneId {
    int lane, g, j;            // lane in warp, group in warp, tube index (lane in group)
    int l0, lnext, lprev;      // absolute lane ids: group lane 0, tube (j+1)%3, tube (j+2)%3
    int w;                     // warp in block
};

// ---- lane exchange -------------------------------------------------------------------------------------------
// The lane exchanges of rhs() and group_sum() go through a per-warp shared-memory mailbox (STS / __syncwarp / LDS), not
// __shfl_sync: a 64-bit shuffle is two SHFL.IDX whose halves ptxas then moves into an aligned register pair (112 of the 941
// instructions of one Dormand-Prince attempt were such copies in the shuffle build of round 1, profiles/r1r_*).
#ifndef CTR_BLOCK
#define CTR_BLOCK 128
#endif
// Mailbox: rows of value PAIRS (one STS.128 / LDS.128 each in float64).
template <typename T>
struct alignas(2 * sizeof(T)) XPair {
    T a, b;
};
// rows: (uz, angle) | (cos, sin) | (ux, uy) | (group_sum partials).  All four share one base register per lane, the row is an
// immediate offset; 34 entries per row so that l0 + 2 of the idle lanes 30 / 31 stays inside the row's storage.
__shared__ XPair<double> s_xp[CTR_BLOCK / 32][4][34];
template <typename T>
__device__ __forceinline__ void xch_put2(const LaneId& id, int row, T a, T b) {
    reinterpret_cast<XPair<T>*>(&s_xp[id.w][row][0])[id.lane] = XPair<T>{a, b};
}
template <typename T>
__device__ __forceinline__ XPair<T> xch_get2(const LaneId& id, int row, int lane) {
    return reinterpret_cast<const XPair<T>*>(&s_xp[id.w][row][0])[lane];
}
template <typename T>
__device__ __forceinline__ T xch_get2a(const LaneId& id, int row, int lane) {
    return reinterpret_cast<const XPair<T>*>(&s_xp[id.w][row][0])[lane].a;
}
template <typename T>
__device__ __forceinline__ T xch_get2b(const LaneId& id, int row, int lane) {
    return reinterpret_cast<const XPair<T>*>(&s_xp[id.w][row][0])[lane].b;
}

__device__ __forceinline__ LaneId lane_id() {
    LaneId id;
    id.lane = threadIdx.x & 31;
    id.g = id.lane / GROUP;
    id.j = id.lane - id.g * GROUP;
    id.l0 = id.g * GROUP;
    id.lnext = (id.l0 + (id.j + 1) % GROUP) & 31;   // & 31: the idle lanes 30 / 31 wrap like a shuffle source lane does
    id.lprev = (id.l0 + (id.j + 2) % GROUP) & 31;
    id.w = threadIdx.x >> 5;
    return id;
}

template <typename T>
__device__ __forceinline__ T pick3(int j, T a0, T a1, T a2) {
    return j == 0 ? a0 : (j == 1 ? a1 : a2);
}

// (p0 + p1) + p2 of the three lanes' partials, identical bits on all three lanes.
template <typename T>
__device__ __forceinline__ T group_sum(const LaneId& id, T p) {
    __syncwarp();              // the previous round's readers of the row are done
    reinterpret_cast<XPair<T>*>(&s_xp[id.w][3][0])[id.lane].a = p;
    __syncwarp();
    const XPair<T>* row = reinterpret_cast<const XPair<T>*>(&s_xp[id.w][3][0]) + id.l0;
    const T p0 = row[0].a, p1 = row[1].a, p2 = row[2].a;
    return (p0 + p1) + p2;
}
// two sums in one exchange
template <typename T>
__device__ __forceinline__ void group_sum2(const LaneId& id, T& p, T& q) {
    __syncwarp();
    xch_put2(id, 3, p, q);
    __syncwarp();
    const XPair<T> r0 = xch_get2<T>(id, 3, id.l0), r1 = xch_get2<T>(id, 3, id.l0 + 1), r2 = xch_get2<T>(id, 3, id.l0 + 2);
    p = (r0.a + r1.a) + r2.a;
    q = (r0.b + r1.b) + r2.b;
}

// ---- math helpers, overloaded on the state type --------------------------------------------------------------
__device__ __forceinline__ double fma_t(double a, double b, double c) { return fma(a, b, c); }
__device__ __forceinline__ float fma_t(float a, float b, float c) { return fmaf(a, b, c); }
__device__ __forceinline__ double abs_t(double a) { return fabs(a); }
__device__ __forceinline__ float abs_t(float a) { return fabsf(a); }
__device__ __forceinline__ double max_t(double a, double b) { return fmax(a, b); }
__device__ __forceinline__ float max_t(float a, float b) { return fmaxf(a, b); }
// min / max by compare-and-select: equal to fmin / fmax whenever neither operand is NaN (the sign of a zero aside), 3 instructions
// instead of the 6 of the NaN-quieting library form (sm_100 has no FP64 min / max instruction).  A NaN operand may come out as NaN:
// every caller sits on a path where a NaN already marks the solve as failed (rk.py:129-130 analogue in fk_dp45).
__device__ __forceinline__ double dmin(double a, double b) { return a < b ? a : b; }
__device__ __forceinline__ double dmax(double a, double b) { return a > b ? a : b; }
// the operand of larger magnitude, sign included: the caller applies |.| inside its FMA, where it is a free operand modifier
// (one DSETP on |a|, |b| and two selects instead of materialising both absolute values first)
__device__ __forceinline__ double absmax_arg(double a, double b) { return fabs(a) > fabs(b) ? a : b; }
__device__ __forceinline__ float absmax_arg(float a, float b) { return fabsf(a) > fabsf(b) ? a : b; }

// 1/x for normal positive x: MUFU.RCP64H seed + two Newton steps (relative error < 2^-52; no special cases, no
// slow-path call -- the compiler's '/' costs ~3x as much and a branch).
__device__ __forceinline__ double rcp_fast(double x) {
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
    double e = fma(-x, r, 1.0);
    r = fma(r, e, r);
    e = fma(-x, r, 1.0);
    r = fma(r, e, r);
    return r;
}
// 1 / scale inside the RMS norms (rk.py:141-142, common.py:105): MUFU.RCP64H seed + ONE Newton step.  Measured on B200
// over 2^28 arguments in [6e-8, 16] (scripts/rcp_probe.cu): seed error 2^-19.9, 9.9e-13 after one step, 1.1e-16 after
// two.  The error norm only has to be exact enough that no accept / reject decision flips, i.e. |norm - 1| < 1e-12 must
// not happen: a probability of ~1e-12 per step attempt (identical nfev on every parity set, tests/test_gpu_parity.py).
__device__ __forceinline__ double rcp_1n(double x) {
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
    const double e = fma(-x, r, 1.0);
    return fma(r, e, r);
}
#ifdef CTR_RCP2
__device__ __forceinline__ double rcp_t(double a) { return rcp_fast(a); }
#else
__device__ __forceinline__ double rcp_t(double a) { return rcp_1n(a); }
#endif
__device__ __forceinline__ float rcp_t(float a) { return 1.0f / a; }

// sqrt(x) for normal positive x: MUFU.RSQ64H seed (2^-20), one Goldschmidt iteration (2^-40), residual correction
// (error squares again: correctly rounded in all but ~1e-7 of the arguments, 1 ulp there; scripts/rcp_probe.cu)
__device__ __forceinline__ double sqrt_fast(double x) {
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
    double g = x * y, h = 0.5 * y;
    double r = fma(-h, g, 0.5);
    g = fma(g, r, g);
    h = fma(h, r, h);
    r = fma(-g, g, x);
    return fma(r, h, g);
}

// x^(-1/10) for x in [1e-30, 1e30]: float MUFU seed (lg2/ex2, ~1e-6) + NEWTON steps on y^-10 = x
// (y <- y (1 + (1 - x y^10) / 10), error -> 5.5 e^2): 5.5e-12 after one step, full double precision after two.  Replaces
// the pow(err, -0.2) / pow(., 0.2) of the step-size controller (rk.py:147, common.py:127), whose argument is err^2 here.
// One step is what the controller needs: a relative 5.5e-12 in the next step size moves the next error norm by ~3e-11,
// i.e. it can flip an accept / reject decision only when that norm lies within 3e-11 of 1 -- the same order as the
// 9.9e-13 of rcp_1n below and far below the tip tolerance either way (a flipped decision moves a tip by < 3e-5 m).
// The float seed is clamped to [1e-30, 1e30] (two FMNMX); the Newton step runs on the unclamped x.  Outside that range
// the result is only used through the controller's own limits (rk.py:10-11 MIN_FACTOR / MAX_FACTOR), which saturate long
// before: x < 1e-30 gives > 1000, x > 1e30 a value below 1e-3 (possibly negative: it meets max(., 0.2) next).
template <int NEWTON>
__device__ __forceinline__ double pow_m01(double x) {
    const float xf = fminf(fmaxf((float)x, 1e-30f), 1e30f);
    double y = (double)exp2f(-0.1f * __log2f(xf));
#pragma unroll
    for (int k = 0; k < NEWTON; ++k) {
        double y2 = y * y, y4 = y2 * y2, y8 = y4 * y4, y10 = y8 * y2;
        double e = fma(-x, y10, 1.0);
        y = fma(y * 0.1, e, y);
    }
    return y;
}

// sin / cos of a double with a 2-term Cody-Waite reduction (exact first product for |n| < 2^20) and the fdlibm
// kernel polynomials on [-pi/4, pi/4]; <= 1 ulp for |x| < 1e3, <= 2 ulp up to 1e5; larger arguments take the
// library path (out of line, so the hot loop carries only a compare and a never-taken branch).  21 FP64 ops.
__device__ __noinline__ void sincos_slow(double x, double* s, double* c) { sincos(x, s, c); }

#define TC(i, v) c_trig[i]
__constant__ double c_trig[16] = {
    0.63661977236758138243, 6755399441055744.0, 1.57079632673412561417e+00, 6.07710050650619224932e-11,
    0.0, 1.58969099521155010221e-10, -2.50507602534068634195e-08, -1.13596475577881948265e-11,
    2.08757232129817482790e-09, 2.75573137070700676789e-06, -2.75573143513906633035e-07, -1.98412698298579493134e-04,
    2.48015872894767294178e-05, 8.33333333332248946124e-03, -1.38888888888741095749e-03, -1.66666666666666324348e-01};

// SLOW is a compile-time switch: a call inside the integration loop (even a never-taken one) makes ptxas guard every
// shuffle group with BRA.DIV and costs 8 % of the step, so the large-argument variant is a separate instantiation
// of the whole FK, selected once per warp (fk_group).
template <bool SLOW>
__device__ __forceinline__ void sincos_t(double x, double* s, double* c) {
    if (SLOW) {
        sincos_slow(x, s, c);
        return;
    }
    const double t = fma(x, TC(0, 0.63661977236758138243), TC(1, 6755399441055744.0));
    const int q = __double2loint(t);
    const double n = t - TC(1, 6755399441055744.0);
    // two-term Cody-Waite: n * pio2_1 (33 bits x < 2^20) is exact and so is x - n pio2_1 (Sterbenz); pio2_1t is the rest of
    // pi/2 rounded to double, its error times n stays below 1e-21
    double r = fma(-n, TC(2, 1.57079632673412561417e+00), x);
    r = fma(-n, TC(3, 6.07710050650619224932e-11), r);
    const double z = r * r;
    double ps = fma(z, TC(5, 1.58969099521155010221e-10), TC(6, -2.50507602534068634195e-08));
    double pc = fma(z, TC(7, -1.13596475577881948265e-11), TC(8, 2.08757232129817482790e-09));
    ps = fma(z, ps, TC(9, 2.75573137070700676789e-06));
    pc = fma(z, pc, TC(10, -2.75573143513906633035e-07));
    ps = fma(z, ps, TC(11, -1.98412698298579493134e-04));
    pc = fma(z, pc, TC(12, 2.48015872894767294178e-05));
    ps = fma(z, ps, TC(13, 8.33333333332248946124e-03));
    pc = fma(z, pc, TC(14, -1.38888888888741095749e-03));
    ps = fma(z, ps, TC(15, -1.66666666666666324348e-01));
    pc = fma(z, pc, 4.16666666666666019037e-02);
    pc = fma(z, pc, -0.5);
    double sn = fma(r * z, ps, r);
    double cs = fma(z, pc, 1.0);
#ifdef CTR_TRIG_FP_SIGN
    if (q & 1) {
        const double tmp = sn;
        sn = cs;
        cs = -tmp;
    }
    if (q & 2) {
        sn = -sn;
        cs = -cs;
    }
    *s = sn;
    *c = cs;
#else
    // quadrant fix-up on the integer pipe (the FP64 pipe is the bound: no DADD negations, no FP64 selects):
    // sin = (q&1 ? cs : sn) * (-1)^(q>>1),  cos = (q&1 ? sn : cs) * (-1)^((q+1)>>1)
    const bool swap = q & 1;
    const int hs = __double2hiint(sn), ls = __double2loint(sn), hc = __double2hiint(cs), lc = __double2loint(cs);
    const int s_hi = (swap ? hc : hs) ^ ((q & 2) << 30);
    const int c_hi = (swap ? hs : hc) ^ (((q + 1) & 2) << 30);
    *s = __hiloint2double(s_hi, swap ? lc : ls);
    *c = __hiloint2double(c_hi, swap ? ls : lc);
#endif
}
template <bool SLOW>
__device__ __forceinline__ void sincos_t(float x, float* s, float* c) { sincosf(x, s, c); }

// Dormand-Prince 5(4) tableau (rk.py: class RK45) in constant memory: used as c[bank][offset] operands, so the
// hot loop materialises no immediates (the previous build spent 18 % of its issue slots on UMOV for these).
enum { A21 = 0, A31, A32, A41, A42, A43, A51, A52, A53, A54, A61, A62, A63, A64, A65, B1, B3, B4, B5, B6, E1, E3, E4, E5,
       E6, E7, DP_NCOEF };
__constant__ double c_dp64[DP_NCOEF] = {
    1.0 / 5, 3.0 / 40, 9.0 / 40, 44.0 / 45, -56.0 / 15, 32.0 / 9, 19372.0 / 6561, -25360.0 / 2187, 64448.0 / 6561,
    -212.0 / 729, 9017.0 / 3168, -355.0 / 33, 46732.0 / 5247, 49.0 / 176, -5103.0 / 18656, 35.0 / 384, 500.0 / 1113,
    125.0 / 192, -2187.0 / 6784, 11.0 / 84, -71.0 / 57600, 71.0 / 16695, -71.0 / 1920, 17253.0 / 339200, -22.0 / 525,
    1.0 / 40};
__constant__ float c_dp32[DP_NCOEF] = {
    (float)(1.0 / 5), (float)(3.0 / 40), (float)(9.0 / 40), (float)(44.0 / 45), (float)(-56.0 / 15), (float)(32.0 / 9),
    (float)(19372.0 / 6561), (float)(-25360.0 / 2187), (float)(64448.0 / 6561), (float)(-212.0 / 729),
    (float)(9017.0 / 3168), (float)(-355.0 / 33), (float)(46732.0 / 5247), (float)(49.0 / 176), (float)(-5103.0 / 18656),
    (float)(35.0 / 384), (float)(500.0 / 1113), (float)(125.0 / 192), (float)(-2187.0 / 6784), (float)(11.0 / 84),
    (float)(-71.0 / 57600), (float)(71.0 / 16695), (float)(-71.0 / 1920), (float)(17253.0 / 339200), (float)(-22.0 / 525),
    (float)(1.0 / 40)};
// Dense-output matrix of RK45 (rk.py: RK45.P, 7 stages x 4 powers; row 1 is zero), used only by the backbone-shape variant.
__constant__ double c_dpP[28] = {
    1, -8048581381.0 / 2820520608.0, 8663915743.0 / 2820520608.0, -12715105075.0 / 11282082432.0,
    0, 0, 0, 0,
    0, 131558114200.0 / 32700410799.0, -68118460800.0 / 10900136933.0, 87487479700.0 / 32700410799.0,
    0, -1754552775.0 / 470086768.0, 14199869525.0 / 1410260304.0, -10690763975.0 / 1880347072.0,
    0, 127303824393.0 / 49829197408.0, -318862633887.0 / 49829197408.0, 701980252875.0 / 199316789632.0,
    0, -282668133.0 / 205662961.0, 2019193451.0 / 616988883.0, -1453857185.0 / 822651844.0,
    0, 40617522.0 / 29380423.0, -110615467.0 / 29380423.0, 69997945.0 / 29380423.0};
constexpr int SHAPE_SAMPLES = 30;      // model.py:143: np.linspace(span[k], span[k+1] - 1e-6, num=30)

// Where the backbone samples of this lane's position component go: sample p of the env at base[p * stride], p < max_pts.
struct ShapeOut {
    double* base;
    int64_t stride;
    int max_pts;
};
// abscissa i of np.linspace(a, b, 30): i * step + a (two roundings), the last one is b itself
__device__ __forceinline__ double shape_abscissa(int i, double a, double step, double b) {
    return (i == SHAPE_SAMPLES - 1) ? b : __dadd_rn(__dmul_rn((double)i, step), a);
}


template <typename T> __device__ __forceinline__ T dpc(int i);
template <> __device__ __forceinline__ double dpc<double>(int i) { return c_dp64[i]; }
template <> __device__ __forceinline__ float dpc<float>(int i) { return c_dp32[i]; }

// ---- section table (Segment.py:6-61) -- IEEE float64, explicit _rn intrinsics so nothing is contracted -------
// Built once per FK by the lane group into its shared-memory slot: the integration spans [t0, tf] of the kept
// sections (model.py:137-151) and, per section, which tubes are present / curved.  The main loop then only indexes it.
constexpr int MAXSEC = 9;
enum { DER_EI = 0, DER_EIUX = 3, DER_EIUY = 6, DER_KX = 9, DER_KY = 12, DER_COUNT = 15 };

struct GroupScratch {            // 312 B: 78 words, so the 10 groups of a warp hit distinct bank pairs
    double t0[MAXSEC], tf[MAXSEC];
    int pat[MAXSEC];             // bits 0-2: tube present, bits 3-5: tube curved
    int nsec;                    // kept sections | (zero-length spans << 8)
    double der[DER_COUNT];       // EI_m, EI_m Ux_m, EI_m Uy_m, (EI_m / GJ_m) Ux_m, (EI_m / GJ_m) Uy_m
    double pad;
};

__device__ __forceinline__ void cswap(double& a, double& b) {
    const double lo = dmin(a, b), hi = dmax(b, a);   // one of each for every pair of values, NaN or not
    a = lo;
    b = hi;
}

// L, Lc: tube lengths / curved lengths; beta: float64 joint extensions.  All lanes of the group compute the same
// values; `writer` (one lane) stores them.  Ends with __syncwarp() so the group can read the table.
// GAP = false: the spans of the BVP model, [S_k-1, S_k] without the 1e-6 gap of model.py:141 (CTR_Model.py:164-167).
// SYNC = false: called by some groups of the warp only; the caller synchronises before the table is read.
template <bool GAP = true, bool SYNC = true>
__device__ __forceinline__ void build_sections(const double* L, const double* Lc, const double* beta, GroupScratch& gs,
                                               bool writer) {
    double p[10], dtip[3], dc[3];
    p[0] = 0.0;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        dtip[i] = __dadd_rn(L[i], beta[i]);        // Segment.py:12
        dc[i] = __dsub_rn(dtip[i], Lc[i]);         // Segment.py:13
        p[1 + i] = beta[i];
        p[4 + i] = dc[i];
        p[7 + i] = dtip[i];
    }
    const double bmin = dmin(dmin(beta[0], beta[1]), beta[2]);   // Segment.py:53 min(base)
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
    double cum = 0.0, prevS = 0.0;
    int m = 0, skipped = 0;
#pragma unroll
    for (int k = 0; k < 9; ++k) {
        const double lo = p[k], hi = p[k + 1];
        const double len = __dmul_rn(1e-5, floor(__dmul_rn(1e5, __dsub_rn(hi, lo))));   // Segment.py:16
        cum = __dadd_rn(cum, len);                                                       // Segment.py:52 (len == 0 adds 0)
        const double S = __dadd_rn(cum, bmin);                                           // Segment.py:53
        if (len != 0.0 && S > 0.0) {                                                     // Segment.py:46,53
            int pat = 0;
#pragma unroll
            for (int i = 0; i < 3; ++i) {
                const bool present = (beta[i] <= lo) && (hi <= dtip[i]);                 // Segment.py:23-41 (value form)
                const bool curved = present && (dc[i] <= lo);                            // Segment.py:42-43
                pat |= (present ? 1 : 0) << i;
                pat |= (curved ? 8 : 0) << i;
            }
            double t0 = prevS;
            double tf = GAP ? __dsub_rn(S, 1e-6) : S;                                    // model.py:141
            if (GAP && tf < t0) {                                                        // model.py:145-151: sorted span
                const double z = t0;
                t0 = tf;
                tf = z;
            }
            prevS = S;
            if (tf == t0) {
                skipped += 1;                  // zero-length span: the solver returns y0 after one RHS call (base.py)
            } else {
                if (writer) {
                    gs.t0[m] = t0;
                    gs.tf[m] = tf;
                    gs.pat[m] = pat;
                }
                m += 1;
            }
        }
    }
    if (writer) gs.nsec = m | (skipped << 8);
    if (SYNC) __syncwarp();
}

// The dense-output samples are not evaluated inside the divergent accept branch (a warp would run that loop for the largest
// number of samples any of its lanes' steps covers, every pass: 60 % of the instructions of the first form of k_fk_shape).
// An accepted step only LOGS what its interpolant needs -- t_old, t_new, r_j(t_old), Q = K^T P for r_j and 1 / h, 8 doubles -- in a
// per-lane log (local memory, SHAPE_LOG_CAP steps); the samples are evaluated afterwards in one warp-uniform loop in which every
// lane walks its own log and its own abscissae (shape_flush: when some lane's log is full and after the last step), so the trip
// count is the largest number of samples + steps of a LANE, not the sum over the passes of the largest per-step count.
constexpr int SHAPE_LOG_CAP = 24;
constexpr int SHAPE_LOG_W = 8;
#ifndef CTR_SHAPE_EMIT
#define CTR_SHAPE_EMIT 2
#endif
constexpr int SHAPE_EMIT = CTR_SHAPE_EMIT;
struct ShapeCursor {
    int sec, i;        // next sample to evaluate: abscissa i of section sec
    int n;             // steps in the log
};
template <bool ON>
struct ShapeLog {
    double e[ON ? SHAPE_LOG_CAP * SHAPE_LOG_W : 1];
};
template <bool ON>
__device__ __forceinline__ void shape_flush(const GroupScratch& gs, int nsec, const ShapeOut& shape, ShapeLog<ON>& lg, ShapeCursor& cur) {
    int s = -1;                                                                  // log entry held in registers
    double a = 0.0, b = 0.0, step = 0.0, te = 0.0;
    double t_old = 0.0, t_new = -1e300, h_d = 0.0, inv_h = 0.0, y_old = 0.0, q0 = 0.0, q1 = 0.0, q2 = 0.0, q3 = 0.0;
    bool have_sec = false;
    double* out = shape.base + (int64_t)(cur.sec * SHAPE_SAMPLES + cur.i) * shape.stride;
    int pt = cur.sec * SHAPE_SAMPLES + cur.i;
    // every iteration either evaluates the lane's next sample or takes its next step from the log
    while (__any_sync(FULL, s < cur.n)) {
        if (s < cur.n) {
            if (!have_sec && cur.sec < nsec) {                                   // abscissae of the section the cursor is in (none after the last)
                a = gs.t0[cur.sec];
                b = gs.tf[cur.sec];
                step = __ddiv_rn(__dsub_rn(b, a), (double)(SHAPE_SAMPLES - 1));
                te = shape_abscissa(cur.i, a, step, b);
                have_sec = true;
            }
            if (have_sec && te <= t_new) {                                       // ivp.py: searchsorted(t_eval, t, 'right')
                // up to SHAPE_EMIT samples of this step per trip: most trips are samples (148 against 13 steps per lane on average),
                // so the loop control (vote, branches, cursor bookkeeping) is shared between them
#pragma unroll
                for (int rep = 0; rep < SHAPE_EMIT; ++rep) {
                    if (rep > 0 && !(have_sec && te <= t_new)) break;
                    const double x = __dmul_rn(__dsub_rn(te, t_old), inv_h);     // differs from numpy's (te - t) / h by <= 1 ulp
                    const double x2 = __dmul_rn(x, x), x3 = __dmul_rn(x2, x), x4 = __dmul_rn(x3, x);   // np.cumprod
                    double acc = __dmul_rn(q0, x);
                    acc = __dadd_rn(acc, __dmul_rn(q1, x2));
                    acc = __dadd_rn(acc, __dmul_rn(q2, x3));
                    acc = __dadd_rn(acc, __dmul_rn(q3, x4));
                    if (pt < shape.max_pts) *out = __dadd_rn(__dmul_rn(h_d, acc), y_old);
                    out += shape.stride;
                    pt += 1;
                    cur.i += 1;
                    if (cur.i == SHAPE_SAMPLES) {
                        cur.i = 0;
                        cur.sec += 1;
                        have_sec = false;
                    } else {
                        te = shape_abscissa(cur.i, a, step, b);
                    }
                }
            } else {                                                             // next step of the log
                s += 1;
                if (s < cur.n) {
                    const double* L = lg.e + s * SHAPE_LOG_W;
                    t_old = L[0];
                    t_new = L[1];
                    y_old = L[2];
                    q0 = L[3];
                    q1 = L[4];
                    q2 = L[5];
                    q3 = L[6];
                    inv_h = L[7];
                    h_d = t_new - t_old;
                }
            }
        }
    }
    cur.n = 0;
}

// ---- per-section constants in the lane-rotated form -----------------------------------------------------------
template <typename T>
struct Section {
    T Ao, Bo;   // A_j, B_j of the lane's own tube:   A_m = EI_m Ux_m / sum(EI),  B_m = EI_m Uy_m / sum(EI)
    T An, Bn;   // tube (j+1)%3
    T Ap, Bp;   // tube (j+2)%3
    T kx, ky;   // (EI_j / GJ_j) * Ux_j , (EI_j / GJ_j) * Uy_j   (0 when tube j is absent or straight: model.py:97-101)
    bool present;  // tube j exists on this section (alpha_j' = uz_j), else alpha_j' = 0
};

// ---- ODE right-hand side (model.py:72-117), one lane's slice ------------------------------------------------
// Y = {uz_j, alpha_j, r_j, R_j0, R_j1, R_j2} -> D = d/ds of the same.  Executed by the full warp.
// The sin / cos of the two angle differences a lane uses -- (co, so) of alpha_j - alpha_{j+1}, computed by the lane, and
// (cp, sp) of alpha_{j-1} - alpha_j, computed by lane j-1 -- stay in row 1 of the mailbox after every call.  When a new
// section starts at the state the last evaluation was made at (FSAL), only the algebra is redone with the new constants:
// do_trig = false re-reads the row instead of carrying four doubles per lane through the whole integration loop.

// The twist integrates to a few radians at most, so the argument range of the sincos is decided once per FK from the
// initial angle differences: true when some environment of this warp needs the library (large-argument) path.
__device__ __forceinline__ bool needs_slow_trig(int j, const double* alpha) {
    const double d = pick3(j, alpha[0] - alpha[1], alpha[1] - alpha[2], alpha[2] - alpha[0]);
    return __any_sync(FULL, !(fabs(d) < 9.0e4));
}

// UY0: every tube has y_curvature == 0 (all four systems of the reference, __init__.py:9-70; domain randomisation never
// touches U_y, model_utils.py:12-20): the B terms and ky vanish -- 5 FP64 operations and 4 registers less per call.
template <typename T, bool SLOW, bool UY0>
__device__ __forceinline__ void rhs(const LaneId& id, const Section<T>& sc, bool do_trig, const T (&Y)[6], T (&D)[6]) {
    // Pair rows 0 | 1 | 2 are written in that order with a warp barrier after each, and every read of a row sits before the
    // barrier that precedes the row's next write (uz_0 of row 0 is therefore fetched before the last barrier of the call).
    xch_put2(id, 0, Y[0], Y[1]);
    __syncwarp();
    T co, so;
    if (do_trig) {   // warp-uniform
        const T a_next = xch_get2b<T>(id, 0, id.lnext);
        sincos_t<SLOW>(Y[1] - a_next, &so, &co);
        xch_put2(id, 1, co, so);
        __syncwarp();
    } else {
        const XPair<T> own = xch_get2<T>(id, 1, id.lane);
        co = own.a;
        so = own.b;
    }
    const XPair<T> pr = xch_get2<T>(id, 1, id.lprev);
    const T cp = pr.a, sp = pr.b;
    // model.py:87-94 with cos(0) = 1, sin(0) = 0 folded; sin(alpha_j - alpha_{j-1}) = -sp
    T ux, uy;
    if (UY0) {
        ux = fma_t(sc.An, co, sc.Ao);
        uy = -sc.An * so;
        ux = fma_t(sc.Ap, cp, ux);
        uy = fma_t(sc.Ap, sp, uy);
        D[0] = -(uy * sc.kx);                                // model.py:101 with U_y = 0
    } else {
        ux = fma_t(sc.An, co, sc.Ao);
        uy = fma_t(-sc.An, so, sc.Bo);
        ux = fma_t(sc.Bn, so, ux);
        uy = fma_t(sc.Bn, co, uy);
        ux = fma_t(sc.Ap, cp, ux);
        uy = fma_t(sc.Ap, sp, uy);
        ux = fma_t(-sc.Bp, sp, ux);
        uy = fma_t(sc.Bp, cp, uy);
        D[0] = fma_t(ux, sc.ky, -(uy * sc.kx));              // model.py:101  (0 when absent: kx = ky = 0)
    }
    D[1] = sc.present ? Y[0] : (T)0;                         // model.py:99,102
    const T uz0 = xch_get2a<T>(id, 0, id.l0);
    xch_put2(id, 2, ux, uy);
    __syncwarp();
    const XPair<T> u0 = xch_get2<T>(id, 2, id.l0);
    const T ux0 = u0.a, uy0 = u0.b;
    D[2] = Y[5];                                         // r' = R e3          (model.py:108)
    D[3] = fma_t(Y[4], uz0, -(Y[5] * uy0));              // R' = R u_hat, row j (model.py:107,109)
    D[4] = fma_t(Y[5], ux0, -(Y[3] * uz0));
    D[5] = fma_t(Y[3], uy0, -(Y[4] * ux0));
}

// 10 * |nextafter(t, +inf) - t| (rk.py:116) for the t of this path (non-negative except the rare reversed first span)
__device__ __forceinline__ double min_step_of(double t) {
    const int e = __double2hiint(t) & 0x7ff00000;
    const int u = e > 0x03400000 ? e - 0x03400000 : 0;
    return 10.0 * __hiloint2double(u, 0);
}

struct FkInfo {
    int nfev, nacc, nrej, nsec, failed;
    int shape_secs;   // sections whose 30 backbone samples were written (SHAPE variant only)
};

// ---- forward kinematics (model.py:30-70, 119-174) for the lane group ----------------------------------------
// Control flow is warp-synchronous: every shuffle sits at a warp-uniform program point; per-environment decisions
// (new section, accept / reject, finished) are predicated.  The stages are unrolled (a rolled stage loop makes ptxas
// copy the whole k-array set at every stage), so rhs() is kept small: no library calls, no slow paths inlined.

template <typename T>
__device__ __forceinline__ void initial_state(int j, const double* alpha, T (&y)[6]) {
    // model.py:56-64: uz = 0, alpha = joint[3:], r = 0, R = Rz(alpha_0)
    double s0, c0;
    sincos(alpha[0], &s0, &c0);
    y[0] = (T)0;
    y[1] = (T)pick3(j, alpha[0], alpha[1], alpha[2]);
    y[2] = (T)0;
    y[3] = (T)pick3(j, c0, s0, 0.0);
    y[4] = (T)pick3(j, -s0, c0, 0.0);
    y[5] = (T)pick3(j, 0.0, 0.0, 1.0);
}

// Next kept section of this lane's environment (model.py:137-151).  Returns true when a section was started.
template <typename T>
__device__ __forceinline__ bool advance_section(const GroupScratch& gs, const double* der, int nsec, int mo, int mn,
                                                int mp, int& ksec, bool& done, double& t, double& tf, Section<T>& sc) {
    if (ksec >= nsec) {
        done = true;
        return false;
    }
    t = gs.t0[ksec];
    tf = gs.tf[ksec];
    const int pat = gs.pat[ksec];
    ksec += 1;
    const double e0 = (pat & 1) ? der[DER_EI + 0] : 0.0;
    const double e1 = (pat & 2) ? der[DER_EI + 1] : 0.0;
    const double e2 = (pat & 4) ? der[DER_EI + 2] : 0.0;
    const double inv = rcp_fast((e0 + e1) + e2);                     // model.py:87 (a kept section has >= 1 tube)
    const bool co = (pat >> (3 + mo)) & 1, cn = (pat >> (3 + mn)) & 1, cp = (pat >> (3 + mp)) & 1;
    sc.Ao = (T)(co ? der[DER_EIUX + mo] * inv : 0.0);
    sc.Bo = (T)(co ? der[DER_EIUY + mo] * inv : 0.0);
    sc.An = (T)(cn ? der[DER_EIUX + mn] * inv : 0.0);
    sc.Bn = (T)(cn ? der[DER_EIUY + mn] * inv : 0.0);
    sc.Ap = (T)(cp ? der[DER_EIUX + mp] * inv : 0.0);
    sc.Bp = (T)(cp ? der[DER_EIUY + mp] * inv : 0.0);
    sc.kx = (T)(co ? der[DER_KX + mo] : 0.0);
    sc.ky = (T)(co ? der[DER_KY + mo] : 0.0);
    sc.present = ((pat >> mo) & 1) != 0;
    return true;
}

// INTEG 0: scipy RK45 restated (rk.py rk_step/_step_impl, common.py select_initial_step), per-env adaptive.
// On return lane j holds y = {uz_j, alpha_j, r_j, R_j0..2} at the tip.
// SHAPE: additionally sample r_j at the 30 abscissae of every section through scipy's dense output (rk.py RkDenseOutput:
// y_old + h * Q . [x, x^2, x^3, x^4], Q = K^T P, x = (s - t_old) / h), as solve_ivp(t_eval=...) does for Model.r
// (model.py:143-158, ivp.py t_eval handling: the samples in (t_old, t_new] after every accepted step).
//
// Register budget (128 per thread at 4 blocks / SM): y, f, k2..k6 and the stage argument are 8 x 6 doubles = 96 registers, so
// everything else that lives across the loop is kept minimal: the loop counts attempts, rejections and started sections
// only (nfev = 6 attempts + 2 started + zero-length spans is derived at the end), the trigonometry of the last evaluation
// stays in the mailbox (rhs), `present` is a predicate.
template <typename T, bool SLOW, bool UY0, bool SHAPE = false>
__device__ __forceinline__ void fk_dp45(const LaneId& id, bool valid, const GroupScratch& gs, const double* der,
                                        const double* alpha, double rtol_d, double atol_d, T (&y)[6], FkInfo& info,
                                        const ShapeOut shape = ShapeOut{nullptr, 0, 0}) {
    const int j = id.j;
    const int mn = (j == 2) ? 0 : j + 1, mp = (j == 0) ? 2 : j - 1;
    initial_state<T>(j, alpha, y);
    const int packed = valid ? gs.nsec : 0;
    const int nsec = packed & 0xff;
    const T rtol = (T)rtol_d, atol = (T)atol_d;
    Section<T> sc;
    sc.Ao = sc.Bo = sc.An = sc.Bn = sc.Ap = sc.Bp = sc.kx = sc.ky = (T)0;
    sc.present = false;
    T f[6];
#pragma unroll
    for (int i = 0; i < 6; ++i) f[i] = (T)0;
    int ksec = 0, attempts = 0, nrej = 0;
    double t = 0.0, tf = 0.0, h_abs = 0.0;
    bool done = !valid, need_sec = true, rejected = false, failed = false;
    bool first = true;   // warp-uniform: no RHS has been evaluated yet, so slot 0 needs the trigonometry
    int sh_sec = -1;                                // SHAPE: sections started
    ShapeLog<SHAPE> sh_log;
    ShapeCursor sh_cur{0, 0, 0};
    info.shape_secs = 0;

    for (;;) {
        bool init = false;
        if (__any_sync(FULL, need_sec && !done)) {
            if (need_sec && !done) {
                init = advance_section<T>(gs, der, nsec, j, mn, mp, ksec, done, t, tf, sc);
                if (init) {
                    need_sec = false;
                    rejected = false;
                    if (SHAPE) sh_sec += 1;
                }
            }
        }
        if (!__any_sync(FULL, !done)) break;
        const bool any_init = __any_sync(FULL, init);
        const bool active = !done;

        // ---- RungeKutta.__init__: f0 and select_initial_step (rk.py:93-101, common.py:68-134); run by the whole warp
        //      whenever at least one environment starts a section (its FSAL derivative belongs to the old constants)
        if (any_init) {
            T D[6], rs[6], Y[6];
            rhs<T, SLOW, UY0>(id, sc, first, y, D);                    // f0 = f(t0, y0); trigonometry reused unless first
            T p0 = 0, p1 = 0;
#pragma unroll
            for (int i = 0; i < 6; ++i) {
                if (init) f[i] = D[i];
                rs[i] = rcp_t(fma_t(abs_t(y[i]), rtol, atol));     // 1 / scale
                const T a = y[i] * rs[i], b = D[i] * rs[i];
                p0 = fma_t(a, a, p0);
                p1 = fma_t(b, b, p1);
            }
            group_sum2(id, p0, p1);
            const double m0 = (double)p0 * (1.0 / 18.0);                  // d0^2 (common.py:63-65 RMS norm)
            const double m1 = (double)p1 * (1.0 / 18.0);                  // d1^2
            const double interval = fabs(tf - t);
            double h0 = (m0 < 1e-10 || m1 < 1e-10) ? 1e-6 : 0.01 * sqrt_fast(m0 * rcp_1n(m1));
            h0 = dmin(h0, interval);
            const T h0t = (T)h0;
#pragma unroll
            for (int i = 0; i < 6; ++i) Y[i] = fma_t(h0t, f[i], y[i]);   // common.py:113
            T D1[6];
            rhs<T, SLOW, UY0>(id, sc, true, Y, D1);
            T p2 = 0;
#pragma unroll
            for (int i = 0; i < 6; ++i) {
                const T c = (D1[i] - f[i]) * rs[i];
                p2 = fma_t(c, c, p2);
            }
            const double m2 = (double)group_sum(id, p2) * (1.0 / 18.0) * rcp_1n(h0 * h0);   // d2^2
            const double M = dmax(m1, m2);
            // h1 = (0.01 / max(d1, d2))^(1/5) = (1e4 M)^(-1/10)
            const double h1 = (M <= 1e-30) ? dmax(1e-6, h0 * 1e-3) : pow_m01<1>(M * 1e4);
            if (init) h_abs = dmin(dmin(100.0 * h0, h1), interval);
        }

        // ---- one step attempt (rk.py:108-178)
        const double min_step = min_step_of(t);
        if (!rejected && h_abs < min_step) h_abs = min_step;
        double t_new = t + h_abs;
        if (t_new - tf > 0.0) t_new = tf;
        const double h_d = t_new - t;
        const T h = (T)h_d;
        h_abs = fabs(h_d);                                  // rk.py:136-137: the factor scales the step actually taken
        T k2[6], k3[6], k4[6], k5[6], k6[6], k7[6], Y[6];
        {
            const T c1 = h * dpc<T>(A21);
#pragma unroll
            for (int i = 0; i < 6; ++i) Y[i] = fma_t(c1, f[i], y[i]);
            rhs<T, SLOW, UY0>(id, sc, true, Y, k2);
        }
        {
            const T c1 = h * dpc<T>(A31), c2 = h * dpc<T>(A32);
#pragma unroll
            for (int i = 0; i < 6; ++i) Y[i] = fma_t(c2, k2[i], fma_t(c1, f[i], y[i]));
            rhs<T, SLOW, UY0>(id, sc, true, Y, k3);
        }
        {
            const T c1 = h * dpc<T>(A41), c2 = h * dpc<T>(A42), c3 = h * dpc<T>(A43);
#pragma unroll
            for (int i = 0; i < 6; ++i) Y[i] = fma_t(c3, k3[i], fma_t(c2, k2[i], fma_t(c1, f[i], y[i])));
            rhs<T, SLOW, UY0>(id, sc, true, Y, k4);
        }
        {
            const T c1 = h * dpc<T>(A51), c2 = h * dpc<T>(A52), c3 = h * dpc<T>(A53), c4 = h * dpc<T>(A54);
#pragma unroll
            for (int i = 0; i < 6; ++i)
                Y[i] = fma_t(c4, k4[i], fma_t(c3, k3[i], fma_t(c2, k2[i], fma_t(c1, f[i], y[i]))));
            rhs<T, SLOW, UY0>(id, sc, true, Y, k5);
        }
        {
            const T c1 = h * dpc<T>(A61), c2 = h * dpc<T>(A62), c3 = h * dpc<T>(A63), c4 = h * dpc<T>(A64),
                    c5 = h * dpc<T>(A65);
#pragma unroll
            for (int i = 0; i < 6; ++i)
                Y[i] = fma_t(c5, k5[i], fma_t(c4, k4[i], fma_t(c3, k3[i], fma_t(c2, k2[i], fma_t(c1, f[i], y[i])))));
            rhs<T, SLOW, UY0>(id, sc, true, Y, k6);
        }
        T e5[6];   // the k1..k6 part of the error estimate (k2 has weight 0); k2..k5 die here
        {
            const T c1 = h * dpc<T>(B1), c3 = h * dpc<T>(B3), c4 = h * dpc<T>(B4), c5 = h * dpc<T>(B5), c6 = h * dpc<T>(B6);
#pragma unroll
            for (int i = 0; i < 6; ++i) {
                Y[i] = fma_t(c6, k6[i], fma_t(c5, k5[i], fma_t(c4, k4[i], fma_t(c3, k3[i], fma_t(c1, f[i], y[i])))));
                e5[i] = fma_t(dpc<T>(E6), k6[i], fma_t(dpc<T>(E5), k5[i], fma_t(dpc<T>(E4), k4[i],
                              fma_t(dpc<T>(E3), k3[i], dpc<T>(E1) * f[i]))));
            }
            rhs<T, SLOW, UY0>(id, sc, true, Y, k7);                    // f(t + h, y_new)
        }
        // ---- error norm and step-size control (rk.py:139-159)
        T pe = 0;
#pragma unroll
        for (int i = 0; i < 6; ++i) {
            const T e = fma_t(dpc<T>(E7), k7[i], e5[i]);
            const T sc_i = fma_t(abs_t(absmax_arg(y[i], Y[i])), rtol, atol);         // rk.py:141
            const T q = e * rcp_t(sc_i);
            pe = fma_t(q, q, pe);
        }
        const double m = (double)group_sum(id, pe) * (h_d * h_d) * (1.0 / 18.0);    // error_norm^2 (rk.py:142)
        const bool accept = m < 1.0;                                                // rk.py:144
        // rk.py:145-158: accepted: min(MAX_FACTOR, SAFETY norm^-1/5), at most 1 after a rejection; rejected: max(MIN_FACTOR, .).
        // One expression serves both: an accepted step has fac > 0.9 (the lower limit is idle), a rejected one fac <= 0.9 (the
        // upper limit is idle); norm == 0 gives fac > 900, i.e. MAX_FACTOR as rk.py:146 states it; a NaN fac gives MIN_FACTOR.
        const double fac = 0.9 * pow_m01<1>(m);                                     // SAFETY * norm^(-1/5)
        const double factor = dmin(dmax(fac, 0.2), rejected ? 1.0 : 10.0);
        if (active) {
            attempts += 1;
            h_abs *= factor;
            if (accept) {
                if (SHAPE) {
                    // Q for the lane's position component r_j (index 2): Q[p] = sum_s K_s[2] P[s][p]
                    double Q[4];
#pragma unroll
                    for (int pw = 0; pw < 4; ++pw) {
                        double q = __dmul_rn((double)f[2], c_dpP[pw]);
                        q = __dadd_rn(q, __dmul_rn((double)k2[2], c_dpP[4 + pw]));
                        q = __dadd_rn(q, __dmul_rn((double)k3[2], c_dpP[8 + pw]));
                        q = __dadd_rn(q, __dmul_rn((double)k4[2], c_dpP[12 + pw]));
                        q = __dadd_rn(q, __dmul_rn((double)k5[2], c_dpP[16 + pw]));
                        q = __dadd_rn(q, __dmul_rn((double)k6[2], c_dpP[20 + pw]));
                        Q[pw] = __dadd_rn(q, __dmul_rn((double)k7[2], c_dpP[24 + pw]));
                    }
                    double* L = sh_log.e + sh_cur.n * SHAPE_LOG_W;           // sh_cur.n < SHAPE_LOG_CAP: a full log is flushed below
                    L[0] = t;
                    L[1] = t_new;
                    L[2] = (double)y[2];
                    L[3] = Q[0];
                    L[4] = Q[1];
                    L[5] = Q[2];
                    L[6] = Q[3];
                    L[7] = __ddiv_rn(1.0, h_d);                                  // one division per accepted step
                    sh_cur.n += 1;
                    info.shape_secs = sh_sec + 1;
                }
#pragma unroll
                for (int i = 0; i < 6; ++i) {
                    y[i] = Y[i];
                    f[i] = k7[i];                                                   // FSAL (rk.py:170)
                }
                t = t_new;
                rejected = false;
                if (t == tf) need_sec = true;
            } else {
                rejected = true;
                nrej += 1;
                if (h_abs < min_step || !(m == m)) {   // rk.py:129-130 TOO_SMALL_STEP; NaN state (model.py:69)
                    failed = true;
                    done = true;
                }
            }
            if (attempts >= MAX_ATTEMPTS) {
                failed = true;
                done = true;
            }
        }
        first = false;
        if (SHAPE) {
            if (__any_sync(FULL, sh_cur.n == SHAPE_LOG_CAP)) shape_flush<SHAPE>(gs, nsec, shape, sh_log, sh_cur);
        }
    }
    if (SHAPE) shape_flush<SHAPE>(gs, nsec, shape, sh_log, sh_cur);
    // zero-length spans cost one RHS call each and no step (base.py); a started section costs 2 calls (select_initial_step),
    // an attempt 6
    const int skipped = packed >> 8;
    info.nsec = skipped + ksec;
    info.nfev = skipped + 2 * ksec + 6 * attempts;
    info.nrej = nrej;
    info.nacc = attempts - nrej;
    info.failed = failed ? 1 : 0;
}

// INTEG 1: classical RK4, rk4_steps fixed steps per section (throughput mode; not the parity mode).
template <typename T, bool SLOW, bool UY0>
__device__ __forceinline__ void fk_rk4(const LaneId& id, bool valid, const GroupScratch& gs, const double* der,
                                       const double* alpha, int rk4_steps, T (&y)[6], FkInfo& info) {
    const int j = id.j;
    const int mn = (j == 2) ? 0 : j + 1, mp = (j == 0) ? 2 : j - 1;
    initial_state<T>(j, alpha, y);
    const int packed = valid ? gs.nsec : 0;
    const int nsec = packed & 0xff;
    info.nfev = info.nsec = packed >> 8;
    info.nacc = info.nrej = info.failed = 0;
    Section<T> sc;
    sc.Ao = sc.Bo = sc.An = sc.Bn = sc.Ap = sc.Bp = sc.kx = sc.ky = (T)0;
    sc.present = false;
    int ksec = 0, steps_left = 0;
    double t = 0.0, tf = 0.0;
    T h = (T)0;
    bool done = !valid;
    for (;;) {
        if (__any_sync(FULL, steps_left == 0 && !done)) {
            if (steps_left == 0 && !done) {
                if (advance_section<T>(gs, der, nsec, j, mn, mp, ksec, done, t, tf, sc)) {
                    h = (T)((tf - t) / (double)rk4_steps);
                    steps_left = rk4_steps;
                    info.nsec += 1;
                }
            }
        }
        if (!__any_sync(FULL, !done)) break;
        const bool active = !done;
        T Y[6], acc[6];
#pragma unroll
        for (int i = 0; i < 6; ++i) {
            Y[i] = y[i];
            acc[i] = (T)0;
        }
#pragma unroll 1
        for (int s = 0; s < 4; ++s) {
            T D[6];
            rhs<T, SLOW, UY0>(id, sc, true, Y, D);
            const T w = (s == 0 || s == 3) ? (T)(1.0 / 6.0) : (T)(1.0 / 3.0);
            const T c = (s == 2) ? h : h * (T)0.5;
#pragma unroll
            for (int i = 0; i < 6; ++i) {
                acc[i] = fma_t(w, D[i], acc[i]);
                Y[i] = fma_t(c, D[i], y[i]);
            }
        }
        if (active) {
#pragma unroll
            for (int i = 0; i < 6; ++i) y[i] = fma_t(h, acc[i], y[i]);
            info.nfev += 4;
            info.nacc += 1;
            steps_left -= 1;
        }
    }
}

// Per-env derived constants (domain randomisation / explicit tube parameters): C = L, Lc, EI, GJ, Ux, Uy (x3 each).
__device__ __forceinline__ void store_derived(const double (&C)[18], GroupScratch& gs) {
#pragma unroll
    for (int m = 0; m < 3; ++m) {
        const double ei = C[6 + m], gj = C[9 + m], ux = C[12 + m], uy = C[15 + m];
        const double k = (ei != 0.0) ? ei / gj : 0.0;            // model.py:101
        gs.der[DER_EI + m] = ei;
        gs.der[DER_EIUX + m] = ei * ux;
        gs.der[DER_EIUY + m] = ei * uy;
        gs.der[DER_KX + m] = k * ux;
        gs.der[DER_KY + m] = k * uy;
    }
}

// beta/alpha: float64 joint values (float32 joints promoted exactly, model.py:50-61).  L / Lc: tube lengths;
// der: the 15 derived constants of this environment's tubes (shared memory: per system, or gs.der per env).
template <typename T, int INTEG>
__device__ __forceinline__ void fk_group(const LaneId& id, bool valid, GroupScratch& gs, const double* L, const double* Lc,
                                         const double* der, const double* beta, const double* alpha, double rtol_d,
                                         double atol_d, int rk4_steps, T (&y)[6], FkInfo& info) {
    build_sections(L, Lc, beta, gs, id.j == 0);
    const bool slow = sizeof(T) == 8 && needs_slow_trig(id.j, alpha);   // warp-uniform
    constexpr bool UY0 = (INTEG & 2) != 0;     // INTEG: bit 0 = integrator (0 dp45, 1 rk4), bit 1 = all U_y are zero
    if ((INTEG & 1) == 0) {
        if (slow) fk_dp45<T, true, UY0>(id, valid, gs, der, alpha, rtol_d, atol_d, y, info);
        else fk_dp45<T, false, UY0>(id, valid, gs, der, alpha, rtol_d, atol_d, y, info);
    } else {
        if (slow) fk_rk4<T, true, UY0>(id, valid, gs, der, alpha, rk4_steps, y, info);
        else fk_rk4<T, false, UY0>(id, valid, gs, der, alpha, rk4_steps, y, info);
    }
    __syncwarp();   // the scratch slot may be rebuilt by the caller's next FK
}

// FK with backbone samples (float64 parity integrator only): same section table and integrator as fk_group.
template <bool UY0>
__device__ __forceinline__ void fk_group_shape(const LaneId& id, bool valid, GroupScratch& gs, const double* L,
                                               const double* Lc, const double* der, const double* beta, const double* alpha,
                                               double rtol_d, double atol_d, double (&y)[6], FkInfo& info,
                                               const ShapeOut shape) {
    build_sections(L, Lc, beta, gs, id.j == 0);
    if (needs_slow_trig(id.j, alpha)) fk_dp45<double, true, UY0, true>(id, valid, gs, der, alpha, rtol_d, atol_d, y, info, shape);
    else fk_dp45<double, false, UY0, true>(id, valid, gs, der, alpha, rtol_d, atol_d, y, info, shape);
    __syncwarp();
}

// ---- joint limits of one environment (obs.py:24-29,50-76) ------------------------------------------------------------
// The reference derives them from the tube lengths of the SYSTEM; with per-env tube lengths (ctr_set_tube_params, or
// domain_rand_length > 0 -- an extension, the reference forbids length randomisation, model_utils.py:12-14) they follow the
// env's own lengths with the same float32 roundings: low_i = f32(-L_i + 1e-3), d10 = f32(L_1 - L_0), d21 = f32(L_2 - L_1).
struct JointLimits {
    float low[6], high[6];   // Box bounds (beta: [-L + 1e-3, 0]; alpha: +-pi or +-inf)
    float d10, d21;          // ordering constants of obs.py:181
    float Lf[3];             // f32(L_i): goal-sampling predicate (obs.py:198-200)
};
__device__ __forceinline__ void limits_of_system(const SysConst& sc, JointLimits& lim) {
#pragma unroll
    for (int i = 0; i < 6; ++i) {
        lim.low[i] = sc.low[i];
        lim.high[i] = sc.high[i];
    }
    lim.d10 = sc.d10;
    lim.d21 = sc.d21;
#pragma unroll
    for (int i = 0; i < 3; ++i) lim.Lf[i] = sc.Lf[i];
}
// L: this env's tube lengths (float64).  For L equal to the system's lengths the result is bit-identical to limits_of_system.
__device__ __forceinline__ void limits_of_lengths(const SysConst& sc, const double* L, JointLimits& lim) {
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        lim.low[i] = (float)__dadd_rn(-L[i], 1e-3);
        lim.high[i] = 0.f;
        lim.low[3 + i] = sc.low[3 + i];
        lim.high[3 + i] = sc.high[3 + i];
        lim.Lf[i] = (float)L[i];
    }
    lim.d10 = (float)__dsub_rn(L[1], L[0]);
    lim.d21 = (float)__dsub_rn(L[2], L[1]);
}

// ---- Obs.set_action x n_substeps (obs.py:166-183, ctr_reach_env.py:134-135), float32 arithmetic -------------
__device__ __forceinline__ void set_action(float (&q)[6], const float (&a)[6], const JointLimits& lim, int n_substeps) {
    for (int s = 0; s < n_substeps; ++s) {
#pragma unroll
        for (int i = 0; i < 6; ++i) q[i] = fminf(fmaxf(__fadd_rn(q[i], a[i]), lim.low[i]), lim.high[i]);   // obs.py:172
        q[0] = fminf(q[0], q[1]);                                   // obs.py:180 (i = 1)
        q[0] = fmaxf(q[0], __fadd_rn(lim.d10, q[1]));               // obs.py:181
        q[1] = fminf(q[1], q[2]);                                   // i = 2
        q[1] = fmaxf(q[1], __fadd_rn(lim.d21, q[2]));
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
__device__ __forceinline__ int sample_goal(const JointLimits& sc, uint2 key, unsigned env_lo, unsigned env_hi,
                                           unsigned episode, unsigned stream, float (&q)[6], bool& ok) {
    int tries = 0;
    ok = false;
    const float PI_F = 3.14159274101257324f;   // float32(np.pi): the Box bounds (obs.py:60-63)
    while (tries < 1001) {                     // obs.py:204: counter > 1000 -> ValueError
        uint4 r = philox4x32(make_uint4(env_lo, env_hi, episode, (stream << 16) | (unsigned)tries), key);
        unsigned rr[3] = {r.x, r.y, r.z};
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            double lo = (double)sc.low[i];             // sample-space low of beta = the Box low (obs.py:59-63)
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
