/* CPU oracle (plain C) for the CTR-Reach env-step path -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * A restatement of the reference algorithm (keshaviyengar/gym-ctr-reach, all citations relative
 * to /root/reference/ctr_reach_envs/) plus the third-party integrator it calls
 * (scipy.integrate.solve_ivp, method RK45, scipy 1.18.1: scipy/integrate/_ivp/{rk,common,ivp}.py).
 * It is the fast twin of oracle/ctr_oracle.py and is validated against it and against the
 * reference-generated fixtures in tests/golden/ (parity PINNED, see tests/test_oracle.py).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg may load
 * the library built from this file.  Build: `make -C oracle` (gcc -O2 -ffp-contract=off -pthread).
 */
#include <math.h>
#include <stdint.h>
#include <string.h>
#include <pthread.h>

#define NT 3
#define NY 18
#define MAXSEC 9

/* ---- Tube.py:14-15, Segment.py:57-59: derived per-tube constants --------------------------------
 * sys[3][8] = length, length_curved, d_inner, d_outer, E, G, Ux, Uy
 * tc[18]    = L[3], Lc[3], EI[3], GJ[3], Ux[3], Uy[3] */
void ctr_oracle_derive(const double *sys, double *tc) {
    for (int i = 0; i < NT; ++i) {
        const double *p = sys + 8 * i;
        double di = p[2], d_o = p[3];
        double J = (M_PI * (pow(d_o, 4) - pow(di, 4))) / 32;
        double I = (M_PI * (pow(d_o, 4) - pow(di, 4))) / 64;
        tc[0 + i] = p[0];
        tc[3 + i] = p[1];
        tc[6 + i] = p[4] * I;
        tc[9 + i] = p[5] * J;
        tc[12 + i] = p[6];
        tc[15 + i] = p[7];
    }
}

/* ---- Segment.py:6-61 (value-based restatement, see ctr_oracle.py: segment) ----------------------
 * returns m (number of kept sections); S[m], and EI/GJ/Ux/Uy as [3][MAXSEC] */
static int segment(const double *tc, const double *beta, double *S, double EI[NT][MAXSEC], double GJ[NT][MAXSEC],
                   double Ux[NT][MAXSEC], double Uy[NT][MAXSEC]) {
    double dtip[NT], dc[NT], pts[10];
    pts[0] = 0.0;
    for (int i = 0; i < NT; ++i) {
        dtip[i] = tc[i] + beta[i];
        dc[i] = dtip[i] - tc[3 + i];
        pts[1 + i] = beta[i];
        pts[4 + i] = dc[i];
        pts[7 + i] = dtip[i];
    }
    for (int i = 1; i < 10; ++i) { /* insertion sort */
        double v = pts[i];
        int j = i - 1;
        while (j >= 0 && pts[j] > v) { pts[j + 1] = pts[j]; --j; }
        pts[j + 1] = v;
    }
    double bmin = beta[0];
    for (int i = 1; i < NT; ++i) if (beta[i] < bmin) bmin = beta[i];
    double cum = 0.0;
    int m = 0;
    for (int k = 0; k < 9; ++k) {
        double len = 1e-5 * floor(1e5 * (pts[k + 1] - pts[k])); /* Segment.py:16 */
        if (len == 0) continue;                                 /* Segment.py:46-50 */
        cum = cum + len;                                        /* Segment.py:52 */
        if (!(cum + bmin > 0)) continue;                        /* Segment.py:53 */
        S[m] = cum + bmin;
        for (int i = 0; i < NT; ++i) {
            int present = (beta[i] <= pts[k]) && (pts[k + 1] <= dtip[i]);
            int curved = present && (dc[i] <= pts[k]);
            EI[i][m] = present ? tc[6 + i] : 0.0;
            GJ[i][m] = present ? tc[9 + i] : 0.0;
            Ux[i][m] = curved ? tc[12 + i] : 0.0;
            Uy[i][m] = curved ? tc[15 + i] : 0.0;
        }
        ++m;
    }
    return m;
}

int ctr_oracle_segment(const double *tc, const double *beta, double *S, double *tab /* [4][3][MAXSEC] */) {
    double EI[NT][MAXSEC], GJ[NT][MAXSEC], Ux[NT][MAXSEC], Uy[NT][MAXSEC];
    int m = segment(tc, beta, S, EI, GJ, Ux, Uy);
    memcpy(tab, EI, sizeof EI);
    memcpy(tab + NT * MAXSEC, GJ, sizeof GJ);
    memcpy(tab + 2 * NT * MAXSEC, Ux, sizeof Ux);
    memcpy(tab + 3 * NT * MAXSEC, Uy, sizeof Uy);
    return m;
}

/* ---- model.py:72-117: ODE right-hand side ------------------------------------------------------- */
typedef struct { double ei[NT], gj[NT], ux0[NT], uy0[NT]; } sec_t;

static void rhs(const sec_t *c, const double *y, double *dy) {
    const double *a = y + 3;
    double inv = 1 / (c->ei[0] + c->ei[1] + c->ei[2]);
    double ux[NT], uy[NT];
    for (int i = 0; i < NT; ++i) {
        double sx = 0.0, sy = 0.0;
        for (int j = 0; j < NT; ++j) {
            double cs = cos(a[i] - a[j]), sn = sin(a[i] - a[j]);
            sx = sx + (c->ei[j] * c->ux0[j] * cs + c->ei[j] * c->uy0[j] * sn);   /* model.py:87-90 */
            sy = sy + (-c->ei[j] * c->ux0[j] * sn + c->ei[j] * c->uy0[j] * cs);  /* model.py:91-94 */
        }
        ux[i] = inv * sx;
        uy[i] = inv * sy;
    }
    for (int j = 0; j < NT; ++j) {
        if (c->ei[j] == 0) { dy[j] = 0.0; dy[3 + j] = 0.0; }                      /* model.py:97-99 */
        else {
            dy[j] = (c->ei[j] / c->gj[j]) * (ux[j] * c->uy0[j] - uy[j] * c->ux0[j]); /* model.py:101 */
            dy[3 + j] = y[j];                                                      /* model.py:102 */
        }
    }
    const double *R = y + 9;
    double u0 = ux[0], u1 = uy[0], u2 = y[0];
    dy[6] = R[2]; dy[7] = R[5]; dy[8] = R[8];                                      /* model.py:108 */
    for (int r = 0; r < 3; ++r) {                                                  /* model.py:107,109 */
        double r0 = R[3 * r], r1 = R[3 * r + 1], r2 = R[3 * r + 2];
        dy[9 + 3 * r] = r1 * u2 - r2 * u1;
        dy[10 + 3 * r] = -r0 * u2 + r2 * u0;
        dy[11 + 3 * r] = r0 * u1 - r1 * u0;
    }
}

/* ---- scipy RK45 (rk.py: RK45 tableau, rk_step, RungeKutta._step_impl; common.py: norm,
 *      select_initial_step; ivp.py: last t_eval sample = dense output at x = 1) ------------------- */
static const double A21 = 1.0 / 5;
static const double A3[2] = {3.0 / 40, 9.0 / 40};
static const double A4[3] = {44.0 / 45, -56.0 / 15, 32.0 / 9};
static const double A5[4] = {19372.0 / 6561, -25360.0 / 2187, 64448.0 / 6561, -212.0 / 729};
static const double A6[5] = {9017.0 / 3168, -355.0 / 33, 46732.0 / 5247, 49.0 / 176, -5103.0 / 18656};
static const double Bc[6] = {35.0 / 384, 0.0, 500.0 / 1113, 125.0 / 192, -2187.0 / 6784, 11.0 / 84};
static const double Ec[7] = {-71.0 / 57600, 0.0, 71.0 / 16695, -71.0 / 1920, 17253.0 / 339200, -22.0 / 525, 1.0 / 40};
static const double Pc[7][4] = {
    {1, -8048581381.0 / 2820520608.0, 8663915743.0 / 2820520608.0, -12715105075.0 / 11282082432.0},
    {0, 0, 0, 0},
    {0, 131558114200.0 / 32700410799.0, -68118460800.0 / 10900136933.0, 87487479700.0 / 32700410799.0},
    {0, -1754552775.0 / 470086768.0, 14199869525.0 / 1410260304.0, -10690763975.0 / 1880347072.0},
    {0, 127303824393.0 / 49829197408.0, -318862633887.0 / 49829197408.0, 701980252875.0 / 199316789632.0},
    {0, -282668133.0 / 205662961.0, 2019193451.0 / 616988883.0, -1453857185.0 / 822651844.0},
    {0, 40617522.0 / 29380423.0, -110615467.0 / 29380423.0, 69997945.0 / 29380423.0}};

static double rms_scaled(const double *x, const double *scale) {
    double s = 0.0;
    for (int i = 0; i < NY; ++i) { double v = x[i] / scale[i]; s += v * v; }
    return sqrt(s) / sqrt((double)NY);
}

typedef struct { int nfev, nacc, nrej; } cnt_t;

/* backbone samples of one section (model.py:137-151: t_eval = np.linspace(span[k], span[k+1] - 1e-6, 30), values from scipy's
 * RK45 dense output, rk.py RkDenseOutput._call_impl): te = the 30 abscissae, out = where the next sample's r goes */
#define NEVAL 30
typedef struct { double te[NEVAL]; int i; double *out; } shape_t;
static shape_t *g_shape = 0;   /* set by ctr_oracle_fk_shape around its own dp45_section calls (single-threaded entry) */

/* integrate one section from t0 to tf (tf >= t0); y updated in place; returns 0 ok, -1 step too small */
static int dp45_section(const sec_t *c, double t0, double tf, double *y, double rtol, double atol, cnt_t *cnt) {
    double K[7][NY], ytmp[NY], ynew[NY], scale[NY], tmp[NY];
    shape_t *sh = g_shape;
    rhs(c, y, K[0]);
    cnt->nfev += 1;
    if (tf == t0) {                                          /* ConstantDenseOutput: every sample is y0 (base.py) */
        if (sh) for (; sh->i < NEVAL; ++sh->i) for (int d = 0; d < 3; ++d) sh->out[3 * sh->i + d] = y[6 + d];
        return 0;
    }
    /* select_initial_step (common.py:68-134) */
    double interval = fabs(tf - t0);
    for (int i = 0; i < NY; ++i) scale[i] = atol + fabs(y[i]) * rtol;
    double d0 = rms_scaled(y, scale), d1 = rms_scaled(K[0], scale);
    double h0 = (d0 < 1e-5 || d1 < 1e-5) ? 1e-6 : 0.01 * d0 / d1;
    if (interval < h0) h0 = interval;
    for (int i = 0; i < NY; ++i) ytmp[i] = y[i] + h0 * K[0][i];
    rhs(c, ytmp, K[1]);
    cnt->nfev += 1;
    for (int i = 0; i < NY; ++i) tmp[i] = K[1][i] - K[0][i];
    double d2 = rms_scaled(tmp, scale) / h0;
    double h1;
    if (d1 <= 1e-15 && d2 <= 1e-15) h1 = fmax(1e-6, h0 * 1e-3);
    else h1 = pow(0.01 / fmax(d1, d2), 1.0 / 5);
    double h_abs = fmin(fmin(100 * h0, h1), interval);
    double t = t0;
    while (t != tf) {
        double min_step = 10 * fabs(nextafter(t, INFINITY) - t);
        if (h_abs < min_step) h_abs = min_step;
        int rejected = 0;
        double t_new, h;
        for (;;) {
            if (h_abs < min_step) return -1;
            t_new = t + h_abs;
            if (t_new - tf > 0) t_new = tf;
            h = t_new - t;
            h_abs = fabs(h);
            for (int i = 0; i < NY; ++i) ytmp[i] = y[i] + (K[0][i] * A21) * h;
            rhs(c, ytmp, K[1]);
            for (int i = 0; i < NY; ++i) ytmp[i] = y[i] + (K[0][i] * A3[0] + K[1][i] * A3[1]) * h;
            rhs(c, ytmp, K[2]);
            for (int i = 0; i < NY; ++i) ytmp[i] = y[i] + (K[0][i] * A4[0] + K[1][i] * A4[1] + K[2][i] * A4[2]) * h;
            rhs(c, ytmp, K[3]);
            for (int i = 0; i < NY; ++i)
                ytmp[i] = y[i] + (K[0][i] * A5[0] + K[1][i] * A5[1] + K[2][i] * A5[2] + K[3][i] * A5[3]) * h;
            rhs(c, ytmp, K[4]);
            for (int i = 0; i < NY; ++i)
                ytmp[i] = y[i] + (K[0][i] * A6[0] + K[1][i] * A6[1] + K[2][i] * A6[2] + K[3][i] * A6[3] + K[4][i] * A6[4]) * h;
            rhs(c, ytmp, K[5]);
            for (int i = 0; i < NY; ++i) {
                double s = 0.0;
                for (int j = 0; j < 6; ++j) s += K[j][i] * Bc[j];
                ynew[i] = y[i] + h * s;
            }
            rhs(c, ynew, K[6]);
            cnt->nfev += 6;
            for (int i = 0; i < NY; ++i) {
                double s = 0.0;
                for (int j = 0; j < 7; ++j) s += K[j][i] * Ec[j];
                tmp[i] = s * h;
                scale[i] = atol + fmax(fabs(y[i]), fabs(ynew[i])) * rtol;
            }
            double err = rms_scaled(tmp, scale);
            if (err < 1) {
                double factor = (err == 0) ? 10.0 : fmin(10.0, 0.9 * pow(err, -0.2));
                if (rejected) factor = fmin(1.0, factor);
                h_abs *= factor;
                cnt->nacc += 1;
                break;
            }
            h_abs *= fmax(0.2, 0.9 * pow(err, -0.2));
            rejected = 1;
            cnt->nrej += 1;
        }
        if (sh) {                                            /* ivp.py: t_eval points in (t_old, t_new] via searchsorted 'right' */
            while (sh->i < NEVAL && sh->te[sh->i] <= t_new) {
                const double x = (sh->te[sh->i] - t) / h;
                double pw[4];
                pw[0] = x; pw[1] = pw[0] * x; pw[2] = pw[1] * x; pw[3] = pw[2] * x;   /* np.cumprod */
                for (int d = 0; d < 3; ++d) {
                    double q = 0.0;
                    for (int p = 0; p < 4; ++p) {
                        double sq = 0.0;
                        for (int j = 0; j < 7; ++j) sq += K[j][6 + d] * Pc[j][p];     /* Q = K.T.dot(P) */
                        q += sq * pw[p];
                    }
                    sh->out[3 * sh->i + d] = h * q + y[6 + d];
                }
                sh->i += 1;
            }
        }
        if (t_new == tf) { /* stored sample at t_bound is the dense-output polynomial at x == 1 */
            for (int i = 0; i < NY; ++i) {
                double q = 0.0;
                for (int p = 0; p < 4; ++p) {
                    double s = 0.0;
                    for (int j = 0; j < 7; ++j) s += K[j][i] * Pc[j][p];
                    q += s;
                }
                y[i] = h * q + y[i];
            }
        } else {
            memcpy(y, ynew, sizeof ynew);
        }
        memcpy(K[0], K[6], sizeof K[0]);
        t = t_new;
    }
    return 0;
}

/* ---- model.py:30-70,119-174: forward kinematics --------------------------------------------------
 * q = [b0,b1,b2,a0,a1,a2] (float64 values; float32 joints are promoted exactly, model.py:50-61)
 * out[18] final ODE state (uz, alpha, r, R); info = {nfev, accepted, rejected, sections}; returns 0 / -1 */
int ctr_oracle_fk(const double *tc, const double *q, double rtol, double atol, double *out, int *info) {
    double S[MAXSEC], EI[NT][MAXSEC], GJ[NT][MAXSEC], Ux[NT][MAXSEC], Uy[NT][MAXSEC];
    int m = segment(tc, q, S, EI, GJ, Ux, Uy);
    double y[NY];
    memset(y, 0, sizeof y);
    y[3] = q[3]; y[4] = q[4]; y[5] = q[5];
    double ca = cos(q[3]), sa = sin(q[3]);                   /* model.py:57-60 */
    y[9] = ca; y[10] = -sa; y[12] = sa; y[13] = ca; y[17] = 1.0;
    cnt_t cnt = {0, 0, 0};
    int status = 0;
    double prev = 0.0;
    for (int k = 0; k < m; ++k) {
        sec_t c;
        for (int i = 0; i < NT; ++i) { c.ei[i] = EI[i][k]; c.gj[i] = GJ[i][k]; c.ux0[i] = Ux[i][k]; c.uy0[i] = Uy[i][k]; }
        double t0 = prev, tf = S[k] - 1e-6;                  /* model.py:141 */
        if (tf < t0) { double z = t0; t0 = tf; tf = z; }     /* model.py:145-151 */
        if (dp45_section(&c, t0, tf, y, rtol, atol, &cnt)) status = -1;
        prev = S[k];
    }
    memcpy(out, y, sizeof y);
    if (info) { info[0] = cnt.nfev; info[1] = cnt.nacc; info[2] = cnt.nrej; info[3] = m; }
    return status;
}

/* backbone shape (model.py:66-68,137-172): r[npts][3] with npts = 30 x sections, tip_idx[3] = tip_pos (model.py:166-172).
 * Not thread-safe (g_shape). Returns the number of points, or -1. */
int ctr_oracle_fk_shape(const double *tc, const double *q, double rtol, double atol, double *r, int *tip_idx) {
    double S[MAXSEC], EI[NT][MAXSEC], GJ[NT][MAXSEC], Ux[NT][MAXSEC], Uy[NT][MAXSEC];
    int m = segment(tc, q, S, EI, GJ, Ux, Uy);
    double y[NY], len[MAXSEC * NEVAL];
    memset(y, 0, sizeof y);
    y[3] = q[3]; y[4] = q[4]; y[5] = q[5];
    double ca = cos(q[3]), sa = sin(q[3]);
    y[9] = ca; y[10] = -sa; y[12] = sa; y[13] = ca; y[17] = 1.0;
    cnt_t cnt = {0, 0, 0};
    int status = 0;
    double prev = 0.0;
    shape_t sh;
    for (int k = 0; k < m; ++k) {
        sec_t c;
        for (int i = 0; i < NT; ++i) { c.ei[i] = EI[i][k]; c.gj[i] = GJ[i][k]; c.ux0[i] = Ux[i][k]; c.uy0[i] = Uy[i][k]; }
        const double a = prev, b = S[k] - 1e-6;              /* np.linspace(a, b, 30): i * step + a, last = b */
        const double step = (b - a) / (NEVAL - 1);
        for (int i = 0; i < NEVAL; ++i) sh.te[i] = (step == 0.0) ? ((double)i / (NEVAL - 1)) * (b - a) + a : (double)i * step + a;
        sh.te[NEVAL - 1] = b;
        if (b < a) {                                         /* model.py:145-148: np.sort(s_span) */
            for (int i = 0; i < NEVAL / 2; ++i) { double z = sh.te[i]; sh.te[i] = sh.te[NEVAL - 1 - i]; sh.te[NEVAL - 1 - i] = z; }
        }
        for (int i = 0; i < NEVAL; ++i) len[k * NEVAL + i] = sh.te[i];
        sh.i = 0;
        sh.out = r + 3 * NEVAL * k;
        double t0 = a, tf = b;
        if (tf < t0) { double z = t0; t0 = tf; tf = z; }
        g_shape = &sh;
        if (dp45_section(&c, t0, tf, y, rtol, atol, &cnt)) status = -1;
        g_shape = 0;
        if (sh.i != NEVAL) status = -1;
        prev = S[k];
    }
    for (int k = 0; k < NT; ++k) {                           /* model.py:166-172 */
        const double d_tip = tc[k] + q[k];
        int b = 0;
        for (int i = 0; i < m * NEVAL; ++i) if (len[i] >= d_tip - 1e-3) { b = i; break; }
        tip_idx[k] = b;
    }
    return status ? -1 : m * NEVAL;
}

/* per-section step attempts of one FK (analysis of the GPU kernel's lock-step cost, oracle/lockstep_sim.py):
 * att[MAXSEC] = accepted + rejected steps of each kept section, returns the section count */
int ctr_oracle_fk_attempts(const double *tc, const double *q, double rtol, double atol, int *att) {
    double S[MAXSEC], EI[NT][MAXSEC], GJ[NT][MAXSEC], Ux[NT][MAXSEC], Uy[NT][MAXSEC];
    int m = segment(tc, q, S, EI, GJ, Ux, Uy);
    double y[NY];
    memset(y, 0, sizeof y);
    y[3] = q[3]; y[4] = q[4]; y[5] = q[5];
    double ca = cos(q[3]), sa = sin(q[3]);
    y[9] = ca; y[10] = -sa; y[12] = sa; y[13] = ca; y[17] = 1.0;
    double prev = 0.0;
    for (int k = 0; k < MAXSEC; ++k) att[k] = 0;
    for (int k = 0; k < m; ++k) {
        sec_t c;
        cnt_t cnt = {0, 0, 0};
        for (int i = 0; i < NT; ++i) { c.ei[i] = EI[i][k]; c.gj[i] = GJ[i][k]; c.ux0[i] = Ux[i][k]; c.uy0[i] = Uy[i][k]; }
        double t0 = prev, tf = S[k] - 1e-6;
        if (tf < t0) { double z = t0; t0 = tf; tf = z; }
        dp45_section(&c, t0, tf, y, rtol, atol, &cnt);
        att[k] = cnt.nacc + cnt.nrej;
        prev = S[k];
    }
    return m;
}
typedef struct { const double *tc; const double *q; int32_t *att; } att_ctx;
static void att_range(void *p, int64_t lo, int64_t hi, int *bad);
int ctr_oracle_fk_attempts_batch(const double *tc, const double *q, int64_t n, int32_t *att, int nthreads);

/* ---- minimal pthread parallel-for (libgomp is not in this image) ------------------------------- */
typedef void (*range_fn)(void *ctx, int64_t lo, int64_t hi, int *bad);
typedef struct { range_fn fn; void *ctx; int64_t lo, hi; int bad; } job_t;
static void *job_main(void *p) { job_t *j = (job_t *)p; j->fn(j->ctx, j->lo, j->hi, &j->bad); return 0; }
static int parallel_for(range_fn fn, void *ctx, int64_t n, int nthreads) {
    if (nthreads < 1) nthreads = 1;
    if (nthreads > 256) nthreads = 256;
    if ((int64_t)nthreads > n) nthreads = n > 0 ? (int)n : 1;
    job_t jobs[256];
    pthread_t th[256];
    int64_t chunk = (n + nthreads - 1) / nthreads;
    int bad = 0;
    for (int k = 0; k < nthreads; ++k) {
        jobs[k].fn = fn; jobs[k].ctx = ctx; jobs[k].bad = 0;
        jobs[k].lo = k * chunk; jobs[k].hi = (k + 1) * chunk < n ? (k + 1) * chunk : n;
        if (k > 0) pthread_create(&th[k], 0, job_main, &jobs[k]);
    }
    job_main(&jobs[0]);
    for (int k = 1; k < nthreads; ++k) pthread_join(th[k], 0);
    for (int k = 0; k < nthreads; ++k) bad += jobs[k].bad;
    return bad;
}

/* batch FK: tcs[nsys][18], system[n] (or NULL -> 0), q[n][6] -> state[n][18], info[n][4]; pthreads over n */
typedef struct { const double *tcs; const int32_t *system; const double *q; double rtol, atol; double *state; int32_t *info; } fk_ctx;
static void fk_range(void *p, int64_t lo, int64_t hi, int *bad) {
    fk_ctx *c = (fk_ctx *)p;
    for (int64_t i = lo; i < hi; ++i) {
        int inf[4];
        int s = c->system ? c->system[i] : 0;
        if (ctr_oracle_fk(c->tcs + 18 * s, c->q + 6 * i, c->rtol, c->atol, c->state + NY * i, inf)) *bad += 1;
        if (c->info) for (int k = 0; k < 4; ++k) c->info[4 * i + k] = inf[k];
    }
}
int ctr_oracle_fk_batch(const double *tcs, const int32_t *system, const double *q, int64_t n, double rtol, double atol,
                        double *state, int32_t *info, int nthreads) {
    fk_ctx c = {tcs, system, q, rtol, atol, state, info};
    return parallel_for(fk_range, &c, n, nthreads);
}

static void att_range(void *p, int64_t lo, int64_t hi, int *bad) {
    att_ctx *c = (att_ctx *)p;
    (void)bad;
    for (int64_t i = lo; i < hi; ++i) {
        int a[MAXSEC];
        ctr_oracle_fk_attempts(c->tc, c->q + 6 * i, 1e-3, 1e-6, a);
        for (int k = 0; k < MAXSEC; ++k) c->att[MAXSEC * i + k] = a[k];
    }
}
int ctr_oracle_fk_attempts_batch(const double *tc, const double *q, int64_t n, int32_t *att, int nthreads) {
    att_ctx c = {tc, q, att};
    return parallel_for(att_range, &c, n, nthreads);
}

/* ---- obs.py:166-183 x n_substeps (ctr_reach_env.py:134-135), float32 arithmetic ------------------
 * lim[12] = low[6], high[6] (float32, obs.py:59-75); d10 = f32(L1 - L0), d21 = f32(L2 - L1) */
void ctr_oracle_set_action(float *j, const float *a, const float *lim, float d10, float d21, int n_substeps) {
    for (int s = 0; s < n_substeps; ++s) {
        for (int k = 0; k < 6; ++k) {
            volatile float v = j[k] + a[k];
            float w = v;
            if (w < lim[k]) w = lim[k];
            if (w > lim[6 + k]) w = lim[6 + k];
            j[k] = w;
        }
        volatile float c0, c1;
        if (j[1] < j[0]) j[0] = j[1];
        c0 = d10 + j[1];
        if (c0 > j[0]) j[0] = c0;
        if (j[2] < j[1]) j[1] = j[2];
        c1 = d21 + j[2];
        if (c1 > j[1]) j[1] = c1;
    }
}

void ctr_oracle_set_action_batch(float *j, const float *a, const int32_t *system, const float *lims /*[nsys][12]*/,
                                 const float *dl /*[nsys][2]*/, int n_substeps, int64_t n) {
    for (int64_t i = 0; i < n; ++i) {
        int s = system ? system[i] : 0;
        ctr_oracle_set_action(j + 6 * i, a + 6 * i, lims + 12 * s, dl[2 * s], dl[2 * s + 1], n_substeps);
    }
}

/* ---- a batched port of one env step (constraints -> FK -> flags), used as a timed CPU baseline ----
 * joints[n][6] f32 in/out, action[n][6] f32, desired[n][3]; outputs achieved[n][3], reward[n], done[n], t[n] */
typedef struct {
    const double *tc; const float *lim; float d10, d21; int n_substeps, max_steps; double tol; float *joints;
    const float *action; const double *desired; int32_t *t; double *achieved, *reward; uint8_t *done;
} step_ctx;
static void step_range(void *p, int64_t lo, int64_t hi, int *bad) {
    step_ctx *c = (step_ctx *)p;
    for (int64_t i = lo; i < hi; ++i) {
        double q[6], st[NY];
        ctr_oracle_set_action(c->joints + 6 * i, c->action + 6 * i, c->lim, c->d10, c->d21, c->n_substeps);
        for (int k = 0; k < 6; ++k) q[k] = (double)c->joints[6 * i + k];
        if (ctr_oracle_fk(c->tc, q, 1e-3, 1e-6, st, 0)) *bad += 1;
        double dx = st[6] - c->desired[3 * i], dy = st[7] - c->desired[3 * i + 1], dz = st[8] - c->desired[3 * i + 2];
        double d = sqrt(dx * dx + dy * dy + dz * dz);
        c->achieved[3 * i] = st[6]; c->achieved[3 * i + 1] = st[7]; c->achieved[3 * i + 2] = st[8];
        c->t[i] += 1;
        c->reward[i] = (d > c->tol) ? -1.0 : -0.0;
        c->done[i] = (uint8_t)(!(d > c->tol) || c->t[i] >= c->max_steps);
    }
}
int ctr_oracle_step_batch(const double *tc, const float *lim, float d10, float d21, int n_substeps, int max_steps,
                          double tol, float *joints, const float *action, const double *desired, int32_t *t,
                          double *achieved, double *reward, uint8_t *done, int64_t n, int nthreads) {
    step_ctx c = {tc, lim, d10, d21, n_substeps, max_steps, tol, joints, action, desired, t, achieved, reward, done};
    return parallel_for(step_range, &c, n, nthreads);
}
