/* CPU oracle (plain C) for the BVP kinematics back-end -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * A restatement of one shooting pass of the reference's full concentric-tube model with tip load
 * (/root/reference/ctr_reach_envs/envs/CTR_Python/CTR_Model.py: ode_eq :56-128, ode_solver :130-236) and of the
 * third-party integrator it calls (scipy.integrate.solve_ivp, method 'RK23', max_step = Tol; scipy 1.18.1:
 * scipy/integrate/_ivp/rk.py RK23 / rk_step / RungeKutta._step_impl, common.py select_initial_step).
 * Pinned against the reference class itself, run in the build container (oracle/make_golden_bvp.py ->
 * tests/golden/bvp_reference.npz, tests/test_oracle.py).
 *
 * One deliberate difference: the reference leaves d(u_iz)/ds and d(alpha_i)/ds of a tube that is ABSENT on a section
 * unassigned (`dydt = np.empty(...)`, CTR_Model.py:63,80-82), i.e. whatever the allocator hands back.  Those two states
 * of an absent tube feed nothing (every use is multiplied by EI_i = 0 or GJ_i = 0), but they do enter RK23's error norm.
 * Here they are 0.  The fixtures show what that costs: identical step counts on all but a few configurations, tips
 * within the tolerance stated in tests/test_oracle.py.
 */
#include <math.h>
#include <stdint.h>
#include <string.h>

#define NT 3
#define MAXSEC 9
#define NB 20 /* u1x u1y | u1z u2z u3z | a1 a2 a3 | r[3] | R[9] */

int ctr_oracle_segment(const double *tc, const double *beta, double *S, double *tab /* [4][3][MAXSEC] */);

typedef struct {
    double ei[NT], gj[NT], ux[NT], uy[NT], f[3];
} bsec_t;

/* CTR_Model.py:56-128 */
static void bvp_rhs(const bsec_t *c, const double *y, double *d) {
    double u[9];
    u[0] = y[0]; u[1] = y[1]; u[2] = y[2];
    for (int i = 1; i < NT; ++i) {                     /* u_i = Rz(alpha_i)^T u_1 (xy), own twist (:64-73) */
        double ct = cos(y[5 + i]), st = sin(y[5 + i]);
        u[3 * i + 0] = ct * y[0] + st * y[1];
        u[3 * i + 1] = -st * y[0] + ct * y[1];
        u[3 * i + 2] = y[2 + i];
    }
    for (int i = 0; i < NT; ++i) {                     /* :76-82; an absent tube's entries are unassigned there, 0 here */
        if (c->gj[i] != 0) {
            d[2 + i] = (c->ei[i] / c->gj[i]) * (u[3 * i] * c->uy[i] - u[3 * i + 1] * c->ux[i]);
            d[5 + i] = y[2 + i] - y[2];
        } else {
            d[2 + i] = 0.0;
            d[5 + i] = 0.0;
        }
    }
    double du0 = 0.0, du1 = 0.0, sei = 0.0;
    for (int i = 0; i < NT; ++i) {                     /* :84-112, term by term */
        double ca = cos(y[5 + i]), sa = sin(y[5 + i]);
        double ei = c->ei[i], gj = c->gj[i], uz = u[3 * i + 2], ux = u[3 * i], uyy = u[3 * i + 1];
        sei += ei;
        du0 += d[5 + i] * ei * y[1] + c->uy[i] * ei * uz * ca + c->ux[i] * ei * uz * sa - ei * uyy * uz * ca + gj * uyy * uz * ca -
               ei * ux * uz * sa + gj * ux * uz * sa;
        du1 += -d[5 + i] * ei * y[0] + c->uy[i] * ei * uz * sa - c->ux[i] * ei * uz * ca + ei * ux * uz * ca - gj * ux * uz * ca -
               ei * uyy * uz * sa + gj * uyy * uz * sa;
    }
    const double *R = y + 11;
    double g0 = R[0] * c->f[0] + R[3] * c->f[1] + R[6] * c->f[2];   /* R^T f */
    double g1 = R[1] * c->f[0] + R[4] * c->f[1] + R[7] * c->f[2];
    d[0] = -du0 / sei + g1 / sei;                      /* :114-118: Du = -K^-1 du - K^-1 [e3]x R^T f */
    d[1] = -du1 / sei - g0 / sei;
    d[8] = R[2]; d[9] = R[5]; d[10] = R[8];            /* r' = R e3 */
    for (int r = 0; r < 3; ++r) {                      /* R' = R u_hat(u_1) */
        const double a = R[3 * r], b = R[3 * r + 1], cc = R[3 * r + 2];
        d[11 + 3 * r + 0] = b * y[2] - cc * y[1];
        d[11 + 3 * r + 1] = cc * y[0] - a * y[2];
        d[11 + 3 * r + 2] = a * y[1] - b * y[0];
    }
}

static double rms(const double *x, const double *scale) {
    double s = 0.0;
    for (int i = 0; i < NB; ++i) s += (x[i] / scale[i]) * (x[i] / scale[i]);
    return sqrt(s / NB);
}

/* track u_z of tubes 2, 3 at the stored point nearest their tips (CTR_Model.py:231-235: argmin |Length - d_tip|, first hit) */
typedef struct { double best[2], uz[2]; } tipuz_t;
static void tip_note(tipuz_t *tz, const double *dtip, double t, const double *y) {
    for (int i = 1; i < NT; ++i) {
        double dist = fabs(t - dtip[i]);
        if (dist < tz->best[i - 1]) { tz->best[i - 1] = dist; tz->uz[i - 1] = y[2 + i]; }
    }
}

/* solve_ivp(fun, [t0, tf], y, method='RK23', max_step=max_step): y advanced in place; returns accepted steps or -1 */
static int rk23_section(const bsec_t *c, double t0, double tf, double *y, double rtol, double atol, double max_step,
                        tipuz_t *tz, const double *dtip, int *nfev) {
    static const double A21 = 0.5, A32 = 0.75, B1 = 2.0 / 9, B2 = 1.0 / 3, B3 = 4.0 / 9, E1 = 5.0 / 72, E2 = -1.0 / 12,
                        E3 = -1.0 / 9, E4 = 1.0 / 8;
    double f[NB], k2[NB], k3[NB], k4[NB], yn[NB], tmp[NB], scale[NB], t = t0;
    int steps = 0;
    bvp_rhs(c, y, f);
    *nfev += 1;
    tip_note(tz, dtip, t, y);                          /* s.t[0] */
    double h_abs;
    {                                                  /* common.py select_initial_step, order = 2 */
        double interval = fabs(tf - t0);
        if (interval == 0.0) return 0;
        for (int i = 0; i < NB; ++i) scale[i] = atol + fabs(y[i]) * rtol;
        double d0 = rms(y, scale), d1 = rms(f, scale);
        double h0 = (d0 < 1e-5 || d1 < 1e-5) ? 1e-6 : 0.01 * d0 / d1;
        if (h0 > interval) h0 = interval;
        for (int i = 0; i < NB; ++i) tmp[i] = y[i] + h0 * f[i];
        bvp_rhs(c, tmp, k2);
        *nfev += 1;
        for (int i = 0; i < NB; ++i) tmp[i] = k2[i] - f[i];
        double d2 = rms(tmp, scale) / h0;
        double h1 = (d1 <= 1e-15 && d2 <= 1e-15) ? fmax(1e-6, h0 * 1e-3) : pow(0.01 / fmax(d1, d2), 1.0 / 3.0);
        h_abs = fmin(fmin(100 * h0, h1), fmin(interval, max_step));
    }
    while (t != tf) {                                  /* rk.py _step_impl */
        double min_step = 10 * fabs(nextafter(t, INFINITY) - t);
        if (h_abs > max_step) h_abs = max_step;
        else if (h_abs < min_step) h_abs = min_step;
        int rejected = 0;
        for (;;) {
            if (h_abs < min_step) return -1;
            double t_new = t + h_abs;
            if (t_new - tf > 0) t_new = tf;
            double h = t_new - t;
            h_abs = fabs(h);
            for (int i = 0; i < NB; ++i) tmp[i] = y[i] + (f[i] * A21) * h;
            bvp_rhs(c, tmp, k2);
            for (int i = 0; i < NB; ++i) tmp[i] = y[i] + (f[i] * 0.0 + k2[i] * A32) * h;
            bvp_rhs(c, tmp, k3);
            for (int i = 0; i < NB; ++i) yn[i] = y[i] + h * (f[i] * B1 + k2[i] * B2 + k3[i] * B3);
            bvp_rhs(c, yn, k4);
            *nfev += 3;
            double s = 0.0;
            for (int i = 0; i < NB; ++i) {
                double sc = atol + fmax(fabs(y[i]), fabs(yn[i])) * rtol;
                double e = (f[i] * E1 + k2[i] * E2 + k3[i] * E3 + k4[i] * E4) * h / sc;
                s += e * e;
            }
            double err = sqrt(s / NB);
            if (err < 1) {
                double factor = (err == 0) ? 10.0 : fmin(10.0, 0.9 * pow(err, -1.0 / 3.0));
                if (rejected) factor = fmin(1.0, factor);
                h_abs *= factor;
                t = t_new;
                memcpy(y, yn, sizeof yn);
                memcpy(f, k4, sizeof k4);
                steps += 1;
                tip_note(tz, dtip, t, y);
                break;
            }
            h_abs *= fmax(0.2, 0.9 * pow(err, -1.0 / 3.0));
            rejected = 1;
        }
        if (steps > 100000) return -1;
    }
    return steps;
}

/* CTR_Model.ode_solver(u_init) with method = 1 (CTR_Model.py:130-236).
 * tc[18]: L, Lc, EI, GJ, Ux, Uy of the three tubes; q[6] = joints + q_0 (beta[3], alpha[3]); f[3] tip force;
 * u_init[5] = u1x, u1y, u1z, u2z, u3z at s = 0.
 * out[38] = tip r[3] | R[9] | cost[5] | final y[20] | accepted steps (all sections).  Returns 0, or -1 on failure. */
int ctr_oracle_bvp_pass(const double *tc, const double *q, const double *f, const double *u_init, double max_step,
                        double rtol, double atol, double *out) {
    double S[MAXSEC], tab[4 * NT * MAXSEC];
    const int m = ctr_oracle_segment(tc, q, S, tab);
    const double(*EI)[MAXSEC] = (const double(*)[MAXSEC])tab;
    const double(*GJ)[MAXSEC] = (const double(*)[MAXSEC])(tab + NT * MAXSEC);
    const double(*Ux)[MAXSEC] = (const double(*)[MAXSEC])(tab + 2 * NT * MAXSEC);
    const double(*Uy)[MAXSEC] = (const double(*)[MAXSEC])(tab + 3 * NT * MAXSEC);
    double y[NB];
    memset(y, 0, sizeof y);
    y[0] = u_init[0]; y[1] = u_init[1];
    y[2] = u_init[2]; y[3] = u_init[3]; y[4] = u_init[4];
    const double a1 = q[3];
    y[5] = q[3] - a1; y[6] = q[4] - a1; y[7] = q[5] - a1;               /* :31: twist relative to tube 1 */
    y[11] = cos(a1); y[12] = -sin(a1); y[14] = sin(a1); y[15] = cos(a1); y[19] = 1.0;
    double dtip[NT];
    for (int i = 0; i < NT; ++i) dtip[i] = tc[i] + q[i];
    tipuz_t tz = {{INFINITY, INFINITY}, {0.0, 0.0}};
    int total = 0, nfev = 0;
    double prev = 0.0;
    bsec_t c;
    memcpy(c.f, f, sizeof c.f);
    for (int seg = 0; seg < m; ++seg) {
        for (int i = 0; i < NT; ++i) { c.ei[i] = EI[i][seg]; c.gj[i] = GJ[i][seg]; c.ux[i] = Ux[i][seg]; c.uy[i] = Uy[i][seg]; }
        int st = rk23_section(&c, prev, S[seg], y, rtol, atol, max_step, &tz, dtip, &nfev);
        if (st < 0) return -1;
        total += st;
        prev = S[seg];
        if (seg < m - 1) {                              /* :186-226: moment continuity -> u1_xy of the next section */
            double sx = 0.0, sy = 0.0, sein = 0.0;
            for (int i = 0; i < NT; ++i) {
                const double th = (i == 0) ? 0.0 : y[5 + i];
                const double ct = cos(th), st_ = sin(th);
                const double uix = ct * y[0] + st_ * y[1], uiy = -st_ * y[0] + ct * y[1];     /* Rz(th)^T u1 */
                const double vx = EI[i][seg] * (uix - Ux[i][seg]) + EI[i][seg + 1] * Ux[i][seg + 1];
                const double vy = EI[i][seg] * (uiy - Uy[i][seg]) + EI[i][seg + 1] * Uy[i][seg + 1];
                sx += ct * vx - st_ * vy;                                                       /* Rz(th) v */
                sy += st_ * vx + ct * vy;
                sein += EI[i][seg + 1];
            }
            y[0] = sx / sein;
            y[1] = sy / sein;
        }
    }
    memcpy(out, y + 8, 3 * sizeof(double));
    memcpy(out + 3, y + 11, 9 * sizeof(double));
    out[12] = y[0] - Ux[0][m - 1];                      /* :228-235 */
    out[13] = y[1] - Uy[0][m - 1];
    out[14] = y[2];
    out[15] = tz.uz[0];
    out[16] = tz.uz[1];
    memcpy(out + 17, y, sizeof y);
    out[37] = (double)total;
    return 0;
}

/* jac (:251-262, eps 1e-4 on q) and comp (:264-275, 0.05 N on f) by forward differences over passes with fixed u_init.
 * J[3][6], C[3][3] row-major; tip[3] of the unperturbed pass. */
int ctr_oracle_bvp_jac_comp(const double *tc, const double *q, const double *f, const double *u_init, double max_step,
                            double rtol, double atol, double *J, double *C, double *tip) {
    double o[38], p[38], qq[6], ff[3];
    if (ctr_oracle_bvp_pass(tc, q, f, u_init, max_step, rtol, atol, o)) return -1;
    memcpy(tip, o, 3 * sizeof(double));
    memcpy(qq, q, sizeof qq);
    for (int i = 0; i < 6; ++i) {
        qq[i] = qq[i] + 1e-4;
        if (ctr_oracle_bvp_pass(tc, qq, f, u_init, max_step, rtol, atol, p)) return -1;
        for (int r = 0; r < 3; ++r) J[r * 6 + i] = (p[r] - o[r]) / 1e-4;
        qq[i] = qq[i] - 1e-4;                           /* the reference restores by subtraction (:261): q drifts by an ulp */
    }
    memcpy(ff, f, sizeof ff);
    if (C) {
        for (int i = 0; i < 3; ++i) {
            ff[i] = ff[i] + 0.05;
            if (ctr_oracle_bvp_pass(tc, q, ff, u_init, max_step, rtol, atol, p)) return -1;
            for (int r = 0; r < 3; ++r) C[r * 3 + i] = (p[r] - o[r]) / 0.05;
            ff[i] = ff[i] - 0.05;
        }
    }
    return 0;
}
