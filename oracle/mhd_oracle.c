/*
 * oracle/mhd_oracle.c -- CPU restatement of the reference's 2-D MUSCL-Hancock +
 * HLLC + mixed-GLM ideal-MHD step (MHD_codes/MUSCL_NEW.cpp).
 *
 * TEST INFRASTRUCTURE ONLY.  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may link or call this file.
 * The product path (lsc_internship_b200/csrc) never does.
 *
 * Parity pin: this restatement is checked bit-for-bit against the REAL
 * reference compiled from /root/reference by oracle/build_ref.py
 * (oracle/_ref/muscl_ref), which itself reproduces the reference's committed
 * golden files MHD_plot_data/MHD*.dat byte-for-byte
 * (tests/test_oracle_vs_ref.py, tests/golden/).
 *
 * All arithmetic is IEEE binary64 with the reference's operation order;
 * compile with -O2 -ffp-contract=off and no -march (no FMA).
 *
 * Layout: AoS, u[(i*(ny+4) + j)*9 + k], i = x (outer), j = y (inner), two
 * ghost cells per side -- exactly the reference's
 * std::vector<std::vector<std::array<double,9>>> (MUSCL_NEW.cpp:673-674).
 * Component order (MUSCL_NEW.cpp:11-19): rho, mx, my, mz, E, Bx, By, Bz, psi.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

#define NV 9

typedef struct {
    int nx, ny;
    double x0, x1, y0, y1;
    double gamma, cfl;
    int bcx; /* 0 periodic (MUSCL_NEW.cpp:609-614), 1 transmissive (:617-622) */
    int bcy; /* 0 reflective (:625-638), 1 periodic (:641-646), 2 transmissive (:649-654) */
} lsc_oracle_mhd_cfg;

typedef struct {
    lsc_oracle_mhd_cfg c;
    size_t ncell;
    double dx, dy, t, dt, ch;
    long step;
    /* the reference's work arrays, MUSCL_NEW.cpp:673-713 */
    double *u, *v, *up1, *fx, *fy, *ubar;
    double *rxL, *rxR, *ryB, *ryT, *hxL, *hxR, *hyB, *hyT;
} mhd_t;

/* std::min / std::max semantics (return first argument on ties / unordered) */
static double smin(double a, double b) { return (b < a) ? b : a; }
static double smax(double a, double b) { return (a < b) ? b : a; }

/* MUSCL_NEW.cpp:9-21 */
void lsc_oracle_mhd_p2c(const double *p, double g, double *c) {
    c[0] = p[0];
    c[1] = p[0] * p[1];
    c[2] = p[0] * p[2];
    c[3] = p[0] * p[3];
    c[5] = p[5];
    c[6] = p[6];
    c[7] = p[7];
    c[4] = p[4] / (g - 1) + 0.5 * p[0] * (p[1] * p[1] + p[2] * p[2] + p[3] * p[3]) +
           0.5 * (p[5] * p[5] + p[6] * p[6] + p[7] * p[7]);
    c[8] = p[8];
}

/* MUSCL_NEW.cpp:22-34 */
void lsc_oracle_mhd_c2p(const double *u, double g, double *v) {
    v[0] = u[0];
    v[1] = u[1] / u[0];
    v[2] = u[2] / u[0];
    v[3] = u[3] / u[0];
    v[5] = u[5];
    v[6] = u[6];
    v[7] = u[7];
    v[4] = (u[4] - 0.5 * v[0] * (v[1] * v[1] + v[2] * v[2] + v[3] * v[3]) -
            0.5 * (u[5] * u[5] + u[6] * u[6] + u[7] * u[7])) * (g - 1);
    v[8] = u[8];
}

/* MUSCL_NEW.cpp:36-49 */
void lsc_oracle_mhd_fluxdef_x(const double *u, double g, double ch, double *f) {
    double v[NV];
    lsc_oracle_mhd_c2p(u, g, v);
    f[0] = u[1];
    f[1] = u[0] * v[1] * v[1] + v[4] + 0.5 * (u[5] * u[5] + u[6] * u[6] + u[7] * u[7]) - u[5] * u[5];
    f[2] = u[0] * v[1] * v[2] - u[5] * u[6];
    f[3] = u[0] * v[1] * v[3] - u[5] * u[7];
    f[4] = (u[4] + v[4] + 0.5 * (u[5] * u[5] + u[6] * u[6] + u[7] * u[7])) * v[1] -
           (v[1] * v[5] + v[2] * v[6] + v[3] * v[7]) * v[5];
    f[5] = u[8];
    f[6] = v[6] * v[1] - v[5] * v[2];
    f[7] = v[7] * v[1] - v[5] * v[3];
    f[8] = ch * ch * u[5];
}

/* MUSCL_NEW.cpp:51-64 */
void lsc_oracle_mhd_fluxdef_y(const double *u, double g, double ch, double *f) {
    double v[NV];
    lsc_oracle_mhd_c2p(u, g, v);
    f[0] = u[2];
    f[1] = u[0] * v[1] * v[2] - v[5] * v[6];
    f[2] = u[0] * v[2] * v[2] + v[4] + 0.5 * (u[5] * u[5] + u[6] * u[6] + u[7] * u[7]) - u[6] * u[6];
    f[3] = u[0] * v[2] * v[3] - u[6] * u[7];
    f[4] = (u[4] + v[4] + 0.5 * (u[5] * u[5] + u[6] * u[6] + u[7] * u[7])) * v[2] -
           (v[1] * v[5] + v[2] * v[6] + v[3] * v[7]) * v[6];
    f[5] = v[5] * v[2] - v[6] * v[1];
    f[6] = u[8];
    f[7] = v[7] * v[2] - v[6] * v[3];
    f[8] = ch * ch * u[6];
}

/* mixed-GLM interface values, MUSCL_NEW.cpp:121-138; n = 5 (x) or 6 (y) */
static void glm_interface(const double *L, const double *R, double ch, int n, double *BnT, double *psiT) {
    double psiL = L[8], psiR = R[8], BL = L[n], BR = R[n];
    *BnT = 0.5 * (BL + BR) - (1.0 / (2.0 * ch)) * (psiR - psiL);
    *psiT = 0.5 * (psiL + psiR) - (ch / 2.0) * (BR - BL);
}

/* fast magnetosonic speed squared, MUSCL_NEW.cpp:140-165 / 197-222 (uses the
 * ORIGINAL normal field; the B?Tilde argument of the reference is unused) */
static double fast_speed_sq(const double *x, double g, int n) {
    double pr[NV];
    lsc_oracle_mhd_c2p(x, g, pr);
    double rho = x[0], p = pr[4];
    double Bx = x[5], By = x[6], Bz = x[7], Bn = x[n];
    double B2 = Bx * Bx + By * By + Bz * Bz;
    double cs2 = g * p / rho;
    return 0.5 * (cs2 + B2 / rho + sqrt((cs2 + B2 / rho) * (cs2 + B2 / rho) - 4.0 * cs2 * Bn * Bn / rho));
}

/* MUSCL_NEW.cpp:167-195 / 224-253; m = normal velocity index (1|2), n = normal B index (5|6) */
static void wavespeeds(const double *x, const double *y, double g, int m, int n, double *sL, double *qS, double *sR) {
    double cfa2L = fast_speed_sq(x, g, n), cfa2R = fast_speed_sq(y, g, n);
    double pL[NV], pR[NV];
    lsc_oracle_mhd_c2p(x, g, pL);
    lsc_oracle_mhd_c2p(y, g, pR);
    *sL = smin(pL[m], pR[m]) - smax(sqrt(cfa2L), sqrt(cfa2R));
    *sR = smax(pL[m], pR[m]) + smax(sqrt(cfa2L), sqrt(cfa2R));
    double B2L = x[5] * x[5] + x[6] * x[6] + x[7] * x[7];
    double pTL = pL[4] + 0.5 * B2L;
    double pModL = pTL - x[n] * x[n];
    double B2R = y[5] * y[5] + y[6] * y[6] + y[7] * y[7];
    double pTR = pR[4] + 0.5 * B2R;
    double pModR = pTR - y[n] * y[n];
    *qS = (y[0] * pR[m] * (*sR - pR[m]) - x[0] * pL[m] * (*sL - pL[m]) + pModL - pModR) /
          (y[0] * (*sR - pR[m]) - x[0] * (*sL - pL[m]));
}

/* MUSCL_NEW.cpp:255-307 */
static void hll_state(const double *x, const double *y, double g, double ch, int dir, double *out) {
    double sL, qS, sR, FL[NV], FR[NV], us[NV];
    if (dir == 0) {
        wavespeeds(x, y, g, 1, 5, &sL, &qS, &sR);
        lsc_oracle_mhd_fluxdef_x(x, g, ch, FL);
        lsc_oracle_mhd_fluxdef_x(y, g, ch, FR);
    } else {
        wavespeeds(x, y, g, 2, 6, &sL, &qS, &sR);
        lsc_oracle_mhd_fluxdef_y(x, g, ch, FL);
        lsc_oracle_mhd_fluxdef_y(y, g, ch, FR);
    }
    for (int k = 0; k < NV; ++k) us[k] = (sR * y[k] - sL * x[k] - FR[k] + FL[k]) / (sR - sL);
    const double *src = (sL >= 0) ? x : ((sL < 0 && sR > 0) ? us : y);
    memcpy(out, src, sizeof(double) * NV);
}

/* MUSCL_NEW.cpp:309-395 */
static void star_states_x(const double *x, const double *y, double g, double ch, double BxT, double psiT,
                          double *usL, double *usR) {
    double sL, qS, sR, pL[NV], pR[NV], hll[NV];
    wavespeeds(x, y, g, 1, 5, &sL, &qS, &sR);
    lsc_oracle_mhd_c2p(x, g, pL);
    lsc_oracle_mhd_c2p(y, g, pR);
    hll_state(x, y, g, ch, 0, hll);
    double rhoL = x[0], uxL = pL[1], uyL = pL[2], uzL = pL[3], EL = x[4], prL = pL[4];
    double BxL = BxT, ByL = x[6], BzL = x[7];
    double rhoR = y[0], uxR = pR[1], uyR = pR[2], uzR = pR[3], ER = y[4], prR = pR[4];
    double BxR = BxT, ByR = y[6], BzR = y[7];
    double ByH = hll[6], BzH = hll[7];
    double B2L = BxL * BxL + ByL * ByL + BzL * BzL;
    double B2R = BxR * BxR + ByR * ByR + BzR * BzR;
    double pTL = prL + 0.5 * B2L, pTR = prR + 0.5 * B2R;
    double BxS = (sR * BxR - sL * BxL) / (sR - sL);
    double pSL = rhoL * (sL - uxL) * (qS - uxL) + pTL - BxL * BxL + BxS * BxS;
    double pSR = rhoR * (sR - uxR) * (qS - uxR) + pTR - BxR * BxR + BxS * BxS;
    double rSL = rhoL * (sL - uxL) / (sL - qS);
    double rSR = rhoR * (sR - uxR) / (sR - qS);
    double mxSL = rSL * qS;
    double mySL = uyL * rSL - (BxS * ByH - BxL * ByL) / (sL - qS);
    double mzSL = uzL * rSL - (BxS * BzH - BxL * BzL) / (sL - qS);
    double mxSR = rSR * qS;
    double mySR = uyR * rSR - (BxS * ByH - BxR * ByR) / (sR - qS);
    double mzSR = uzR * rSR - (BxS * BzH - BxR * BzR) / (sR - qS);
    double vBL = BxL * uxL + ByL * uyL + BzL * uzL;
    double vBR = BxR * uxR + ByR * uyR + BzR * uzR;
    double vBHL = (BxS * mxSL + ByH * mySL + BzH * mzSL) / rSL;
    double vBHR = (BxS * mxSR + ByH * mySR + BzH * mzSR) / rSR;
    double ESL = EL * (rSL / rhoL) + ((pSL * qS - pTL * uxL) - (BxS * vBHL - BxL * vBL)) / (sL - qS);
    double ESR = ER * (rSR / rhoR) + ((pSR * qS - pTR * uxR) - (BxS * vBHR - BxR * vBR)) / (sR - qS);
    usL[0] = rSL; usL[1] = mxSL; usL[2] = mySL; usL[3] = mzSL; usL[4] = ESL;
    usL[5] = BxS; usL[6] = ByH; usL[7] = BzH; usL[8] = psiT;
    usR[0] = rSR; usR[1] = mxSR; usR[2] = mySR; usR[3] = mzSR; usR[4] = ESR;
    usR[5] = BxS; usR[6] = ByH; usR[7] = BzH; usR[8] = psiT;
}

/* MUSCL_NEW.cpp:397-481.  NOTE :442 -- the right z-momentum is built from the
 * LEFT quantities in the reference; kept, it is load-bearing for parity. */
static void star_states_y(const double *x, const double *y, double g, double ch, double ByT, double psiT,
                          double *usL, double *usR) {
    double sL, qS, sR, pL[NV], pR[NV], hll[NV];
    wavespeeds(x, y, g, 2, 6, &sL, &qS, &sR);
    lsc_oracle_mhd_c2p(x, g, pL);
    lsc_oracle_mhd_c2p(y, g, pR);
    hll_state(x, y, g, ch, 1, hll);
    double rhoL = x[0], uxL = pL[1], uyL = pL[2], uzL = pL[3], EL = x[4], prL = pL[4];
    double BxL = x[5], ByL = ByT, BzL = x[7];
    double rhoR = y[0], uxR = pR[1], uyR = pR[2], uzR = pR[3], ER = y[4], prR = pR[4];
    double BxR = y[5], ByR = ByT, BzR = y[7];
    double BxH = hll[5], BzH = hll[7];
    double B2L = BxL * BxL + ByL * ByL + BzL * BzL;
    double B2R = BxR * BxR + ByR * ByR + BzR * BzR;
    double pTL = prL + 0.5 * B2L, pTR = prR + 0.5 * B2R;
    double ByS = (sR * ByR - sL * ByL) / (sR - sL);
    double pSL = rhoL * (sL - uyL) * (qS - uyL) + pTL - ByL * ByL + ByS * ByS;
    double pSR = rhoR * (sR - uyR) * (qS - uyR) + pTR - ByR * ByR + ByS * ByS;
    double rSL = rhoL * (sL - uyL) / (sL - qS);
    double rSR = rhoR * (sR - uyR) / (sR - qS);
    double mxSL = rSL * uxL - (BxH * ByS - BxL * ByL) / (sL - qS);
    double mySL = rSL * qS;
    double mzSL = rSL * uzL - (ByS * BzH - ByL * BzL) / (sL - qS);
    double mxSR = rSR * uxR - (BxH * ByS - BxR * ByR) / (sR - qS);
    double mySR = rSR * qS;
    double mzSR = rSL * uzL - (ByS * BzH - ByL * BzL) / (sL - qS); /* sic, :442 */
    double vBL = BxL * uxL + ByL * uyL + BzL * uzL;
    double vBR = BxR * uxR + ByR * uyR + BzR * uzR;
    double vBHL = (BxH * mxSL + ByS * mySL + BzH * mzSL) / rSL;
    double vBHR = (BxH * mxSR + ByS * mySR + BzH * mzSR) / rSR;
    double ESL = EL * (rSL / rhoL) + ((pSL * qS - pTL * uyL) - (ByS * vBHL - ByL * vBL)) / (sL - qS);
    double ESR = ER * (rSR / rhoR) + ((pSR * qS - pTR * uyR) - (ByS * vBHR - ByR * vBR)) / (sR - qS);
    usL[0] = rSL; usL[1] = mxSL; usL[2] = mySL; usL[3] = mzSL; usL[4] = ESL;
    usL[5] = BxH; usL[6] = ByS; usL[7] = BzH; usL[8] = psiT;
    usR[0] = rSR; usR[1] = mxSR; usR[2] = mySR; usR[3] = mzSR; usR[4] = ESR;
    usR[5] = BxH; usR[6] = ByS; usR[7] = BzH; usR[8] = psiT;
}

/* MUSCL_NEW.cpp:483-549; dir 0 = x, 1 = y */
void lsc_oracle_mhd_hllc(const double *x, const double *y, double g, double ch, int dir, double *F) {
    const int n = dir == 0 ? 5 : 6;
    double BnT, psiT, usL[NV], usR[NV], sL, qS, sR, x1[NV], y1[NV], FL[NV], FR[NV];
    glm_interface(x, y, ch, n, &BnT, &psiT);
    if (dir == 0) {
        star_states_x(x, y, g, ch, BnT, psiT, usL, usR);
        wavespeeds(x, y, g, 1, 5, &sL, &qS, &sR);
    } else {
        star_states_y(x, y, g, ch, BnT, psiT, usL, usR);
        wavespeeds(x, y, g, 2, 6, &sL, &qS, &sR);
    }
    memcpy(x1, x, sizeof x1);
    memcpy(y1, y, sizeof y1);
    x1[n] = y1[n] = BnT;
    x1[8] = y1[8] = psiT;
    if (dir == 0) {
        lsc_oracle_mhd_fluxdef_x(x1, g, ch, FL);
        lsc_oracle_mhd_fluxdef_x(y1, g, ch, FR);
    } else {
        lsc_oracle_mhd_fluxdef_y(x1, g, ch, FL);
        lsc_oracle_mhd_fluxdef_y(y1, g, ch, FR);
    }
    if (sL >= 0) {
        memcpy(F, FL, sizeof FL);
    } else if (sL <= 0 && qS >= 0) {
        for (int k = 0; k < NV; ++k) F[k] = FL[k] + sL * (usL[k] - x1[k]);
    } else if (qS <= 0 && sR >= 0) {
        for (int k = 0; k < NV; ++k) F[k] = FR[k] + sR * (usR[k] - y1[k]);
    } else {
        memcpy(F, FR, sizeof FR);
    }
}

/* ---------------------------------------------------------------------- */
/* grid-level pieces                                                       */
/* ---------------------------------------------------------------------- */
#define AT(a, i, j) ((a) + ((size_t)(i) * (size_t)(m->c.ny + 4) + (size_t)(j)) * NV)

/* MUSCL_NEW.cpp:607-655 (loop bounds `< n + 2` kept: they stop two short) */
static void apply_bc(const mhd_t *m, double *a) {
    const int nx = m->c.nx, ny = m->c.ny;
    const size_t cell = sizeof(double) * NV;
    if (m->c.bcx == 0) {
        for (int j = 0; j < ny + 2; ++j) {
            memcpy(AT(a, 0, j), AT(a, nx, j), cell);
            memcpy(AT(a, 1, j), AT(a, nx + 1, j), cell);
            memcpy(AT(a, nx + 2, j), AT(a, 2, j), cell);
            memcpy(AT(a, nx + 3, j), AT(a, 3, j), cell);
        }
    } else {
        for (int j = 0; j < ny + 2; ++j) {
            memcpy(AT(a, 0, j), AT(a, 2, j), cell);
            memcpy(AT(a, 1, j), AT(a, 3, j), cell);
            memcpy(AT(a, nx + 2, j), AT(a, nx, j), cell);
            memcpy(AT(a, nx + 3, j), AT(a, nx + 1, j), cell);
        }
    }
    for (int i = 0; i < nx + 2; ++i) {
        if (m->c.bcy == 1) {
            memcpy(AT(a, i, 0), AT(a, i, ny), cell);
            memcpy(AT(a, i, 1), AT(a, i, ny + 1), cell);
            memcpy(AT(a, i, ny + 2), AT(a, i, 2), cell);
            memcpy(AT(a, i, ny + 3), AT(a, i, 3), cell);
        } else {
            memcpy(AT(a, i, 0), AT(a, i, 2), cell);
            memcpy(AT(a, i, 1), AT(a, i, 3), cell);
            memcpy(AT(a, i, ny + 2), AT(a, i, ny), cell);
            memcpy(AT(a, i, ny + 3), AT(a, i, ny + 1), cell);
            if (m->c.bcy == 0) {
                AT(a, i, 0)[2] = -AT(a, i, 2)[2];
                AT(a, i, 0)[6] = -AT(a, i, 2)[6];
                AT(a, i, 1)[2] = -AT(a, i, 3)[2];
                AT(a, i, 1)[6] = -AT(a, i, 3)[6];
                AT(a, i, ny + 2)[2] = -AT(a, i, ny)[2];
                AT(a, i, ny + 2)[6] = -AT(a, i, ny)[6];
                AT(a, i, ny + 3)[2] = -AT(a, i, ny + 1)[2];
                AT(a, i, ny + 3)[6] = -AT(a, i, ny + 1)[6];
            }
        }
    }
}

/* MUSCL_NEW.cpp:71-120; range i in [1,nx+1], j in [1,ny+1] (`< size-2`) */
static double max_speed(const mhd_t *m) {
    const double g = m->c.gamma;
    double best = 0.0;
    for (int i = 1; i < m->c.nx + 2; ++i)
        for (int j = 1; j < m->c.ny + 2; ++j) {
            const double *q = AT(m->u, i, j);
            double rho = q[0], Bx = q[5], By = q[6], Bz = q[7];
            double ux = q[1] / rho, uy = q[2] / rho, uz = q[3] / rho;
            double ke = 0.5 * rho * (ux * ux + uy * uy + uz * uz) + 0.5 * (Bx * Bx + By * By + Bz * Bz);
            double B2 = Bx * Bx + By * By + Bz * Bz;
            double p = (g - 1.0) * (q[4] - ke);
            double a = sqrt(g * p / rho);
            double s = a * a + B2 / rho; /* pow(s,2) == s*s */
            double cfx = sqrt(0.5 * (a * a + B2 / rho + sqrt(s * s - 4 * a * a * Bx * Bx / rho)));
            double cfy = sqrt(0.5 * (a * a + B2 / rho + sqrt(s * s - 4 * a * a * By * By / rho)));
            double cfz = sqrt(0.5 * (a * a + B2 / rho + sqrt(s * s - 4 * a * a * Bz * Bz / rho)));
            double sp = smax(smax(fabs(ux) + cfx, fabs(uy) + cfy), fabs(uz) + cfz);
            if (sp > best) best = sp;
        }
    return best;
}

/* limiter + linear reconstruction of one component, MUSCL_NEW.cpp:845-859 */
static void recon1(double pm, double p0, double pp, double um, double u0, double up, double *lo, double *hi) {
    double dPlus = pp - p0, dMinus = p0 - pm;
    double r = dMinus / (dPlus + 1e-8);
    dPlus = up - u0;
    dMinus = u0 - um;
    double xi;
    if (r <= 0) xi = 0;
    else xi = fmin(2.0 * r / (1 + r), 2.0 / (1 + r));
    double Delta = 0.5 * dMinus + 0.5 * dPlus;
    *lo = u0 - 0.5 * xi * Delta;
    *hi = u0 + 0.5 * xi * Delta;
}

/* one pass of the do{}while body, MUSCL_NEW.cpp:831-971 */
static void one_step(mhd_t *m) {
    const int nx = m->c.nx, ny = m->c.ny;
    const double g = m->c.gamma, dx = m->dx, dy = m->dy;
    double ms = max_speed(m);
    double dspace = smin(dx, dy);
    double dt = m->c.cfl * dspace / ms, ch = ms;
    m->dt = dt; m->ch = ch; m->t += dt;
    apply_bc(m, m->u);
    /* x reconstruct: v refreshed INSIDE the i loop => v[i+1] is one step stale */
    for (int i = 1; i < nx + 3; ++i)
        for (int j = 0; j < ny + 4; ++j) {
            lsc_oracle_mhd_c2p(AT(m->u, i, j), g, AT(m->v, i, j));
            for (int k = 0; k < NV; ++k)
                recon1(AT(m->v, i - 1, j)[k], AT(m->v, i, j)[k], AT(m->v, i + 1, j)[k],
                       AT(m->u, i - 1, j)[k], AT(m->u, i, j)[k], AT(m->u, i + 1, j)[k],
                       &AT(m->rxL, i, j)[k], &AT(m->rxR, i, j)[k]);
        }
    for (int i = 1; i < nx + 3; ++i)
        for (int j = 1; j < ny + 3; ++j) {
            double fl[NV], fr[NV];
            lsc_oracle_mhd_fluxdef_x(AT(m->rxL, i, j), g, ch, fl);
            lsc_oracle_mhd_fluxdef_x(AT(m->rxR, i, j), g, ch, fr);
            for (int k = 0; k < NV; ++k) {
                AT(m->hxL, i, j)[k] = AT(m->rxL, i, j)[k] - 0.5 * (dt / dx) * (fr[k] - fl[k]);
                AT(m->hxR, i, j)[k] = AT(m->rxR, i, j)[k] - 0.5 * (dt / dx) * (fr[k] - fl[k]);
            }
        }
    apply_bc(m, m->hxL);
    apply_bc(m, m->hxR);
    for (int i = 1; i < nx + 3; ++i)
        for (int j = 1; j < ny + 3; ++j)
            lsc_oracle_mhd_hllc(AT(m->hxR, i, j), AT(m->hxL, i + 1, j), g, ms, 0, AT(m->fx, i, j));
    for (int i = 2; i < nx + 2; ++i)
        for (int j = 2; j < ny + 2; ++j)
            for (int k = 0; k < NV; ++k)
                AT(m->ubar, i, j)[k] = AT(m->u, i, j)[k] - (dt / dx) * (AT(m->fx, i, j)[k] - AT(m->fx, i - 1, j)[k]);
    apply_bc(m, m->ubar);
    {
        double f = exp(-1.0 * (0.5 * dt) * ms / 0.18); /* psiUpdate, :66-69 with dt := 0.5*dt */
        for (size_t c = 0; c < m->ncell; ++c) m->ubar[c * NV + 8] = m->ubar[c * NV + 8] * f;
    }
    /* y reconstruct: limiter ratio from prim(u), slopes from uBar */
    for (int i = 0; i < nx + 4; ++i)
        for (int j = 1; j < ny + 3; ++j) {
            lsc_oracle_mhd_c2p(AT(m->u, i, j), g, AT(m->v, i, j));
            for (int k = 0; k < NV; ++k)
                recon1(AT(m->v, i, j - 1)[k], AT(m->v, i, j)[k], AT(m->v, i, j + 1)[k],
                       AT(m->ubar, i, j - 1)[k], AT(m->ubar, i, j)[k], AT(m->ubar, i, j + 1)[k],
                       &AT(m->ryB, i, j)[k], &AT(m->ryT, i, j)[k]);
        }
    for (int i = 1; i < nx + 3; ++i)
        for (int j = 1; j < ny + 3; ++j) {
            double fb[NV], ft[NV];
            lsc_oracle_mhd_fluxdef_y(AT(m->ryB, i, j), g, ch, fb);
            lsc_oracle_mhd_fluxdef_y(AT(m->ryT, i, j), g, ch, ft);
            for (int k = 0; k < NV; ++k) {
                AT(m->hyB, i, j)[k] = AT(m->ryB, i, j)[k] - 0.5 * (dt / dy) * (ft[k] - fb[k]);
                AT(m->hyT, i, j)[k] = AT(m->ryT, i, j)[k] - 0.5 * (dt / dy) * (ft[k] - fb[k]);
            }
        }
    apply_bc(m, m->hyT);
    apply_bc(m, m->hyB);
    for (int i = 1; i < nx + 3; ++i)
        for (int j = 1; j < ny + 3; ++j)
            lsc_oracle_mhd_hllc(AT(m->hyT, i, j), AT(m->hyB, i, j + 1), g, ms, 1, AT(m->fy, i, j));
    for (int i = 2; i < nx + 2; ++i)
        for (int j = 2; j < ny + 2; ++j)
            for (int k = 0; k < NV; ++k)
                AT(m->up1, i, j)[k] = AT(m->ubar, i, j)[k] - (dt / dy) * (AT(m->fy, i, j)[k] - AT(m->fy, i, j - 1)[k]);
    apply_bc(m, m->up1);
    {
        double f = exp(-1.0 * (0.5 * dt) * ms / 0.18);
        for (size_t c = 0; c < m->ncell; ++c) m->up1[c * NV + 8] = m->up1[c * NV + 8] * f;
    }
    memcpy(m->u, m->up1, m->ncell * NV * sizeof(double));
    apply_bc(m, m->u);
    m->step++;
}

/* ---------------------------------------------------------------------- */
/* C entry points used by the tests through ctypes                         */
/* ---------------------------------------------------------------------- */

/* initial conditions, MUSCL_NEW.cpp:719-825; prob 0 = KH (:800-808), 1 = OT (:811-819) */
void lsc_oracle_mhd_ic(const lsc_oracle_mhd_cfg *c, int prob, double *u) {
    const double pi = 3.14159265358979323846;
    const double g = c->gamma;
    const double dx = (c->x1 - c->x0) / c->nx, dy = (c->y1 - c->y0) / c->ny;
    for (int i = 0; i < c->nx + 4; ++i)
        for (int j = 0; j < c->ny + 4; ++j) {
            double x = c->x0 + (i - 0.5) * dx, y = c->y0 + (j - 0.5) * dy, pr[NV];
            if (prob == 0) {
                pr[0] = 1.0;
                pr[1] = 0.5 * tanh(20 * y);
                pr[2] = 0.01 * sin(2.0 * pi * x) * exp(-y * y / (0.01));
                pr[3] = 0.0;
                pr[4] = 1.0 / g;
                pr[5] = 0.1 * cos(pi / 3.0);
                pr[6] = 0.0;
                pr[7] = 0.1 * sin(pi / 3.0);
                pr[8] = 0;
            } else {
                pr[0] = g * g;
                pr[1] = -sin(2.0 * pi * y);
                pr[2] = sin(2.0 * pi * x);
                pr[3] = 0.0;
                pr[4] = g;
                pr[5] = -sin(2.0 * pi * y);
                pr[6] = sin(4.0 * pi * x);
                pr[7] = 0.0;
                pr[8] = 0;
            }
            lsc_oracle_mhd_p2c(pr, g, u + ((size_t)i * (c->ny + 4) + j) * NV);
        }
}

void *lsc_oracle_mhd_create(const lsc_oracle_mhd_cfg *c, const double *u0) {
    mhd_t *m = (mhd_t *)calloc(1, sizeof(mhd_t));
    m->c = *c;
    m->ncell = (size_t)(c->nx + 4) * (size_t)(c->ny + 4);
    m->dx = (c->x1 - c->x0) / c->nx;
    m->dy = (c->y1 - c->y0) / c->ny;
    double **arr[] = {&m->u, &m->v, &m->up1, &m->fx, &m->fy, &m->ubar, &m->rxL, &m->rxR,
                      &m->ryB, &m->ryT, &m->hxL, &m->hxR, &m->hyB, &m->hyT};
    for (size_t a = 0; a < sizeof(arr) / sizeof(arr[0]); ++a) *arr[a] = (double *)calloc(m->ncell * NV, sizeof(double));
    memcpy(m->u, u0, m->ncell * NV * sizeof(double));
    return m;
}

/* advance nsteps; steplog (may be NULL) receives (t, dt, ch) per step */
void lsc_oracle_mhd_step(void *h, long nsteps, double *steplog) {
    mhd_t *m = (mhd_t *)h;
    for (long s = 0; s < nsteps; ++s) {
        one_step(m);
        if (steplog) { steplog[3 * s] = m->t; steplog[3 * s + 1] = m->dt; steplog[3 * s + 2] = m->ch; }
    }
}

/* the reference's do{...}while(t<tStop) (:831, :1003): no clamp, overshoots */
long lsc_oracle_mhd_run_until(void *h, double t_stop) {
    mhd_t *m = (mhd_t *)h;
    long n = 0;
    do { one_step(m); ++n; } while (m->t < t_stop);
    return n;
}

void lsc_oracle_mhd_get(void *h, double *u, double *t, double *dt, double *ch) {
    mhd_t *m = (mhd_t *)h;
    if (u) memcpy(u, m->u, m->ncell * NV * sizeof(double));
    if (t) *t = m->t;
    if (dt) *dt = m->dt;
    if (ch) *ch = m->ch;
}

void lsc_oracle_mhd_destroy(void *h) {
    mhd_t *m = (mhd_t *)h;
    double *arr[] = {m->u, m->v, m->up1, m->fx, m->fy, m->ubar, m->rxL, m->rxR, m->ryB, m->ryT, m->hxL, m->hxR, m->hyB, m->hyT};
    for (size_t a = 0; a < sizeof(arr) / sizeof(arr[0]); ++a) free(arr[a]);
    free(m);
}
