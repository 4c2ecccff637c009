/*
 * oracle/euler_oracle.c -- CPU restatement of the reference's 2-D Euler SLIC step
 * (Euler_codes/SLIC_euler_2d.cpp = S, transmissive variant Euler_codes/generaliseSLIC.cpp = G).
 *
 * and of the tabulated-EoS variant MHD_codes/TabEoS.cpp = T (same sweeps, EoS by table lookup, T:147-438).
 *
 * TEST INFRASTRUCTURE ONLY (see mhd_oracle.c).  Parity pin: bit-identical to the real reference programs
 * compiled by oracle/build_ref_euler.py (oracle/_ref/slic_ref, gslic_ref); slic_ref reproduces the
 * reference's committed MHD_plot_data/euler_{1..20}.dat byte-for-byte.  Table path: bit-identical to
 * oracle/_ref/tabeos_ref (the real TabEoS.cpp) run on the synthetic table of tests/oracle_util.py; the real
 * Plasma19_EoS.txt is not in the reference tree, so that path is NOT pinned against the committed SLIC2d.dat.
 *
 * Layout: AoS u[(i*(ny+4)+j)*4 + k], k = rho, rho*vx, rho*vy, E (S:10-17), two ghosts per side.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

#define NE 4

typedef struct {
    int nx, ny;
    double x0, x1, y0, y1;
    double gamma, cfl, omega;
    int bc; /* 0 periodic (S:198-226), 1 transmissive (G:199-215) */
} lsc_oracle_euler_cfg;

/* Tabulated EoS of MHD_codes/TabEoS.cpp (= T): 500 densities, 500 pressures, sound speed and specific
 * internal energy as [pressure index][density index] (T:18-133).  NULL = ideal gas (S / G). */
#define NT 500
typedef struct {
    const double *dens, *pres, *sound, *en;
} eos_table_t;

typedef struct {
    lsc_oracle_euler_cfg c;
    size_t ncell;
    double dx, dy, t, dt;
    long step;
    double *u, *A, *B, *flux, *ubar, *rL, *rR, *hL, *hR;
    eos_table_t tab;  /* tab.dens == NULL: ideal gas */
    double *tab_store;
} euler_t;

#define EAT(a, i, j) ((a) + ((size_t)(i) * (size_t)(ny + 4) + (size_t)(j)) * NE)

/* S:10-17 */
void lsc_oracle_euler_p2c(const double *p, double g, double *c) {
    c[0] = p[0];
    c[1] = p[0] * p[1];
    c[2] = p[0] * p[2];
    c[3] = p[3] / (g - 1) + (0.5 * p[0] * (p[1] * p[1] + p[2] * p[2]));
}

/* S:30-44 (dir 0) and S:46-60 (dir 1) */
void lsc_oracle_euler_fluxdef(const double *x, double g, int dir, double *f) {
    double rho = x[0];
    double vx = x[1] / x[0];
    double vy = x[2] / x[0];
    double KE = 0.5 * rho * (vx * vx + vy * vy);
    double p = (g - 1) * (x[3] - KE);
    if (dir == 0) {
        f[0] = x[1];
        f[1] = rho * vx * vx + p;
        f[2] = rho * vx * vy;
        f[3] = (x[3] + p) * vx;
    } else {
        f[0] = x[2];
        f[1] = rho * vx * vy;
        f[2] = rho * vy * vy + p;
        f[3] = (x[3] + p) * vy;
    }
}

/* ---- tabulated EoS, T:147-404 ------------------------------------------------------------------------ */
/* T:147-153 */
static double tab_bilinear(double f_NE, double f_SE, double f_SW, double f_NW, double x, double y) {
    double f_x_y1 = (1 - x) * f_SW + x * f_SE;
    double f_x_y2 = (1 - x) * f_NW + x * f_NE;
    return (1 - y) * f_x_y1 + y * f_x_y2;
}

/* T:156-163 */
static double tab_reverse_bilinear(double f_NE, double f_SE, double f_SW, double f_NW, double x, double f_x_y) {
    double f_x_y1 = (1 - x) * f_SW + x * f_SE;
    double f_x_y2 = (1 - x) * f_NW + x * f_NE;
    double y = (f_x_y - f_x_y1) / (f_x_y2 - f_x_y1);
    if (f_x_y2 == f_x_y1) y = 1;
    return y;
}

/* the bisection every lookup starts with (T:174-191, :211-228, :257-275, :339-356, :364-384): brackets `v` in
 * the ascending axis `a`; point2 starts at 499, or at 500 in SoundSpeed's pressure search (T:366) */
static void tab_bracket(const double *a, double v, int point2, int *lo, int *hi) {
    int point1 = 0, half;
    while (abs(point1 - point2) > 1) {
        half = (point1 + point2) / 2;
        if (v > a[half]) point1 = half;
        else point2 = half;
    }
    if (point1 >= 499) point1 = 499;
    if (point2 >= 499) point2 = 499;
    if (point1 <= 0) point1 = 0;
    if (point2 <= 0) point2 = 0;
    *lo = point1;
    *hi = point2;
}

/* PrimativeToConservative, T:166-246 */
void lsc_oracle_tabeos_p2c(const eos_table_t *tb, const double *prim, double *consv) {
    consv[0] = prim[0];
    consv[1] = prim[0] * prim[1];
    consv[2] = prim[0] * prim[2];
    double rho = prim[0], p = prim[3];
    int db, dt_, pb, pt;
    tab_bracket(tb->dens, rho, 499, &db, &dt_);
    double density_ratio = (rho - tb->dens[db]) / (tb->dens[dt_] - tb->dens[db]);
    if (tb->dens[db] == tb->dens[dt_]) density_ratio = 1;
    tab_bracket(tb->pres, p, 499, &pb, &pt);
    double pressure_ratio = (p - tb->pres[pb]) / (tb->pres[pt] - tb->pres[pb]);
    if (tb->pres[pb] == tb->pres[pt]) pressure_ratio = 1;
    double energy = tab_bilinear(tb->en[pt * NT + dt_], tb->en[pb * NT + dt_], tb->en[pb * NT + db], tb->en[pt * NT + db],
                                 density_ratio, pressure_ratio);
    consv[3] = energy * rho + 0.5 * rho * (prim[1] * prim[1] + prim[2] * prim[2]);
}

/* ConservativeToPrimative, T:248-327.  The pressure bracket is captured BEFORE the clamps (T:303-310). */
void lsc_oracle_tabeos_c2p(const eos_table_t *tb, const double *consv, double *prim) {
    prim[0] = consv[0];
    prim[1] = consv[1] / consv[0];
    prim[2] = consv[2] / consv[0];
    double rho = prim[0];
    int db, dt_;
    tab_bracket(tb->dens, rho, 499, &db, &dt_);
    double density_ratio = (rho - tb->dens[db]) / (tb->dens[dt_] - tb->dens[db]);
    if (tb->dens[db] == tb->dens[dt_]) density_ratio = 1;
    double e = (consv[3] - 0.5 * rho * (prim[1] * prim[1] + prim[2] * prim[2])) / rho;
    int point1 = 0, point2 = 499, half;
    while (abs(point1 - point2) > 1) {
        half = (point1 + point2) / 2;
        if (e > (tb->en[half * NT + db] + density_ratio * (tb->en[half * NT + dt_] - tb->en[half * NT + db]))) point1 = half;
        else point2 = half;
    }
    int pt = point2, pb = point1;
    double pressure_ratio = tab_reverse_bilinear(tb->en[pt * NT + dt_], tb->en[pb * NT + dt_], tb->en[pb * NT + db],
                                                 tb->en[pt * NT + db], density_ratio, e);
    prim[3] = pressure_ratio * (tb->pres[pt] - tb->pres[pb]) + tb->pres[pb];
}

/* SoundSpeed, T:330-404 */
double lsc_oracle_tabeos_sound(const eos_table_t *tb, const double *u) {
    double prim[NE];
    lsc_oracle_tabeos_c2p(tb, u, prim);
    double rho = u[0], p = prim[3];
    int db, dt_, pb, pt;
    tab_bracket(tb->dens, rho, 499, &db, &dt_);
    double density_ratio = (rho - tb->dens[db]) / (tb->dens[dt_] - tb->dens[db]);
    if (tb->dens[db] == tb->dens[dt_]) density_ratio = 1;
    tab_bracket(tb->pres, p, 500, &pb, &pt); /* T:366: starts at 500; index 500 is never read */
    double pressure_ratio = (p - tb->pres[pb]) / (tb->pres[pt] - tb->pres[pb]);
    if (tb->pres[pb] == tb->pres[pt]) pressure_ratio = 1;
    return tab_bilinear(tb->sound[pt * NT + dt_], tb->sound[pb * NT + dt_], tb->sound[pb * NT + db],
                        tb->sound[pt * NT + db], density_ratio, pressure_ratio);
}

/* fluxX_def / fluxY_def with the table, T:408-438 */
void lsc_oracle_tabeos_fluxdef(const eos_table_t *tb, const double *x, int dir, double *f) {
    double prim[NE];
    lsc_oracle_tabeos_c2p(tb, x, prim);
    double rho = prim[0], v_x = prim[1], v_y = prim[2], p = prim[3];
    if (dir == 0) {
        f[0] = x[1];
        f[1] = rho * v_x * v_x + p;
        f[2] = rho * v_x * v_y;
        f[3] = (x[3] + p) * v_x;
    } else {
        f[0] = x[2];
        f[1] = rho * v_x * v_y;
        f[2] = rho * v_y * v_y + p;
        f[3] = (x[3] + p) * v_y;
    }
}

static void fluxdef_eos(const eos_table_t *tb, const double *x, double g, int dir, double *f) {
    if (tb && tb->dens) lsc_oracle_tabeos_fluxdef(tb, x, dir, f);
    else lsc_oracle_euler_fluxdef(x, g, dir, f);
}

/* FORCE flux, S:111-136 (x) / S:139-163 (y), T:440-492; h = dx or dy */
static void force_eos(const eos_table_t *tb, const double *x, const double *y, double h, double dt, double g, int dir,
                      double *F) {
    double f1[NE], f2[NE], uh[NE], ri[NE];
    fluxdef_eos(tb, x, g, dir, f1);
    fluxdef_eos(tb, y, g, dir, f2);
    for (int k = 0; k < NE; ++k) uh[k] = 0.5 * (x[k] + y[k]) - 0.5 * (dt / h) * (f2[k] - f1[k]);
    fluxdef_eos(tb, uh, g, dir, ri);
    for (int k = 0; k < NE; ++k) {
        double lf = (0.5 * h / dt) * (x[k] - y[k]) + 0.5 * (f1[k] + f2[k]);
        F[k] = 0.5 * (lf + ri[k]);
    }
}

void lsc_oracle_euler_force(const double *x, const double *y, double h, double dt, double g, int dir, double *F) {
    force_eos(NULL, x, y, h, dt, g, dir, F);
}

/* S:198-226 (periodic, full loop ranges) / G:199-215 (transmissive) */
static void ebc(const euler_t *m, double *a) {
    const int nx = m->c.nx, ny = m->c.ny;
    const size_t cell = sizeof(double) * NE;
    for (int j = 0; j < ny + 4; ++j) {
        if (m->c.bc == 0) {
            memcpy(EAT(a, 0, j), EAT(a, nx, j), cell);
            memcpy(EAT(a, 1, j), EAT(a, nx + 1, j), cell);
            memcpy(EAT(a, nx + 2, j), EAT(a, 2, j), cell);
            memcpy(EAT(a, nx + 3, j), EAT(a, 3, j), cell);
        } else {
            memcpy(EAT(a, 0, j), EAT(a, 2, j), cell);
            memcpy(EAT(a, 1, j), EAT(a, 2, j), cell);
            memcpy(EAT(a, nx + 2, j), EAT(a, nx + 1, j), cell);
            memcpy(EAT(a, nx + 3, j), EAT(a, nx + 1, j), cell);
        }
    }
    for (int i = 0; i < nx + 4; ++i) {
        if (m->c.bc == 0) {
            memcpy(EAT(a, i, 0), EAT(a, i, ny), cell);
            memcpy(EAT(a, i, 1), EAT(a, i, ny + 1), cell);
            memcpy(EAT(a, i, ny + 2), EAT(a, i, 2), cell);
            memcpy(EAT(a, i, ny + 3), EAT(a, i, 3), cell);
        } else {
            memcpy(EAT(a, i, 0), EAT(a, i, 2), cell);
            memcpy(EAT(a, i, 1), EAT(a, i, 2), cell);
            memcpy(EAT(a, i, ny + 2), EAT(a, i, ny + 1), cell);
            memcpy(EAT(a, i, ny + 3), EAT(a, i, ny + 1), cell);
        }
    }
}

/* S:165-195: range i in [2,nx+2], j in [2,ny+2] (`<= size-2`) */
static double euler_dt(const euler_t *m) {
    const int nx = m->c.nx, ny = m->c.ny;
    const double g = m->c.gamma;
    double dspace = m->dx;
    if (m->dy < m->dx) dspace = m->dy;
    double best = 0.0;
    for (int i = 2; i <= nx + 2; ++i)
        for (int j = 2; j <= ny + 2; ++j) {
            const double *q = EAT(m->u, i, j);
            if (m->tab.dens) { /* T:505-521 */
                double prim[NE];
                lsc_oracle_tabeos_c2p(&m->tab, q, prim);
                double speed = sqrt(prim[1] * prim[1] + prim[2] * prim[2]);
                double sound_speed = lsc_oracle_tabeos_sound(&m->tab, q);
                double new_speed = speed + sound_speed;
                if (new_speed > best) best = new_speed;
                continue;
            }
            double rho = q[0], vx = q[1] / rho, vy = q[2] / rho;
            double speed = sqrt(vx * vx + vy * vy);
            double KE = 0.5 * rho * speed * speed;
            double p = (g - 1.0) * (q[3] - KE);
            double a = sqrt(g * p / rho);
            double s = speed + a;
            if (s > best) best = s;
        }
    return m->c.cfl * dspace / best;
}

/* limiter + reconstruction of one component, S:262-283 */
static void erecon1(double um, double u0, double up, double omega, double *lo, double *hi) {
    double DeltaPlus = up - u0;
    double DeltaMinus = u0 - um;
    double r = DeltaMinus / (DeltaPlus + 1e-8);
    double xi_R = 2.0 / (1 + r);
    double xi;
    double Delta = 0.5 * (1 + omega) * DeltaMinus + 0.5 * (1 - omega) * DeltaPlus;
    if (r <= 0) xi = 0;
    else if (r > 0 && r <= 1) xi = r;
    else xi = fmin(1, xi_R);
    *lo = u0 - 0.5 * xi * Delta;
    *hi = u0 + 0.5 * xi * Delta;
}

/* One directional SLIC sweep of `in` into `out` (interior only), as each half of XthenY / YthenX does:
 * reconstruct (S:260-287), half step (S:289-296), BC on the half states (S:297-298), FORCE flux on
 * i,j in [1,n+2] (S:300-304), flux difference (S:306-312).  The work arrays are zeroed first because the
 * reference allocates them afresh in every call (S:232-254). */
static void sweep(euler_t *m, const double *in, double *out, int dir) {
    const int nx = m->c.nx, ny = m->c.ny;
    const double g = m->c.gamma, dt = m->dt;
    const double h = dir == 0 ? m->dx : m->dy;
    const size_t bytes = m->ncell * NE * sizeof(double);
    const int di = dir == 0, dj = dir == 1;
    memset(m->rL, 0, bytes); memset(m->rR, 0, bytes);
    memset(m->hL, 0, bytes); memset(m->hR, 0, bytes);
    memset(m->flux, 0, bytes);
    for (int j = 2; j < ny + 2; ++j)
        for (int i = 2; i < nx + 2; ++i)
            for (int k = 0; k < NE; ++k)
                erecon1(EAT(in, i - di, j - dj)[k], EAT(in, i, j)[k], EAT(in, i + di, j + dj)[k], m->c.omega,
                        &EAT(m->rL, i, j)[k], &EAT(m->rR, i, j)[k]);
    for (int j = 2; j < ny + 2; ++j)
        for (int i = 2; i < nx + 2; ++i) {
            double fl[NE], fr[NE];
            fluxdef_eos(&m->tab, EAT(m->rL, i, j), g, dir, fl);
            fluxdef_eos(&m->tab, EAT(m->rR, i, j), g, dir, fr);
            for (int k = 0; k < NE; ++k) {
                EAT(m->hL, i, j)[k] = EAT(m->rL, i, j)[k] - 0.5 * (dt / h) * (fr[k] - fl[k]);
                EAT(m->hR, i, j)[k] = EAT(m->rR, i, j)[k] - 0.5 * (dt / h) * (fr[k] - fl[k]);
            }
        }
    ebc(m, m->hL);
    ebc(m, m->hR);
    for (int j = 1; j < ny + 3; ++j)
        for (int i = 1; i < nx + 3; ++i)
            force_eos(&m->tab, EAT(m->hR, i, j), EAT(m->hL, i + di, j + dj), h, dt, g, dir, EAT(m->flux, i, j));
    for (int j = 2; j < ny + 2; ++j)
        for (int i = 2; i < nx + 2; ++i)
            for (int k = 0; k < NE; ++k)
                EAT(out, i, j)[k] = EAT(in, i, j)[k] - (dt / h) * (EAT(m->flux, i, j)[k] - EAT(m->flux, i - di, j - dj)[k]);
}

/* body of the do-loop S:600-616 */
static void euler_step(euler_t *m) {
    const int nx = m->c.nx, ny = m->c.ny;
    const size_t bytes = m->ncell * NE * sizeof(double);
    m->dt = euler_dt(m);
    m->t += m->dt;
    ebc(m, m->u);
    /* XthenY (S:230-376): x sweep, BC, y sweep */
    memset(m->ubar, 0, bytes);
    memset(m->A, 0, bytes);
    sweep(m, m->u, m->ubar, 0);
    ebc(m, m->ubar);
    sweep(m, m->ubar, m->A, 1);
    /* YthenX (S:379-525): y sweep; its x sweep stores the flux in fluxY (S:510) but updates from the all-zero
     * fluxX (S:519), so the result is uPlus1Bar - (dt/dx)*(0 - 0) */
    memset(m->ubar, 0, bytes);
    memset(m->B, 0, bytes);
    sweep(m, m->u, m->ubar, 1);
    for (int j = 2; j < ny + 2; ++j)
        for (int i = 2; i < nx + 2; ++i)
            for (int k = 0; k < NE; ++k)
                EAT(m->B, i, j)[k] = EAT(m->ubar, i, j)[k] - (m->dt / m->dx) * (0.0 - 0.0);
    for (int j = 2; j < ny + 2; ++j)
        for (int i = 2; i < nx + 2; ++i)
            for (int k = 0; k < NE; ++k) EAT(m->u, i, j)[k] = 0.5 * (EAT(m->A, i, j)[k] + EAT(m->B, i, j)[k]);
    ebc(m, m->u);
    m->step++;
}

/* IC: 0 = S:585-589 (Euler "Orszag-Tang"), 1 = G:594-617 (four quadrants), 2 = cylindrical explosion of
 * BASELINE config 3 (not in the reference).  Cell centres x0+(i-1.5)dx (S:556-557);
 * the boundary function is applied once after the fill (S:596). */
void lsc_oracle_euler_ic(const lsc_oracle_euler_cfg *c, int ic, double *u) {
    const double pi = 3.14159265358979323846;
    const int nx = c->nx, ny = c->ny;
    const double g = c->gamma;
    const double dx = (c->x1 - c->x0) / nx, dy = (c->y1 - c->y0) / ny;
    for (int i = 0; i < nx + 4; ++i)
        for (int j = 0; j < ny + 4; ++j) {
            double x = c->x0 + (i - 1.5) * dx, y = c->y0 + (j - 1.5) * dy, pr[NE];
            if (ic == 0) {
                pr[0] = g * g; pr[1] = -sin(2.0 * pi * y); pr[2] = sin(2.0 * pi * x); pr[3] = g;
            } else if (ic == 1) {
                if (x > 0.5 && y > 0.5) { pr[0] = 1.5; pr[1] = 0; pr[2] = 0; pr[3] = 1.5; }
                else if (x <= 0.5 && y > 0.5) { pr[0] = 0.5323; pr[1] = 1.206; pr[2] = 0; pr[3] = 0.3; }
                else if (x <= 0.5 && y <= 0.5) { pr[0] = 0.138; pr[1] = 1.206; pr[2] = 1.206; pr[3] = 0.029; }
                else { pr[0] = 0.5323; pr[1] = 0; pr[2] = 1.206; pr[3] = 0.3; }
            } else {
                if ((x - 0.5) * (x - 0.5) + (y - 0.5) * (y - 0.5) <= 0.2 * 0.2) { pr[0] = 1.0; pr[1] = 0; pr[2] = 0; pr[3] = 1.0; }
                else { pr[0] = 0.125; pr[1] = 0; pr[2] = 0; pr[3] = 0.1; }
            }
            lsc_oracle_euler_p2c(pr, g, EAT(u, i, j));
        }
    euler_t tmp;
    tmp.c = *c;
    ebc(&tmp, u);
}

/* IC of TabEoS.cpp main(), T:918-941: four quadrants at rest, rho = 1, p = 0.01e5 (x > 0.5) / 1000e5 (x <= 0.5),
 * PrimativeToConservative through the table, then the transmissive boundary fill (T:954). */
void lsc_oracle_tabeos_ic(const lsc_oracle_euler_cfg *c, const double *dens, const double *pres, const double *sound,
                          const double *en, double *u) {
    eos_table_t tb = {dens, pres, sound, en};
    const int nx = c->nx, ny = c->ny;
    const double dx = (c->x1 - c->x0) / nx, dy = (c->y1 - c->y0) / ny;
    for (int i = 0; i < nx + 4; ++i)
        for (int j = 0; j < ny + 4; ++j) {
            double x = c->x0 + (i - 1.5) * dx, y = c->y0 + (j - 1.5) * dy, pr[NE];
            pr[0] = 1.0; pr[1] = 0; pr[2] = 0;
            if (x > 0.5 && y > 0.5) pr[3] = 0.01 * 1e5;
            else if (x <= 0.5 && y > 0.5) pr[3] = 1000 * 1e5;
            else if (x <= 0.5 && y <= 0.5) pr[3] = 1000 * 1e5;
            else pr[3] = 0.01 * 1e5;
            lsc_oracle_tabeos_p2c(&tb, pr, EAT(u, i, j));
        }
    euler_t tmp;
    tmp.c = *c;
    tmp.c.bc = 1;
    ebc(&tmp, u);
}

/* attach a copy of the four tables to an oracle instance: every flux and the time step then go through T's EoS */
void lsc_oracle_euler_set_table(void *h, const double *dens, const double *pres, const double *sound, const double *en) {
    euler_t *m = (euler_t *)h;
    free(m->tab_store);
    m->tab_store = (double *)malloc(sizeof(double) * (2 * NT + 2 * NT * NT));
    memcpy(m->tab_store, dens, sizeof(double) * NT);
    memcpy(m->tab_store + NT, pres, sizeof(double) * NT);
    memcpy(m->tab_store + 2 * NT, sound, sizeof(double) * NT * NT);
    memcpy(m->tab_store + 2 * NT + NT * NT, en, sizeof(double) * NT * NT);
    m->tab.dens = m->tab_store;
    m->tab.pres = m->tab_store + NT;
    m->tab.sound = m->tab_store + 2 * NT;
    m->tab.en = m->tab_store + 2 * NT + NT * NT;
}

void lsc_oracle_tabeos_c2p_flat(const double *dens, const double *pres, const double *sound, const double *en,
                                const double *consv, double *prim) {
    eos_table_t tb = {dens, pres, sound, en};
    lsc_oracle_tabeos_c2p(&tb, consv, prim);
}

void *lsc_oracle_euler_create(const lsc_oracle_euler_cfg *c, const double *u0) {
    euler_t *m = (euler_t *)calloc(1, sizeof(euler_t));
    m->c = *c;
    m->ncell = (size_t)(c->nx + 4) * (size_t)(c->ny + 4);
    m->dx = (c->x1 - c->x0) / c->nx;
    m->dy = (c->y1 - c->y0) / c->ny;
    double **arr[] = {&m->u, &m->A, &m->B, &m->flux, &m->ubar, &m->rL, &m->rR, &m->hL, &m->hR};
    for (size_t a = 0; a < sizeof(arr) / sizeof(arr[0]); ++a) *arr[a] = (double *)calloc(m->ncell * NE, sizeof(double));
    memcpy(m->u, u0, m->ncell * NE * sizeof(double));
    return m;
}

void lsc_oracle_euler_step(void *h, long nsteps, double *steplog) {
    euler_t *m = (euler_t *)h;
    for (long s = 0; s < nsteps; ++s) {
        euler_step(m);
        if (steplog) { steplog[2 * s] = m->t; steplog[2 * s + 1] = m->dt; }
    }
}

void lsc_oracle_euler_get(void *h, double *u, double *t, double *dt) {
    euler_t *m = (euler_t *)h;
    if (u) memcpy(u, m->u, m->ncell * NE * sizeof(double));
    if (t) *t = m->t;
    if (dt) *dt = m->dt;
}

void lsc_oracle_euler_destroy(void *h) {
    euler_t *m = (euler_t *)h;
    double *arr[] = {m->u, m->A, m->B, m->flux, m->ubar, m->rL, m->rR, m->hL, m->hR};
    for (size_t a = 0; a < sizeof(arr) / sizeof(arr[0]); ++a) free(arr[a]);
    free(m->tab_store);
    free(m);
}
