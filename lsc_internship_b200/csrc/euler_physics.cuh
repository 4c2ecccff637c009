// euler_physics.cuh -- per-cell / per-interface device arithmetic of the 2-D Euler SLIC scheme
// (Euler_codes/SLIC_euler_2d.cpp = S; transmissive variant generaliseSLIC.cpp = G).  Bit-faithful: the
// reference's association, IEEE division (fp64_div.cuh), no FMA contraction.
#pragma once
#include <cuda_runtime.h>

#include "fp64_div.cuh"

namespace lsc {

constexpr int NE = 4;  // rho, rho*vx, rho*vy, E  (S:10-17)

// fluxX_def S:30-44 (DIR 0) / fluxY_def S:46-60 (DIR 1).  The reference evaluates this 16 times per cell
// on identical inputs (S:292-293); here once per distinct state.
template <int DIR, class D>
__device__ __forceinline__ void euler_flux(D& dv, const double (&x)[NE], double gm1, double (&f)[NE]) {
    const double rho = x[0];
    const Rcp ir = dv.rcp(x[0]);
    const double vx = dv.div(x[1], ir);
    const double vy = dv.div(x[2], ir);
    const double KE = 0.5 * rho * (vx * vx + vy * vy);
    const double p = gm1 * (x[3] - KE);
    if (DIR == 0) {
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

// Limiter + linear reconstruction of one conserved component, S:262-283.  cm = 0.5*(1+omega),
// cp = 0.5*(1-omega) (host scalars).  xi_R = 2/(1+r) is only consumed when r > 1 (or r is NaN), so the
// division is issued only then; fmin keeps the reference's NaN behaviour.
template <class D>
__device__ __forceinline__ void euler_reconstruct1(D& dv, double um, double u0, double up, double cm, double cp,
                                                   double& lo, double& hi) {
    const double dPlus = up - u0;
    const double dMinus = u0 - um;
    const double r = dv.div(dMinus, dPlus + 1e-8);
    const double Delta = cm * dMinus + cp * dPlus;
    double xi;
    if (r <= 0) {
        xi = 0;
    } else if (r <= 1) {
        xi = r;
    } else {
        xi = fmin(1.0, dv.div(2.0, 1 + r));
    }
    lo = u0 - 0.5 * xi * Delta;
    hi = u0 + 0.5 * xi * Delta;
}

// Reconstruct cell (um,u0,up) and advance both face states by half a step, S:260-296.  hdtdh = 0.5*(dt/h).
template <int DIR, class D>
__device__ __forceinline__ void euler_halfstates(D& dv, const double (&um)[NE], const double (&u0)[NE],
                                                 const double (&up)[NE], double gm1, double cm, double cp,
                                                 double hdtdh, double (&lo)[NE], double (&hi)[NE]) {
#pragma unroll
    for (int k = 0; k < NE; ++k) euler_reconstruct1(dv, um[k], u0[k], up[k], cm, cp, lo[k], hi[k]);
    double fl[NE], fh[NE];
    euler_flux<DIR>(dv, lo, gm1, fl);
    euler_flux<DIR>(dv, hi, gm1, fh);
#pragma unroll
    for (int k = 0; k < NE; ++k) {
        const double d = hdtdh * (fh[k] - fl[k]);
        lo[k] = lo[k] - d;
        hi[k] = hi[k] - d;
    }
}

// FORCE flux, getFluxX S:111-136 / getFluxY S:139-163.  hdtdh = 0.5*(dt/h), hhdt = 0.5*h/dt.
template <int DIR, class D>
__device__ __forceinline__ void euler_force(D& dv, const double (&a)[NE], const double (&b)[NE], double gm1,
                                            double hdtdh, double hhdt, double (&F)[NE]) {
    double f1[NE], f2[NE], uh[NE], ri[NE];
    euler_flux<DIR>(dv, a, gm1, f1);
    euler_flux<DIR>(dv, b, gm1, f2);
#pragma unroll
    for (int k = 0; k < NE; ++k) uh[k] = 0.5 * (a[k] + b[k]) - hdtdh * (f2[k] - f1[k]);
    euler_flux<DIR>(dv, uh, gm1, ri);
#pragma unroll
    for (int k = 0; k < NE; ++k) {
        const double lf = hhdt * (a[k] - b[k]) + 0.5 * (f1[k] + f2[k]);
        F[k] = 0.5 * (lf + ri[k]);
    }
}

// Signal speed of ComputeTimeStep, S:175-184.
template <class D>
__device__ __forceinline__ double euler_cfl_speed(D& dv, const double (&q)[NE], double gamma) {
    const double rho = q[0];
    const Rcp ir = dv.rcp(rho);
    const double vx = dv.div(q[1], ir);
    const double vy = dv.div(q[2], ir);
    const double speed = sqrt(vx * vx + vy * vy);
    const double KE = 0.5 * rho * speed * speed;
    const double p = (gamma - 1.0) * (q[3] - KE);
    const double a = sqrt(dv.div(gamma * p, ir));
    return speed + a;
}

}  // namespace lsc
