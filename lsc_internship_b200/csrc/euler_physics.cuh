// euler_physics.cuh -- per-cell / per-interface device arithmetic of the 2-D Euler SLIC scheme
// (Euler_codes/SLIC_euler_2d.cpp = S; transmissive variant generaliseSLIC.cpp = G; tabulated-EoS variant
// MHD_codes/TabEoS.cpp = T).  Bit-faithful: the
// reference's association, IEEE division (fp64_div.cuh), no FMA contraction.  Multiplications by 0.5 are exact,
// so where the reference adds such a product the pair is issued as one explicit FMA (escaled_add; the argument,
// its guard for tiny operands and the literal forms of the IEEE fix instantiation are those in the header of
// mhd_physics.cuh).
#pragma once
#ifndef EULER_BRANCHFREE_XI
#define EULER_BRANCHFREE_XI 0
#endif
#include <cuda_runtime.h>

#include "fp64_div.cuh"

namespace lsc {

constexpr int NE = 4;  // rho, rho*vx, rho*vy, E  (S:10-17)

// fl(c + k*w) for k = +-0.5: the exact product and the addition's single rounding
__device__ __forceinline__ double escaled_add(double k, double w, double c) { return __fma_rn(k, w, c); }

// Guards on the one-fma scalings other than the limiter product xi*Delta (which is always guarded:
// euler_reconstruct1).  Off by default, because for this scheme they are redundant: a fused form can differ from the
// reference's two roundings only if the scaled operand w is a non-zero value below 2^-1021 AND the addend is below
// ~2^-960 as well.  (a) E - 0.5*rho*(v.v): w that small means |rho*v| < 2^-510*sqrt(rho), and every momentum
// component of every state the pressure is taken of has just been a DIVIDEND of v = m/rho in the same policy object
// -- a non-zero dividend below 2^-500 flags the block.  (b) 0.5*(a+b) and 0.5*(f1+f2) in the FORCE flux: density and
// energy sums are of normal size; a momentum (flux) sum or the difference a-b in the addend that small needs both
// face momenta below 2^-900 and one of them non-zero, and both are dividends in euler_flux(a) / euler_flux(b).
// What is left are states with rho > 2^20 together with a total energy below 2^-960, which no run reaches.
// A/B on one B200, cylinder 4096^2: 1.152 ms/step without, 1.181 with (profiles/r2/ab_scaled_guards_level1.log).
#ifndef EULER_AUX_SCALED_GUARDS
#define EULER_AUX_SCALED_GUARDS 0
#endif
template <class D>
__device__ __forceinline__ void euler_note_aux(D& dv, double w) {
#if EULER_AUX_SCALED_GUARDS
    dv.note_scaled_aux(w);
#endif
}

// ---- equation of state policies ---------------------------------------------------------------------------
// The sweeps are templated on the EoS: IdealEos is S / G (p = (gamma-1)(E - KE)), TableEos is
// MHD_codes/TabEoS.cpp = T (bilinear lookup in a 500x500 table, T:147-404).  Both take the state, the shared
// reciprocal of rho and the velocities the flux function has already formed.

struct IdealEos {
    double gm1, gamma;
    static constexpr bool kHasCflBound = true;
    __device__ __forceinline__ void load_hints(const int*) const {}
    __device__ __forceinline__ void store_hints(int*) const {}

    // Single-precision UPPER BOUND of cfl_speed(q) / R for the y sweep's warp filter (the construction and its
    // safety argument are those of cfl_speed_over_max_f32 in mhd_physics.cuh): everything in units of the running
    // maximum R (inv_r = 1/R rounded up) so that the estimate is invariant under rescaling of the state; the
    // pressure -- the one cancellation-prone term -- gets +1e-6*(|E| + kinetic) before use; the total is inflated
    // by 2^-10; scale factors outside [1e-25, 1e25] and any NaN/inf (negative pressure, R = 0) give NaN, and the
    // caller evaluates exactly unless `bound < 1`.  tests/test_euler_gpu.py::test_euler_cfl_upper_bound.
    __device__ __forceinline__ float cfl_speed_over_max_f32(const double (&q)[NE], float inv_r) const {
        const float gamma_up = __double2float_ru(gamma), gm1_up = __double2float_ru(gamma - 1.0);
        const float rho = (float)q[0];
        float ir;
        asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(ir) : "f"(rho));
        const float kv = ir * inv_r;  // momentum -> velocity / R
        const float k = kv * inv_r;   // energy density -> speed^2 / R^2
        const float vx = fabsf((float)q[1]) * kv, vy = fabsf((float)q[2]) * kv;
        const float v2 = vx * vx + vy * vy;
        const float e = (float)q[3] * k;
        const float ke = 0.5f * v2;
        const float pk = gm1_up * ((e - ke) + 1e-6f * (fabsf(e) + ke));  // p / (rho R^2), upper bound
        float sp, a;
        asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(sp) : "f"(v2));
        asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(a) : "f"(gamma_up * pk));
        const float best = (sp + a) * 1.0009765625f;
        const bool scales_ok = fminf(k, kv) > 1e-25f && fmaxf(k, kv) < 1e25f;
        return scales_ok ? best : __int_as_float(0x7fc00000);
    }

    template <class D>
    __device__ __forceinline__ double pressure(D& dv, const double (&x)[NE], const Rcp&, double vx, double vy) const {
        // KE = 0.5*rho*(vx*vx+vy*vy) = 0.5*fl(rho*(...)) exactly; p = (gamma-1)*(E - KE), S:37-38
        if (literal_scalings<D>::value) {  // S:35-38 as written
            const double KE = 0.5 * x[0] * (vx * vx + vy * vy);
            return gm1 * (x[3] - KE);
        }
        euler_note_aux(dv, x[3]);  // exact while the energy the halved term is taken from is not tiny
        return gm1 * escaled_add(-0.5, x[0] * (vx * vx + vy * vy), x[3]);
    }

    // Signal speed of ComputeTimeStep, S:175-184.
    template <class D>
    __device__ __forceinline__ double cfl_speed(D& dv, const double (&q)[NE]) const {
        const double rho = q[0];
        const Rcp ir = dv.rcp(rho);
        const double vx = dv.div(q[1], ir);
        const double vy = dv.div(q[2], ir);
        const double speed = sqrt(vx * vx + vy * vy);
        // KE = 0.5*rho*speed*speed = 0.5*fl(fl(rho*speed)*speed) exactly
        euler_note_aux(dv, q[3]);
        const double p = literal_scalings<D>::value
                             ? (gamma - 1.0) * (q[3] - 0.5 * rho * speed * speed)  // S:179-180 as written
                             : (gamma - 1.0) * escaled_add(-0.5, rho * speed * speed, q[3]);
        const double a = sqrt(dv.div(gamma * p, ir));
        return speed + a;
    }
};

constexpr int NT = 500;  // table axes, T:12

struct TableEos {
    static constexpr bool kHasCflBound = false;  // every cell evaluates its table lookup
    __device__ __forceinline__ float cfl_speed_over_max_f32(const double (&)[NE], float) const {
        return __int_as_float(0x7fc00000);
    }
    const double* __restrict__ dens;   // [500] ascending densities (file line 10)
    const double* __restrict__ pres;   // [500] ascending pressures (file line 12)
    const double* __restrict__ sound;  // [500][500] sound speed, [pressure index][density index]
    const double* __restrict__ en;     // [500][500] specific internal energy, same indexing
    // Host-verified (eos_table_is_monotone, euler_host.cpp): both axes strictly ascending and every energy column
    // ascending by more than the rounding of the interpolation.  Then each search predicate of T is monotone in
    // the index and the bracket its bisection ends in is unique -- (last index whose entry is below the value,
    // that index + 1) -- so it can be FOUND any way: here by checking the bracket of the thread's previous lookup
    // first (consecutive lookups are the face states of one cell) and bisecting only on a miss.  With
    // warm == 0 every search is the reference's bisection.
    int warm;
    mutable int hint_d = 249, hint_e = 249, hint_p = 249;  // previous brackets (density axis, energy column, pressure axis)

    // The y sweep is one thread per cell: its first lookup would have no previous bracket.  Each block therefore
    // starts from the brackets one of its threads ended with in the previous step (3 ints per block, written
    // at the end of the kernel; any values in range are valid starting points, a stale hint only costs a miss).
    __device__ __forceinline__ void load_hints(const int* h) const {
        if (!warm || !h) return;
        hint_d = min(max(__ldg(h), 0), 498);
        hint_e = min(max(__ldg(h + 1), 0), 498);
        hint_p = min(max(__ldg(h + 2), 0), 499);
    }
    __device__ __forceinline__ void store_hints(int* h) const {
        if (!warm || !h) return;
        h[0] = hint_d;
        h[1] = hint_e;
        h[2] = hint_p;
    }

    // The bisection every lookup of T starts with (T:257-275 etc.): brackets v in the ascending axis; point2
    // starts at 499, or at 500 in SoundSpeed's pressure search (T:366; index 500 is never read).  Returns the
    // unclamped point1 (point2 = point1 + 1 on exit).
    __device__ __forceinline__ static int bisect_axis(const double* __restrict__ a, double v, int point2) {
        int point1 = 0;
        while (point2 - point1 > 1) {
            const int half = (point1 + point2) >> 1;
            if (v > __ldg(a + half)) point1 = half;
            else point2 = half;
        }
        return point1;
    }
    // lo / hi: the clamped bracket (T:271-275); a_lo / a_hi: the axis values there
    __device__ __forceinline__ void bracket(const double* __restrict__ a, double v, int point2, int& hint, int& lo,
                                            int& hi, double& a_lo, double& a_hi) const {
        int p1 = hint;
        bool hit = false;
        if (warm) {
            // p1 is the answer iff (p1 == 0 or v > a[p1]) and (p1 + 1 == point2 or not v > a[p1 + 1]); the two
            // entries read for the test are the bracket's axis values
            a_lo = __ldg(a + p1);
            a_hi = __ldg(a + min(p1 + 1, NT - 1));
            hit = (p1 == 0 || v > a_lo) && (p1 + 1 >= point2 || !(v > a_hi));
        }
        if (!hit) {
            p1 = bisect_axis(a, v, point2);
            a_lo = __ldg(a + p1);
            a_hi = __ldg(a + min(p1 + 1, NT - 1));
        }
        hint = p1;
        lo = p1;  // 0 <= p1 <= 499 and p1 + 1 >= 1: only the upper clamp of T:271-275 can act
        hi = min(p1 + 1, NT - 1);
    }

    // density-interpolated energy at pressure index h (the bisection's comparand, T:293-296)
    __device__ __forceinline__ double energy_at(int h, int db, int dt, double density_ratio) const {
        const double eb = __ldg(en + h * NT + db), et = __ldg(en + h * NT + dt);
        return eb + density_ratio * (et - eb);
    }

    // ConservativeToPrimative's pressure, T:263-324.  Also returns the density bracket for SoundSpeed.
    template <class D>
    __device__ __forceinline__ double pressure_ex(D& dv, const double (&x)[NE], const Rcp& ir, double vx, double vy,
                                                  int& db, int& dt, double& density_ratio) const {
        const double rho = x[0];
        double d_bottom, d_top;
        bracket(dens, rho, 499, hint_d, db, dt, d_bottom, d_top);
        density_ratio = dv.div(rho - d_bottom, d_top - d_bottom);
        if (d_bottom == d_top) density_ratio = 1;
        euler_note_aux(dv, x[3]);
        const double e = literal_scalings<D>::value
                             ? dv.div(x[3] - 0.5 * rho * (vx * vx + vy * vy), ir)  // T:288 as written
                             : dv.div(escaled_add(-0.5, rho * (vx * vx + vy * vy), x[3]), ir);
        // T:291-299: bisection on the density-interpolated energy column (monotone for 0 <= ratio <= 1 when the
        // table passed the host check; outside that range -- extrapolation -- always the reference's bisection).
        // The four entries the warm test reads are the corners of the interpolation cell.
        int point1 = hint_e;
        bool hit = false;
        double f_SW, f_SE, f_NW, f_NE;
        if (warm && density_ratio >= 0.0 && density_ratio <= 1.0) {
            f_SW = __ldg(en + point1 * NT + db);
            f_SE = __ldg(en + point1 * NT + dt);
            f_NW = __ldg(en + (point1 + 1) * NT + db);  // hint_e <= 498
            f_NE = __ldg(en + (point1 + 1) * NT + dt);
            hit = (point1 == 0 || e > (f_SW + density_ratio * (f_SE - f_SW))) &&
                  (point1 + 1 >= 499 || !(e > (f_NW + density_ratio * (f_NE - f_NW))));
        }
        if (!hit) {
            point1 = 0;
            int point2 = 499;
            while (point2 - point1 > 1) {
                const int half = (point1 + point2) >> 1;
                if (e > energy_at(half, db, dt, density_ratio)) point1 = half;
                else point2 = half;
            }
            f_SW = __ldg(en + point1 * NT + db);
            f_SE = __ldg(en + point1 * NT + dt);
            f_NW = __ldg(en + (point1 + 1) * NT + db);
            f_NE = __ldg(en + (point1 + 1) * NT + dt);
        }
        hint_e = point1;
        const int pt = point1 + 1, pb = point1;  // captured before the (no-op) clamps, T:303-310
        // ReverseBilinearInterpolation, T:156-163
        const double f_x_y1 = (1 - density_ratio) * f_SW + density_ratio * f_SE;
        const double f_x_y2 = (1 - density_ratio) * f_NW + density_ratio * f_NE;
        double y = dv.div(e - f_x_y1, f_x_y2 - f_x_y1);
        if (f_x_y2 == f_x_y1) y = 1;
        const double p_bottom = __ldg(pres + pb);
        return y * (__ldg(pres + pt) - p_bottom) + p_bottom;  // T:322
    }

    template <class D>
    __device__ __forceinline__ double pressure(D& dv, const double (&x)[NE], const Rcp& ir, double vx, double vy) const {
        int db, dt;
        double ratio;
        return pressure_ex(dv, x, ir, vx, vy, db, dt, ratio);
    }

    // ComputeTimeStep's speed, T:505-521: |v| + SoundSpeed(u) (T:330-404; its ConservativeToPrimative and density
    // bracket repeat the ones above on identical inputs)
    template <class D>
    __device__ __forceinline__ double cfl_speed(D& dv, const double (&q)[NE]) const {
        const Rcp ir = dv.rcp(q[0]);
        const double vx = dv.div(q[1], ir);
        const double vy = dv.div(q[2], ir);
        const double speed = sqrt(vx * vx + vy * vy);
        int db, dt, pb, pt;
        double density_ratio;
        const double p = pressure_ex(dv, q, ir, vx, vy, db, dt, density_ratio);
        double p_bottom, p_top;
        bracket(pres, p, 500, hint_p, pb, pt, p_bottom, p_top);  // T:366
        double pressure_ratio = dv.div(p - p_bottom, p_top - p_bottom);
        if (p_bottom == p_top) pressure_ratio = 1;
        const double ss_NE = __ldg(sound + pt * NT + dt), ss_SE = __ldg(sound + pb * NT + dt);
        const double ss_SW = __ldg(sound + pb * NT + db), ss_NW = __ldg(sound + pt * NT + db);
        // BilinearInterpolation, T:147-153
        const double f_x_y1 = (1 - density_ratio) * ss_SW + density_ratio * ss_SE;
        const double f_x_y2 = (1 - density_ratio) * ss_NW + density_ratio * ss_NE;
        const double sound_speed = (1 - pressure_ratio) * f_x_y1 + pressure_ratio * f_x_y2;
        return speed + sound_speed;
    }
};

// fluxX_def S:30-44 / T:408-422 (DIR 0), fluxY_def S:46-60 / T:424-438 (DIR 1).  The reference evaluates this 16
// times per cell on identical inputs (S:292-293); here once per distinct state.
template <int DIR, class D, class EOS>
__device__ __forceinline__ void euler_flux(D& dv, const double (&x)[NE], const EOS& eos, double (&f)[NE]) {
    const double rho = x[0];
    const Rcp ir = dv.rcp(x[0]);
    const double vx = dv.div(x[1], ir);
    const double vy = dv.div(x[2], ir);
    const double p = eos.pressure(dv, x, ir, vx, vy);
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
#if EULER_BRANCHFREE_XI
    {
        // no branch around the quotient: it is formed for every cell, with r replaced by 1 where the reference does
        // not look at it (r <= 1; a NaN r lands in the last case either way and yields fmin(1, NaN) = 1 = fmin(1, 2/2))
        const double rs = (r > 1) ? r : 1.0;
        const double q = fmin(1.0, dv.div_pos(2.0, 1 + rs));
        xi = (r <= 0) ? 0.0 : ((r <= 1) ? r : q);
    }
#else
    if (r <= 0) {
        xi = 0;
    } else if (r <= 1) {
        xi = r;
    } else {
        xi = fmin(1.0, dv.div_pos(2.0, 1 + r));  // r > 1 (or NaN): dividend 2, divisor > 2
    }
#endif
    if (literal_scalings<D>::value) {  // S:281-282 as written
        lo = u0 - 0.5 * xi * Delta;
        hi = u0 + 0.5 * xi * Delta;
        return;
    }
    const double T = xi * Delta;  // 0.5*xi*Delta = 0.5*fl(xi*Delta) exactly
    dv.note_scaled(T);
    lo = escaled_add(-0.5, T, u0);
    hi = escaled_add(0.5, T, u0);
}

// Reconstruct cell (um,u0,up) and advance both face states by half a step, S:260-296.  hdtdh = 0.5*(dt/h).
template <int DIR, class D, class EOS>
__device__ __forceinline__ void euler_halfstates(D& dv, const double (&um)[NE], const double (&u0)[NE],
                                                 const double (&up)[NE], const EOS& eos, double cm, double cp,
                                                 double hdtdh, double (&lo)[NE], double (&hi)[NE]) {
#pragma unroll
    for (int k = 0; k < NE; ++k) euler_reconstruct1(dv, um[k], u0[k], up[k], cm, cp, lo[k], hi[k]);
    double fl[NE], fh[NE];
    euler_flux<DIR>(dv, lo, eos, fl);
    euler_flux<DIR>(dv, hi, eos, fh);
#pragma unroll
    for (int k = 0; k < NE; ++k) {
        const double d = hdtdh * (fh[k] - fl[k]);
        lo[k] = lo[k] - d;
        hi[k] = hi[k] - d;
    }
}

// FORCE flux, getFluxX S:111-136 / getFluxY S:139-163.  hdtdh = 0.5*(dt/h), hhdt = 0.5*h/dt.
template <int DIR, class D, class EOS>
__device__ __forceinline__ void euler_force(D& dv, const double (&a)[NE], const double (&b)[NE], const EOS& eos,
                                            double hdtdh, double hhdt, double (&F)[NE]) {
    double f1[NE], f2[NE], uh[NE], ri[NE];
    euler_flux<DIR>(dv, a, eos, f1);
    euler_flux<DIR>(dv, b, eos, f2);
#pragma unroll
    for (int k = 0; k < NE; ++k) {
        const double w = a[k] + b[k];
        euler_note_aux(dv, w);
        uh[k] = sadd<D>(0.5, w, -(hdtdh * (f2[k] - f1[k])));
    }
    euler_flux<DIR>(dv, uh, eos, ri);
#pragma unroll
    for (int k = 0; k < NE; ++k) {
        const double w = f1[k] + f2[k];
        euler_note_aux(dv, w);
        const double lf = sadd<D>(0.5, w, hhdt * (a[k] - b[k]));
        F[k] = 0.5 * (lf + ri[k]);
    }
}

}  // namespace lsc
