// mhd_physics.cuh -- per-cell / per-interface device arithmetic of the 2-D ideal-MHD
// MUSCL-Hancock + HLLC + mixed-GLM update.
//
// Bit-faithful to the reference's expressions (MHD_codes/MUSCL_NEW.cpp, cited per
// function): same association, IEEE division and sqrt, no FMA contraction (the
// translation unit is compiled with -fmad=false).  What differs from the reference is
// only *how often* things are evaluated: the reference recomputes the same pure
// sub-expressions up to 14 times per interface; here each
// distinct value is computed once, and of the two star states only the one the
// HLLC branch selects.  Every common sub-expression that is reused is an IDENTICAL
// IEEE expression on identical inputs, so results are bit-identical.
//
// Exact scalings.  Multiplying by 0.5, 0.25 or 4 is exact in binary floating point (no rounding unless the
// result is subnormal), so fl(0.5*a + 0.5*b) = 0.5*fl(a+b), fl(fl(0.5*x)*y) = 0.5*fl(x*y) and
// fl(c - fl(0.5*w)) = fma(-0.5, w, c): where the reference multiplies by such a constant right before an
// addition, the pair is issued as ONE explicit __fma_rn with the constant as multiplier (scaled_add below).
// That is not a contraction of the reference's arithmetic -- the product is exact, the single rounding is the
// reference's rounding of the sum -- and it removes ~55 of ~815 fp64 instructions per cell and sweep.  The
// identity needs the scaled intermediate to be normal or an exact zero.  That is guarded, not assumed: the scaled
// operand of every such fma whose size the state does not bound from below enters the fast division's window
// accumulator (FastDiv::note_scaled -- a non-zero value below 2^-500 flags the block), and flagged blocks are
// recomputed by the IeeeDiv instantiation, for which every one of these sites is the reference's LITERAL
// expression (literal_scalings / sadd in fp64_div.cuh), two roundings and all.
// tests/test_mhd_gpu.py::test_subnormal_scalings_match_the_reference_literally holds states with u0 == 0 next to
// subnormal xi*Delta.
#pragma once
#include <cuda_runtime.h>

#include "fp64_div.cuh"
#include "mhd_kernels.cuh"

// Square roots of the HLLC wave speeds: the policy's branch-free guarded sequence (fp64_div.cuh: FastDiv::sqrt) in
// the x sweep, plain sqrt() in the y sweep.  A/B on one B200, OT 8192^2, same state hash: x 4.35 -> 4.28 ms with
// it, y 4.61 -> 4.62 ms (the y sweep sits at 127 registers; the extra live accumulator updates cost it what the
// removed branches gain) -- so each sweep gets what is faster for it.  MHD_GUARDED_SQRT: 0 never, 1 x only, 2 both.
#ifndef MHD_HLL_BRANCHFREE
#define MHD_HLL_BRANCHFREE 0
#endif
#ifndef MHD_GUARDED_SQRT
#define MHD_GUARDED_SQRT 1
#endif
#define MHD_SQRT(DIR, dv, x) (((MHD_GUARDED_SQRT == 2) || (MHD_GUARDED_SQRT == 1 && (DIR) == 0)) ? (dv).sqrt(x) : sqrt(x))

namespace lsc {

// fl(c + k*w) for k = +-0.5, +-0.25, +-4: the exact product k*w and the addition's single rounding
__device__ __forceinline__ double scaled_add(double k, double w, double c) { return __fma_rn(k, w, c); }

// std::min / std::max as libstdc++ defines them (first argument on ties / unordered)
__device__ __forceinline__ double std_min(double a, double b) { return (b < a) ? b : a; }
__device__ __forceinline__ double std_max(double a, double b) { return (a < b) ? b : a; }

// Non-trivial primitives of a conserved state: velocities and gas pressure.
// MUSCL_NEW.cpp:22-34 (ConservativeToPrimitive); v0,v5..v8 are copies of u.
struct Prim {
    double vx, vy, vz, p;
    double kw;  // rho*(vx*vx+vy*vy+vz*vz): twice the kinetic term 0.5*rho*(...) of the pressure expression
    double vv;  // vx*vx+vy*vy+vz*vz (only the literal IeeeDiv forms read it)
    Rcp irho;  // shared refined reciprocal of the density (fp64_div.cuh)
};

template <class D>
__device__ __forceinline__ Prim c2p(D& dv, const double (&u)[NV], double gm1) {
    Prim q;
    q.irho = dv.rcp(u[0]);
    q.vx = dv.div(u[1], q.irho);
    q.vy = dv.div(u[2], q.irho);
    q.vz = dv.div(u[3], q.irho);
    // v4 = (u4 - 0.5*v0*(v.v) - 0.5*(B.B))*(gamma-1), MUSCL_NEW.cpp:29-31: 0.5*v0*(v.v) = 0.5*fl(v0*(v.v)) exactly
    q.vv = q.vx * q.vx + q.vy * q.vy + q.vz * q.vz;
    q.kw = u[0] * q.vv;
    if (literal_scalings<D>::value) {  // the reference's expression as written (MUSCL_NEW.cpp:29-31)
        q.p = (u[4] - 0.5 * u[0] * q.vv - 0.5 * (u[5] * u[5] + u[6] * u[6] + u[7] * u[7])) * gm1;
    } else {
        dv.note_scaled_aux(u[4]);  // the two halved terms are subtracted from the energy: exact while it is not tiny
        q.p = scaled_add(-0.5, u[5] * u[5] + u[6] * u[6] + u[7] * u[7], scaled_add(-0.5, q.kw, u[4])) * gm1;
    }
    return q;
}

// Full primitive vector (for the limiter ratio, which differences primitives).
template <class D>
__device__ __forceinline__ void c2p_full(D& dv, const double (&u)[NV], double gm1, double (&v)[NV]) {
    Prim q = c2p(dv, u, gm1);
    v[0] = u[0]; v[1] = q.vx; v[2] = q.vy; v[3] = q.vz; v[4] = q.p;
    v[5] = u[5]; v[6] = u[6]; v[7] = u[7]; v[8] = u[8];
}

// Physical flux of a conserved state.  DIR 0: FluxDefX MUSCL_NEW.cpp:36-49,
// DIR 1: FluxDefY MUSCL_NEW.cpp:51-64.  ch2 = ch*ch.
template <int DIR, class D>
__device__ __forceinline__ void flux_from_prim(const double (&u)[NV], const Prim& q, double ch2, double (&f)[NV]) {
    const double B2 = u[5] * u[5] + u[6] * u[6] + u[7] * u[7];
    const double vB = q.vx * u[5] + q.vy * u[6] + q.vz * u[7];
    if (DIR == 0) {
        const double rvx = u[0] * q.vx;
        f[0] = u[1];
        f[1] = sadd<D>(0.5, B2, rvx * q.vx + q.p) - u[5] * u[5];
        f[2] = rvx * q.vy - u[5] * u[6];
        f[3] = rvx * q.vz - u[5] * u[7];
        f[4] = sadd<D>(0.5, B2, u[4] + q.p) * q.vx - vB * u[5];
        f[5] = u[8];
        f[6] = u[6] * q.vx - u[5] * q.vy;
        f[7] = u[7] * q.vx - u[5] * q.vz;
        f[8] = ch2 * u[5];
    } else {
        const double rvx = u[0] * q.vx;
        const double rvy = u[0] * q.vy;
        f[0] = u[2];
        f[1] = rvx * q.vy - u[5] * u[6];
        f[2] = sadd<D>(0.5, B2, rvy * q.vy + q.p) - u[6] * u[6];
        f[3] = rvy * q.vz - u[6] * u[7];
        f[4] = sadd<D>(0.5, B2, u[4] + q.p) * q.vy - vB * u[6];
        f[5] = u[5] * q.vy - u[6] * q.vx;
        f[6] = u[8];
        f[7] = u[7] * q.vy - u[6] * q.vz;
        f[8] = ch2 * u[6];
    }
}

template <int DIR, class D>
__device__ __forceinline__ void flux_def(D& dv, const double (&u)[NV], double gm1, double ch2, double (&f)[NV]) {
    const Prim q = c2p(dv, u, gm1);
    flux_from_prim<DIR, D>(u, q, ch2, f);
}

// Slope-limited linear reconstruction of one component.  MUSCL_NEW.cpp:845-859
// (x) and :909-923 (y).  (pm,p0,pp) are the PRIMITIVE values the limiter ratio is
// formed from, (um,u0,up) the CONSERVED values that are reconstructed.
//
// xi = fmin(2r/(1+r), 2/(1+r)) is evaluated with ONE division: both quotients share
// the positive denominator d = fl(1+r), 2.0*r is exact, and IEEE division is
// monotone in the numerator, so the minimum is fl(2r/d) when r <= 1 and fl(2/d)
// otherwise (equal at r == 1; for r = +inf the reference's fmin(NaN, 0) = 0 = 2/inf;
// NaN r gives NaN both ways).
// Split in two so that the sweeps can form the nine limiter ratios back to back (nine independent Newton chains
// the scheduler can interleave) before the branchy second quotient of each component.
template <class D>
__device__ __forceinline__ double limiter_ratio(D& dv, double pm, double p0, double pp) {
    return dv.div(p0 - pm, (pp - p0) + 1e-8);
}

// The same ratio for the components that are identically zero in planar problems (v_z, B_z: OT, KH and every
// other committed initial condition): a zero numerator over a non-zero, non-NaN denominator gives r = +-0, which
// the limiter maps to xi = 0 whatever its sign, so the reciprocal refinement and the quotient (8 fp64
// instructions + guards) are skipped behind a branch that is warp-uniform in such problems.  Any other operand
// pair (incl. 0/0 and 0/NaN, which must stay NaN) takes the division.
template <class D>
__device__ __forceinline__ double limiter_ratio_zskip(D& dv, double pm, double p0, double pp) {
    const double num = p0 - pm;
    const double den = (pp - p0) + 1e-8;
#ifndef MHD_NO_ZSKIP
    if (num == 0.0 && (den < 0.0 || den > 0.0)) return 0.0;
#endif
    return dv.div(num, den);
}

// limiter ratio of component k as the sweeps form it
// (k is a compile-time constant after unrolling)
template <class D>
__device__ __forceinline__ double limiter_ratio_k(D& dv, int k, double pm, double p0, double pp) {
    return (k == 3 || k == 7) ? limiter_ratio_zskip(dv, pm, p0, pp) : limiter_ratio(dv, pm, p0, pp);
}

#ifndef MHD_BRANCHFREE_XI
#define MHD_BRANCHFREE_XI 1
#endif
template <class D, bool BRANCHFREE = false>
__device__ __forceinline__ void reconstruct_with_ratio(D& dv, double r, double um, double u0, double up,
                                                       double& lo, double& hi) {
    const double dPlus = up - u0;
    const double dMinus = u0 - um;
    double xi;
#if MHD_BRANCHFREE_XI
    if (BRANCHFREE) {
        // no branch around the second quotient: r <= 0 is replaced by r = +0, for which 2r/(1+r) = 0/1 = +0 = the
        // reference's xi (NaN r stays NaN: NaN <= 0 is false).  The nine quotients of a cell then are nine
        // independent Newton chains in one straight-line region instead of nine branch regions.
        const double rs = (r <= 0) ? 0.0 : r;
        xi = dv.div_pos0((rs <= 1.0) ? 2.0 * rs : 2.0, 1 + rs);
    } else
#endif
    if (r <= 0) {
        xi = 0;
    } else {
        xi = dv.div_pos((r <= 1.0) ? 2.0 * r : 2.0, 1 + r);  // r > 0 (or NaN): 0 < dividend <= 2, divisor >= 1
    }
    // Delta = 0.5*dMinus + 0.5*dPlus = 0.5*S exactly; 0.5*xi*Delta = 0.25*fl(xi*S) exactly (see the header)
    if (literal_scalings<D>::value) {  // MUSCL_NEW.cpp:856-859 as written
        const double Delta = 0.5 * dMinus + 0.5 * dPlus;
        lo = u0 - 0.5 * xi * Delta;
        hi = u0 + 0.5 * xi * Delta;
        return;
    }
    const double S = dMinus + dPlus;
    const double T = xi * S;
    dv.note_scaled(T);
    lo = scaled_add(-0.25, T, u0);
    hi = scaled_add(0.25, T, u0);
}

template <class D>
__device__ __forceinline__ void reconstruct1(D& dv, double pm, double p0, double pp, double um, double u0,
                                             double up, double& lo, double& hi) {
    reconstruct_with_ratio(dv, limiter_ratio(dv, pm, p0, pp), um, u0, up, lo, hi);
}

// MUSCL-Hancock predictor: both face states advanced by half a step with the flux
// difference across the cell.  MUSCL_NEW.cpp:865-874 (x), :929-938 (y).
// hdtdh = 0.5*(dt/dx) (or dy).  In-place on lo/hi.
template <int DIR, class D>
__device__ __forceinline__ void predictor(D& dv, double (&lo)[NV], double (&hi)[NV], double gm1, double ch2,
                                          double hdtdh) {
    double fl[NV], fh[NV];
    flux_def<DIR>(dv, lo, gm1, ch2, fl);
    flux_def<DIR>(dv, hi, gm1, ch2, fh);
#pragma unroll
    for (int k = 0; k < NV; ++k) {
        const double d = hdtdh * (fh[k] - fl[k]);
        lo[k] = lo[k] - d;
        hi[k] = hi[k] - d;
    }
}

// fast magnetosonic speed squared along DIR.  speedsx/speedsy MUSCL_NEW.cpp:140-165,
// :197-222 -- evaluated on the ORIGINAL state (the reference ignores its B?Tilde argument).
template <int DIR, class D>
__device__ __forceinline__ double fast_speed_sq(D& dv, const Rcp& irho, double p, double B2, double Bn, double gamma) {
    const double cs2 = dv.div(gamma * p, irho);
    const double t = cs2 + dv.div(B2, irho);
    // 4.0*cs2*Bn*Bn/rho = 4*fl(fl(fl(cs2*Bn)*Bn)/rho) exactly
    if (literal_scalings<D>::value)  // MUSCL_NEW.cpp:155: t*t - 4.0*cs2*Bn*Bn/rho as written
        return 0.5 * (t + sqrt(t * t - dv.div(4.0 * cs2 * Bn * Bn, irho)));
    return 0.5 * (t + MHD_SQRT(DIR, dv, scaled_add(-4.0, dv.div(cs2 * Bn * Bn, irho), t * t)));
}

// Second half of hllc_flux for the side the HLLC branch selects (LEFT: flux built from the left state, regions
// sL >= 0 and sL <= 0 <= qStar; otherwise from the right state): flux of the modified state, and in the two star
// regions (`star`) the HLL tangential fields, the star state uStarX/uStarY and F = F(S1) + s (U* - S1).
template <int DIR, bool LEFT, class D>
__device__ __forceinline__ void hllc_side(D& dv, const double (&L)[NV], const double (&R)[NV], const Prim& qL,
                                          const Prim& qR, double sL, double sR, double qS, double BnT, double psiT,
                                          bool star, double gm1, double ch2, double (&F)[NV]) {
    constexpr int N = (DIR == 0) ? 5 : 6;  // normal field component
    const double (&S)[NV] = LEFT ? L : R;
    const Prim& qs = LEFT ? qL : qR;
    const double ss = LEFT ? sL : sR;
    const double vn = (DIR == 0) ? qs.vx : qs.vy;
    // modified state S1: S with S1[N]=BnT, S1[8]=psiT
    double S1[NV];
#pragma unroll
    for (int k = 0; k < NV; ++k) S1[k] = S[k];
    S1[N] = BnT;
    S1[8] = psiT;

    // flux of the modified state: FluxDefX/Y(x1) -- velocities unchanged, pressure re-derived
    // with the cleaned normal field (:497-498 / :531-532)
    double FS[NV];
    {
        Prim q1 = qs;  // S1[0..4] == S[0..4]: same velocities and kinetic term, new magnetic pressure
        if (literal_scalings<D>::value)
            q1.p = (S1[4] - 0.5 * S1[0] * qs.vv - 0.5 * (S1[5] * S1[5] + S1[6] * S1[6] + S1[7] * S1[7])) * gm1;
        else
            q1.p = scaled_add(-0.5, S1[5] * S1[5] + S1[6] * S1[6] + S1[7] * S1[7], scaled_add(-0.5, qs.kw, S1[4])) * gm1;
        flux_from_prim<DIR, D>(S1, q1, ch2, FS);
    }

    if (!star) {
#pragma unroll
        for (int k = 0; k < NV; ++k) F[k] = FS[k];
        return;
    }

    // ---- HLL average of the two tangential field components (uHLLx/y; only these are consumed)
    constexpr int T1 = (DIR == 0) ? 6 : 5;  // first tangential field component
    double fT1L, fT1R, f7L, f7R;              // physical flux components T1 and 7 of the ORIGINAL states
    if (DIR == 0) {
        fT1L = L[6] * qL.vx - L[5] * qL.vy;  fT1R = R[6] * qR.vx - R[5] * qR.vy;
        f7L = L[7] * qL.vx - L[5] * qL.vz;   f7R = R[7] * qR.vx - R[5] * qR.vz;
    } else {
        fT1L = L[5] * qL.vy - L[6] * qL.vx;  fT1R = R[5] * qR.vy - R[6] * qR.vx;
        f7L = L[7] * qL.vy - L[6] * qL.vz;   f7R = R[7] * qR.vy - R[6] * qR.vz;
    }
    const Rcp iw = dv.rcp(sR - sL);  // shared by the three quotients over the fan width
    double BtH, BzH;
#if MHD_HLL_BRANCHFREE
    {
        // the two HLL quotients formed unconditionally (subsonic fans take them anyway), the outer cases selected
        const double qt = dv.div(sR * R[T1] - sL * L[T1] - fT1R + fT1L, iw);
        const double qz = dv.div(sR * R[7] - sL * L[7] - f7R + f7L, iw);
        const bool useL = (sL >= 0), useM = !useL && (sL < 0 && sR > 0);
        BtH = useL ? L[T1] : (useM ? qt : R[T1]);
        BzH = useL ? L[7] : (useM ? qz : R[7]);
    }
#else
    if (sL >= 0) {
        BtH = L[T1]; BzH = L[7];
    } else if (sL < 0 && sR > 0) {
        BtH = dv.div(sR * R[T1] - sL * L[T1] - fT1R + fT1L, iw);
        BzH = dv.div(sR * R[7] - sL * L[7] - f7R + f7L, iw);
    } else {
        BtH = R[T1]; BzH = R[7];
    }
#endif

    // ---- star state on the selected side (uStarX / uStarY), normal field = BnT on both sides
    const double BnS = dv.div(sR * BnT - sL * BnT, iw);
    const double Bt = S[T1], Bz = S[7];
    const double vt = (DIR == 0) ? qs.vy : qs.vx;  // tangential in-plane velocity
    const double vz = qs.vz;
    // B2 with the cleaned normal component, summed in x,y,z order (:335-336 / :420-421)
    const double B2s = (DIR == 0) ? (BnT * BnT + Bt * Bt + Bz * Bz) : (Bt * Bt + BnT * BnT + Bz * Bz);
    const double pT = sadd<D>(0.5, B2s, qs.p);
    const double pStar = S[0] * (ss - vn) * (qS - vn) + pT - BnT * BnT + BnS * BnS;
    const Rcp is = dv.rcp(ss - qS);  // shared by the four quotients over (s - qStar)
    const double rhoS = dv.div(S[0] * (ss - vn), is);
    const double mnS = rhoS * qS;
    // tangential momenta (:353-357 / :437-442)
    const double mtS = vt * rhoS - dv.div(BnS * BtH - BnT * Bt, is);
    double mzS;
    if (DIR == 1 && !LEFT) {
        // MUSCL_NEW.cpp:442: momzStarR = rhoStarL*u_zL - (ByStar*BzHLL - ByL*BzL)/(sL - qStar)
        const Rcp isl = dv.rcp(sL - qS);
        const double rhoSL = dv.div(L[0] * (sL - qL.vy), isl);
        mzS = rhoSL * qL.vz - dv.div(BnS * BzH - BnT * L[7], isl);
    } else {
        mzS = vz * rhoS - dv.div(BnS * BzH - BnT * Bz, is);
    }
    // v.B of the side state (x,y,z order) and of the star state
    double vB, vBH;
    if (DIR == 0) {
        vB = BnT * vn + Bt * vt + Bz * vz;
        vBH = dv.div(BnS * mnS + BtH * mtS + BzH * mzS, rhoS);
    } else {
        vB = Bt * vt + BnT * vn + Bz * vz;
        vBH = dv.div(BtH * mtS + BnS * mnS + BzH * mzS, rhoS);
    }
    const double ES = S[4] * dv.div(rhoS, qs.irho) +
                      dv.div((pStar * qS - pT * vn) - (BnS * vBH - BnT * vB), is);

    double US[NV];
    US[0] = rhoS;
    US[(DIR == 0) ? 1 : 2] = mnS;
    US[(DIR == 0) ? 2 : 1] = mtS;
    US[3] = mzS;
    US[4] = ES;
    US[N] = BnS;
    US[T1] = BtH;
    US[7] = BzH;
    US[8] = psiT;
#pragma unroll
    for (int k = 0; k < NV; ++k) F[k] = FS[k] + ss * (US[k] - S1[k]);
}

// HLLC flux with mixed-GLM interface values.  FluxHLLCX MUSCL_NEW.cpp:483-515 /
// FluxHLLCY :517-549, together with everything they call:
//   psiAndBx/psiAndBy :121-138, wavespeedx/y :167-253, uHLLx/y :255-307,
//   uStarX :309-395, uStarY :397-481 (incl. the :442 quirk: the RIGHT star
//   z-momentum in y is built from the LEFT quantities).
// L is the state left/below the interface, R the one right/above.
template <int DIR, class D>
__device__ __forceinline__ void hllc_flux(D& dv, const double (&L)[NV], const double (&R)[NV], double gamma,
                                          double gm1, double ch, double ch2, double inv2ch, double halfch,
                                          double (&F)[NV]) {
    constexpr int N = (DIR == 0) ? 5 : 6;  // normal field component
    // ---- mixed GLM interface values (:126-127 / :135-136)
    // inv2ch = 1.0/(2.0*ch), halfch = ch/2.0: per-step scalars, divided once on the host
    const double sBn = L[N] + R[N], sPsi = L[8] + R[8];
    dv.note_scaled_aux(sBn);
    dv.note_scaled_aux(sPsi);
    const double BnT = sadd<D>(0.5, sBn, -(inv2ch * (R[8] - L[8])));
    const double psiT = sadd<D>(0.5, sPsi, -(halfch * (R[N] - L[N])));

    // ---- primitives of the ORIGINAL states
    const Prim qL = c2p(dv, L, gm1), qR = c2p(dv, R, gm1);
    const double vnL = (DIR == 0) ? qL.vx : qL.vy;
    const double vnR = (DIR == 0) ? qR.vx : qR.vy;
    const double B2L = L[5] * L[5] + L[6] * L[6] + L[7] * L[7];
    const double B2R = R[5] * R[5] + R[6] * R[6] + R[7] * R[7];

    // ---- wave speeds (:178-192 / :235-250)
    const double cfL = MHD_SQRT(DIR, dv, fast_speed_sq<DIR>(dv, qL.irho, qL.p, B2L, L[N], gamma));
    const double cfR = MHD_SQRT(DIR, dv, fast_speed_sq<DIR>(dv, qR.irho, qR.p, B2R, R[N], gamma));
    const double cmax = std_max(cfL, cfR);
    const double sL = std_min(vnL, vnR) - cmax;
    const double sR = std_max(vnL, vnR) + cmax;
    const double pModL = sadd<D>(0.5, B2L, qL.p) - L[N] * L[N];
    const double pModR = sadd<D>(0.5, B2R, qR.p) - R[N] * R[N];
    const double qS = dv.div(R[0] * vnR * (sR - vnR) - L[0] * vnL * (sL - vnL) + pModL - pModR,
                           R[0] * (sR - vnR) - L[0] * (sL - vnL));

    // ---- which of the four HLLC regions (:502-512 / :536-546), evaluated in the reference's order
    const bool regA = (sL >= 0);
    const bool regB = !regA && (sL <= 0 && qS >= 0);
    const bool regC = !regA && !regB && (qS <= 0 && sR >= 0);
    const bool left = regA || regB;  // flux built from the left (modified) state

    // ---- the side the branch selects.  Two instantiations of the remainder behind ONE branch instead of ~20
    // selected doubles (state, primitives, reciprocal, speed): the branch is warp-uniform except where the
    // contact speed changes sign, and each instantiation reads its side's registers directly.
    if (left)
        hllc_side<DIR, true>(dv, L, R, qL, qR, sL, sR, qS, BnT, psiT, regB || regC, gm1, ch2, F);
    else
        hllc_side<DIR, false>(dv, L, R, qL, qR, sL, sR, qS, BnT, psiT, regB || regC, gm1, ch2, F);
}

// Local signal speed for the CFL condition.  ComputeTimeStep MUSCL_NEW.cpp:75-109.
template <class D>
__device__ __forceinline__ double cfl_speed(D& dv, const double (&q)[NV], double gamma) {
    const double rho = q[0], Bx = q[5], By = q[6], Bz = q[7];
    const Rcp irho = dv.rcp(rho);
    const double ux = dv.div(q[1], irho), uy = dv.div(q[2], irho), uz = dv.div(q[3], irho);
    const double B2 = Bx * Bx + By * By + Bz * Bz;
    dv.note_scaled_aux(q[4]);
    if (literal_scalings<D>::value) {  // MUSCL_NEW.cpp:85-101 as written
        const double ke = 0.5 * rho * (ux * ux + uy * uy + uz * uz) + 0.5 * (Bx * Bx + By * By + Bz * Bz);
        const double pl = (gamma - 1.0) * (q[4] - ke);
        const double al = sqrt(dv.div(gamma * pl, irho));
        const double sl = al * al + dv.div(B2, irho);
        const double cx = sqrt(0.5 * (sl + sqrt(sl * sl - dv.div(4 * al * al * Bx * Bx, irho))));
        const double cy = sqrt(0.5 * (sl + sqrt(sl * sl - dv.div(4 * al * al * By * By, irho))));
        const double cz = sqrt(0.5 * (sl + sqrt(sl * sl - dv.div(4 * al * al * Bz * Bz, irho))));
        return std_max(std_max(fabs(ux) + cx, fabs(uy) + cy), fabs(uz) + cz);
    }
    // ke = 0.5*rho*(u.u) + 0.5*B2 = 0.5*fl(rho*(u.u) + B2) exactly
    const double p = (gamma - 1.0) * scaled_add(-0.5, rho * (ux * ux + uy * uy + uz * uz) + B2, q[4]);
    const double a = sqrt(dv.div(gamma * p, irho));
    const double a2 = a * a;
    const double s = a2 + dv.div(B2, irho);
    const double s2 = s * s;  // pow(s, 2)
    // 4*a*a*Bd*Bd/rho = 4*fl(fl(fl(a2*Bd)*Bd)/rho) exactly (4*a*a = 4*a2)
    const double cfx = sqrt(0.5 * (s + sqrt(scaled_add(-4.0, dv.div(a2 * Bx * Bx, irho), s2))));
    const double cfy = sqrt(0.5 * (s + sqrt(scaled_add(-4.0, dv.div(a2 * By * By, irho), s2))));
    const double cfz = sqrt(0.5 * (s + sqrt(scaled_add(-4.0, dv.div(a2 * Bz * Bz, irho), s2))));
    return std_max(std_max(fabs(ux) + cfx, fabs(uy) + cfy), fabs(uz) + cfz);
}

// Single-precision UPPER BOUND of cfl_speed(q) / R for the y sweep's warp filter: a cell with bound < 1 has an
// exact signal speed below the running exact maximum R and cannot raise it, so its warp may skip the fp64
// evaluation.  Everything is formed in units of R (inv_r = 1/R rounded up), which makes the estimate invariant
// under the scalings of the state (rho, m, E -> c*rho, c*m, c*E; B -> sqrt(c)*B; any velocity unit): fp32
// underflow can then only lose contributions that are < 1e-13 of R.  Conservative by construction:
//   * the pressure, the one cancellation-prone term, gets +1e-6*(|E| + kinetic + magnetic) before use;
//   * the inner radicand s^2 - 4 a^2 b_d^2 (cancels when a ~ b with B along d) gets +2e-6*s^2;
//   * the total is inflated by 2^-10, three orders above the accumulated fp32 rounding (~20 ops * 6e-8) and
//     the error of the approximate rcp/sqrt (2^-22);
//   * the scale factors 1/(rho R), 1/(rho R^2) must lie in [1e-25, 1e25], otherwise the result is NaN.
// NaN/inf anywhere (negative pressure, fp32 overflow, R = 0) propagates to the result; the caller evaluates
// exactly unless `bound < 1` holds.  tests/test_mhd_gpu.py::test_cfl_upper_bound checks bound*R >= exact.
__device__ __forceinline__ float cfl_speed_over_max_f32(const double (&q)[NV], float gamma_up, float gm1_up,
                                                        float inv_r) {
    const float rho = (float)q[0];
    float ir;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(ir) : "f"(rho));
    const float kv = ir * inv_r;  // momentum -> velocity / R
    const float k = kv * inv_r;   // energy density -> speed^2 / R^2
    const float vx = fabsf((float)q[1]) * kv, vy = fabsf((float)q[2]) * kv, vz = fabsf((float)q[3]) * kv;
    const float e = (float)q[4] * k;
    const float b1 = (float)q[5], b2 = (float)q[6], b3 = (float)q[7];
    const float bx2 = (b1 * k) * b1, by2 = (b2 * k) * b2, bz2 = (b3 * k) * b3;
    const float B2 = bx2 + by2 + bz2;
    const float ke = 0.5f * ((vx * vx + vy * vy) + vz * vz);
    const float mag = 0.5f * B2;
    const float pk = gm1_up * ((e - ke - mag) + 1e-6f * (fabsf(e) + ke + mag));  // p / (rho R^2), upper bound
    const float a2 = gamma_up * pk;
    const float s = a2 + B2;
    const float s2 = s * s;
    const float pad = 2e-6f * s2;
    const float fa = 4.0f * a2;
    float rx, ry, rz, cx, cy, cz;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(rx) : "f"(fmaxf(s2 - fa * bx2, 0.0f) + pad));
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(ry) : "f"(fmaxf(s2 - fa * by2, 0.0f) + pad));
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(rz) : "f"(fmaxf(s2 - fa * bz2, 0.0f) + pad));
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(cx) : "f"(0.5f * (s + rx)));
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(cy) : "f"(0.5f * (s + ry)));
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(cz) : "f"(0.5f * (s + rz)));
    const float sx = vx + cx, sy = vy + cy, sz = vz + cz;
    float best = fmax3_nan(sx, sy, sz) * 1.0009765625f;  // NaN-propagating maximum
    const bool scales_ok = fminf(k, kv) > 1e-25f && fmaxf(k, kv) < 1e25f;
    return scales_ok ? best : __int_as_float(0x7fc00000);
}

}  // namespace lsc
