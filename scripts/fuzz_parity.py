"""Fuzz of the bit parity CUDA == oracle (TEST INFRASTRUCTURE, run on the GPU box: it imports tests/oracle_util).
Random states in regimes the fixtures do not visit -- piecewise-constant states with jumps (every HLLC region, every
limiter branch), states scaled by powers of two towards the edges of the fast division's operand window (the IEEE
fix path with its literal expressions), fields that vanish in patches, all boundary-condition pairs, ragged sizes.
`slabs`: the MHD cases through a single-process group of 2-4 uneven slabs; `params`: other domains, Courant numbers,
adiabatic indices and limiter weights; `tabeos`: the table equation of state.
`reuse`: contexts uploaded again and again; `long`: 19 steps in one call; `ic`: the initial-condition generators.
usage: fuzz_parity.py [mhd|euler|both|slabs|params|tabeos|reuse|long|ic|run] [first_seed] [count]
prints one line per case, exits 1 on a mismatch"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import lsc_internship_b200 as lsc  # noqa: E402
from oracle_util import OracleEuler, OracleMhd, euler_cfg, mhd_cfg  # noqa: E402

MHD_BCS = [("periodic", "periodic"), ("periodic", "reflective"), ("transmissive", "transmissive"),
           ("transmissive", "reflective"), ("periodic", "transmissive")]


SCHEDULE = (1, 2, 3)   # steps per call; `long` mode: one call of 19 steps (CUDA graphs of 8 + single steps)
REUSE = None           # `reuse` mode: contexts kept per shape / parameters and uploaded again and again


def bits(a):
    return np.ascontiguousarray(a).view(np.uint64)


def solver(key, make):
    """a fresh context, or -- in `reuse` mode -- the one an earlier case with the same key left behind (a new upload
    must wipe everything: the step count, u^{n-1} of the MHD limiter, the exp table, the fix lists)"""
    if REUSE is None:
        return make()
    if key not in REUSE:
        REUSE[key] = make()
        REUSE[key].close = lambda: None
    return REUSE[key]


def blocky(rng, shape, lo, hi, nblocks):
    """piecewise-constant random field: nblocks x nblocks patches with values in [lo, hi), plus small noise"""
    bx = np.sort(rng.choice(np.arange(1, shape[0]), size=min(nblocks, shape[0] - 1), replace=False))
    by = np.sort(rng.choice(np.arange(1, shape[1]), size=min(nblocks, shape[1] - 1), replace=False))
    ix = np.searchsorted(bx, np.arange(shape[0]), side="right")
    iy = np.searchsorted(by, np.arange(shape[1]), side="right")
    vals = rng.uniform(lo, hi, (len(bx) + 1, len(by) + 1))
    return vals[ix][:, iy]


def sizes(rng, seed, small):
    """mostly small ragged grids; every 16th seed one that spans several strips / tiles / blocks"""
    if seed % 16 == 7:
        return int(rng.integers(100, 520)), int(rng.integers(100, 520))
    return int(rng.integers(4, small)), int(rng.integers(4, small))


def slab_bounds(rng, nx):
    """2-4 uneven slabs of at least 4 columns each"""
    n = int(rng.integers(2, 5))
    while n > 1 and nx < 4 * n:
        n -= 1
    cuts = np.sort(rng.choice(np.arange(1, nx // 4), size=n - 1, replace=False)) * 4 if n > 1 else np.array([], int)
    edges = [0] + [int(c) for c in cuts] + [nx]
    return [(edges[k], edges[k + 1] - edges[k]) for k in range(n)]


def mhd_case(seed, slabs=False, params=False):
    rng = np.random.default_rng(seed)
    nx, ny = sizes(rng, seed, 90)
    bcx, bcy = MHD_BCS[seed % len(MHD_BCS)]
    regime = ["jumps", "smooth", "scaled", "patchy", "tiny"][(seed // len(MHD_BCS)) % 5]
    shape = (nx + 4, ny + 4)
    nb = int(rng.integers(1, 6))
    if regime == "smooth":
        x = np.linspace(0, 2 * np.pi, shape[0])[:, None]
        y = np.linspace(0, 2 * np.pi, shape[1])[None, :]
        rho = 1.5 + 0.5 * np.sin(x + rng.uniform(0, 6)) * np.cos(y)
        v = [0.7 * np.sin(y + k) * np.cos(x * (k + 1)) for k in range(3)]
        B = [0.6 * np.cos(x + 2 * k) * np.sin(y * (k + 1)) for k in range(3)]
        p = 1.0 + 0.3 * np.cos(x) * np.cos(y)
        psi = 0.05 * np.sin(x) * np.sin(y)
    else:
        rho = blocky(rng, shape, 0.3, 3.0, nb)
        v = [blocky(rng, shape, -1.0, 1.0, nb) for _ in range(3)]
        B = [blocky(rng, shape, -1.0, 1.0, nb) for _ in range(3)]
        p = blocky(rng, shape, 0.5, 3.0, nb)
        psi = blocky(rng, shape, -0.1, 0.1, nb)
        noise = 1e-3 if regime != "patchy" else 0.0
        rho = rho * (1 + noise * rng.uniform(-1, 1, shape))
    if regime == "tiny":         # out-of-plane fields around 1e-160 (limiter products subnormal) with exact zeros
        k = int(rng.integers(-575, -525))
        for f in (v[2], B[2], psi):
            f *= 2.0 ** k
            f[blocky(rng, shape, 0.0, 1.0, nb) < 0.3] = 0.0
    if regime == "patchy":       # fields that vanish exactly in patches, signed zeros included
        for f in (v[2], B[2], psi, v[0]):
            m = blocky(rng, shape, 0.0, 1.0, nb) < 0.5
            f[m] = 0.0
        B[1][blocky(rng, shape, 0.0, 1.0, nb) < 0.3] = -0.0
    g = 5.0 / 3.0
    u = np.zeros(shape + (9,))
    u[..., 0] = rho
    for k in range(3):
        u[..., 1 + k] = rho * v[k]
        u[..., 5 + k] = B[k]
    u[..., 4] = p / (g - 1) + 0.5 * rho * (v[0] ** 2 + v[1] ** 2 + v[2] ** 2) + 0.5 * (B[0] ** 2 + B[1] ** 2 + B[2] ** 2)
    u[..., 8] = psi
    if regime == "scaled":       # exact power-of-two scaling: densities by 4^k, fields by 2^k (velocities unchanged)
        k = int(rng.choice([-240, -200, -120, 120, 200, 230]))
        u[..., 0:5] = np.ldexp(u[..., 0:5], 2 * k)
        u[..., 5:9] = np.ldexp(u[..., 5:9], k)
    dom, cfl = (0.0, 1.0, 0.0, 1.0), 0.8
    if params:                   # other domains (dx != dy), Courant numbers and adiabatic indices than the committed ones
        prng = np.random.default_rng(777_000 + seed)
        dom = (float(prng.uniform(-2, 0)), float(prng.uniform(0.5, 3)), float(prng.uniform(-1, 0)), float(prng.uniform(0.2, 5)))
        cfl = float(prng.choice([0.3, 0.5, 0.8, 0.9]))
        gnew = float(prng.choice([5.0 / 3.0, 1.4, 2.0]))
        u[..., 4] = (u[..., 4] - 0.5 * (u[..., 1] ** 2 + u[..., 2] ** 2 + u[..., 3] ** 2) / u[..., 0]
                     - 0.5 * (u[..., 5] ** 2 + u[..., 6] ** 2 + u[..., 7] ** 2)) * (g - 1) / (gnew - 1) \
            + 0.5 * (u[..., 1] ** 2 + u[..., 2] ** 2 + u[..., 3] ** 2) / u[..., 0] \
            + 0.5 * (u[..., 5] ** 2 + u[..., 6] ** 2 + u[..., 7] ** 2)
        g = gnew
    cfg = mhd_cfg(nx, ny, dom=dom, gamma=g, cfl=cfl, bcx=bcx, bcy=bcy)
    o = OracleMhd(cfg, u)
    kw = dict(nxCells=nx, nyCells=ny, x0=dom[0], x1=dom[1], y0=dom[2], y1=dom[3], C=cfl, gamma=g, bc_x=bcx, bc_y=bcy,
              problem="ot")
    if slabs and nx >= 8:
        bounds = slab_bounds(np.random.default_rng(555_000 + seed), nx)
        s = lsc.MusclNewGroup.from_parts([lsc.MusclNew(rank=r, nranks=len(bounds), slab=b, **kw)
                                          for r, b in enumerate(bounds)])
        regime += f"/{len(bounds)}sl"
    else:
        s = solver(("mhd",) + tuple(sorted(kw.items())), lambda: lsc.MusclNew(**kw))
    s.upload(u)
    ok, steps = True, 0
    for n in SCHEDULE:
        lg = o.step(n)
        if not np.isfinite(lg).all() or not (lg[:, 2] > 0).all():
            break                # the reference itself has lost the state (NaN speed): nothing to compare
        s.step(n)
        steps += n
        ur, tr = o.state()[:2]
        if not (np.array_equal(bits(s.download()), bits(ur)) and s.scalars()[0] == tr):
            ok = False
            break
    ur = o.state()[0]
    finite = bool(np.isfinite(ur[2:-2, 2:-2]).all())
    redo = s.ieee_fallback_count()
    s.close()
    o.close()
    print(f"mhd   seed {seed:5d} {nx:3d}x{ny:<3d} {bcx[:5]}/{bcy[:5]} {regime:11s} steps {steps} finite {finite} "
          f"ieee-redo {redo:6d}  {'ok' if ok else 'MISMATCH'}", flush=True)
    return ok


def euler_case(seed, params=False, slabs=False):
    rng = np.random.default_rng(10_000 + seed)
    nx, ny = sizes(rng, seed, 100)
    bc = ["periodic", "transmissive"][seed % 2]
    regime = ["jumps", "scaled", "patchy", "front", "tiny"][(seed // 2) % 5]
    shape = (nx + 4, ny + 4)
    nb = int(rng.integers(1, 6))
    rho = blocky(rng, shape, 0.3, 3.0, nb)
    vx, vy = blocky(rng, shape, -1.0, 1.0, nb), blocky(rng, shape, -1.0, 1.0, nb)
    p = blocky(rng, shape, 0.5, 3.0, nb)
    if regime == "patchy":
        vx[blocky(rng, shape, 0.0, 1.0, nb) < 0.5] = 0.0
        vy[blocky(rng, shape, 0.0, 1.0, nb) < 0.5] = -0.0
    if regime == "tiny":         # one or both velocity components around 1e-160, exact zeros in patches
        k = int(rng.integers(-575, -525))
        vy = vy * 2.0 ** k
        if seed % 3 == 0:
            vx = vx * 2.0 ** k
        vy[blocky(rng, shape, 0.0, 1.0, nb) < 0.3] = 0.0
    if regime == "front":        # momenta decaying through 1e-100 ... 1e-320 across the grid, as at an explosion front
        decay = 10.0 ** (-np.linspace(0, 330, shape[0]))[:, None]
        vx, vy = vx * decay, vy * decay
    g, dom, cfl, omega = 1.4, (0.0, 1.0, 0.0, 1.0), 0.8, 0.0
    if params:                   # the limiter's omega (S:272), other domains, Courant numbers, adiabatic indices
        prng = np.random.default_rng(888_000 + seed)
        dom = (float(prng.uniform(-2, 0)), float(prng.uniform(0.5, 3)), float(prng.uniform(-1, 0)), float(prng.uniform(0.2, 5)))
        cfl = float(prng.choice([0.3, 0.5, 0.8, 0.9]))
        g = float(prng.choice([1.4, 5.0 / 3.0, 1.1]))
        omega = float(prng.choice([0.0, 0.5, -0.5, 0.3, 1.0]))
    u = np.stack([rho, rho * vx, rho * vy, p / (g - 1) + 0.5 * rho * (vx * vx + vy * vy)], axis=-1)
    if regime == "scaled":
        k = int(rng.choice([-230, -200, -120, 120, 200, 230]))
        u = np.ldexp(u, 2 * k)
    cfg = euler_cfg(nx, ny, dom, gamma=g, cfl=cfl, omega=omega, bc=bc)
    o = OracleEuler(cfg, u)
    kw = dict(nxCells=nx, nyCells=ny, x0=dom[0], x1=dom[1], y0=dom[2], y1=dom[3], C=cfl, gamma=g, omega=omega, bc=bc)
    nsl = int(np.random.default_rng(666_000 + seed).integers(2, 5))
    if slabs and nx >= 4 * nsl:
        s = lsc.SlicEuler2dGroup(nsl, **kw)
        s.ieee_fallback_count = lambda: sum(q.ieee_fallback_count() for q in s.parts)
        regime += f"/{nsl}sl"
    else:
        s = solver(("euler",) + tuple(sorted(kw.items())), lambda: lsc.SlicEuler2d(**kw))
    s.upload(u)
    ok, steps = True, 0
    for n in SCHEDULE:
        lg = o.step(n)
        if not np.isfinite(lg).all() or not (lg[:, 1] > 0).all():
            break
        s.step(n)
        steps += n
        ur, tr = o.state()[:2]
        if not (np.array_equal(bits(s.download()), bits(ur)) and s.scalars()[0] == tr):
            ok = False
            break
    ur = o.state()[0]
    finite = bool(np.isfinite(ur[2:-2, 2:-2]).all())
    redo = s.ieee_fallback_count()
    s.close()
    o.close()
    print(f"euler seed {seed:5d} {nx:3d}x{ny:<3d} {bc[:5]:5s}       {regime:11s} steps {steps} finite {finite} "
          f"ieee-redo {redo:6d}  {'ok' if ok else 'MISMATCH'}", flush=True)
    return ok


def tabeos_case(seed):
    """TabEoS.cpp's table equation of state (synthetic Plasma19-format table): the program's four-quadrant state with
    patchwise random density / energy factors and momenta, warm-started and plain searches alike"""
    from oracle_util import parsed_eos_table, tabeos_ic
    rng = np.random.default_rng(20_000 + seed)
    nx, ny = int(rng.integers(4, 70)), int(rng.integers(4, 70))
    shape = (nx + 4, ny + 4)
    nb = int(rng.integers(1, 5))
    table = parsed_eos_table()
    cfg = euler_cfg(nx, ny, (0.0, 1.0, 0.0, 1.0), bc="transmissive")
    u = tabeos_ic(cfg, table)
    u[..., 0] *= blocky(rng, shape, 0.6, 1.6, nb)
    u[..., 3] *= blocky(rng, shape, 0.7, 1.5, nb)
    c = np.sqrt(u[..., 3] / u[..., 0])
    u[..., 1] = u[..., 0] * c * blocky(rng, shape, -0.3, 0.3, nb)
    u[..., 2] = u[..., 0] * c * blocky(rng, shape, -0.3, 0.3, nb)
    if seed % 3 == 1:            # operands the fast division has to hand to the IEEE fix kernels
        u[..., 1][blocky(rng, shape, 0.0, 1.0, nb) < 0.4] = -0.0
        u[..., 2] = np.where(blocky(rng, shape, 0.0, 1.0, nb) < 0.5, u[..., 2] * 2.0 ** -550, u[..., 2])
        u[..., 2][blocky(rng, shape, 0.0, 1.0, nb) < 0.2] = 0.0
    o = OracleEuler(cfg, u)
    o.set_table(*table)
    s = lsc.SlicEuler2d(nx, ny, bc="transmissive", problem="tabeos")
    s.set_eos_table(*table)
    if seed % 4 == 3:
        s.set_eos_search(False)
    s.upload(u)
    ok, steps = True, 0
    for n in (1, 2):
        o.step(n)
        s.step(n)
        steps += n
        ur, tr = o.state()[:2]
        if not np.isfinite(ur).all():
            break
        if not (np.array_equal(bits(s.download()), bits(ur)) and s.scalars()[0] == tr):
            ok = False
            break
    finite = bool(np.isfinite(ur).all())
    redo = s.ieee_fallback_count()
    s.close()
    o.close()
    print(f"tabeos seed {seed:5d} {nx:3d}x{ny:<3d} trans       {'table':11s} steps {steps} finite {finite} "
          f"ieee-redo {redo:6d}  {'ok' if ok else 'MISMATCH'}", flush=True)
    return ok


def run_case(seed):
    """The whole loop M:831-1003 through run() -- dump cadence (M:975), stop test, pause / resume of the device-resident
    loop, psi-factor table -- on a smooth random state with a random stop time (5..50 steps, up to three dumps),
    alternately on one context and on a group of uneven slabs: step count, t and state against the oracle's loop."""
    rng = np.random.default_rng(40_000 + seed)
    nx, ny = int(rng.integers(12, 60)), int(rng.integers(12, 60))
    bcx, bcy = MHD_BCS[seed % len(MHD_BCS)]
    shape = (nx + 4, ny + 4)
    x = np.linspace(0, 2 * np.pi, shape[0])[:, None]
    y = np.linspace(0, 2 * np.pi, shape[1])[None, :]
    ph = rng.uniform(0, 6, 8)
    rho = 1.5 + 0.4 * np.sin(x + ph[0]) * np.cos(y + ph[1])
    v = [0.5 * np.sin(y + ph[2 + k]) * np.cos(x * (k + 1)) for k in range(3)]
    B = [0.5 * np.cos(x + ph[5 + k]) * np.sin(y * (k + 1)) for k in range(3)]
    p = 1.0 + 0.3 * np.cos(x) * np.cos(y)
    g = 5.0 / 3.0
    u = np.zeros(shape + (9,))
    u[..., 0] = rho
    for k in range(3):
        u[..., 1 + k] = rho * v[k]
        u[..., 5 + k] = B[k]
    u[..., 4] = p / (g - 1) + 0.5 * rho * (v[0] ** 2 + v[1] ** 2 + v[2] ** 2) + 0.5 * (B[0] ** 2 + B[1] ** 2 + B[2] ** 2)
    cfl = float(rng.choice([0.4, 0.8, 0.9]))
    t_stop = float(rng.uniform(0.04, 0.33))
    cfg = mhd_cfg(nx, ny, cfl=cfl, bcx=bcx, bcy=bcy)
    o = OracleMhd(cfg, u)
    n_ref = o.run_until(t_stop)
    ur, tr = o.state()[:2]
    kw = dict(nxCells=nx, nyCells=ny, x0=0.0, x1=1.0, y0=0.0, y1=1.0, C=cfl, bc_x=bcx, bc_y=bcy, problem="ot")
    if seed % 2 and nx >= 8:
        bounds = slab_bounds(np.random.default_rng(555_000 + seed), nx)
        s = lsc.MusclNewGroup.from_parts([lsc.MusclNew(rank=r, nranks=len(bounds), slab=b, **kw)
                                          for r, b in enumerate(bounds)])
        how = f"{len(bounds)} slabs"
    else:
        s = lsc.MusclNew(**kw)
        how = "1 ctx  "
    s.upload(u)
    dumps = []
    import tempfile
    with tempfile.TemporaryDirectory() as td:
        n = s.run(tStop=t_stop, outdir=td)
        dumps = sorted(os.listdir(td))
    ok = bool(n == n_ref and s.scalars()[0] == tr and np.array_equal(bits(s.download()), bits(ur)))
    # the reference writes MHD<k>.dat when t crosses 0.1 k, plus the final dump
    want = sum(1 for k in range(0, 5) if 0.1 * k < tr) + 1
    ok &= len(dumps) == want
    s.close()
    o.close()
    print(f"run   seed {seed:5d} {nx:3d}x{ny:<3d} {bcx[:5]}/{bcy[:5]} {how} t_stop {t_stop:.4f} steps {n} (ref {n_ref}) "
          f"files {len(dumps)} (want {want})  {'ok' if ok else 'MISMATCH'}", flush=True)
    return ok


def ic_case(seed):
    """the library's initial-condition generators (host libm, MUSCL_NEW.cpp:719-825, S:552-596) against the oracle's on
    random grids and domains -- no GPU work"""
    from oracle_util import euler_ic, mhd_ic
    rng = np.random.default_rng(30_000 + seed)
    nx, ny = int(rng.integers(4, 200)), int(rng.integers(4, 200))
    dom = (float(rng.uniform(-2, 0)), float(rng.uniform(0.5, 3)), float(rng.uniform(-2, 0)), float(rng.uniform(0.2, 5)))
    if seed % 2:
        dom = (0.0, 1.0, 0.0, 1.0) if seed % 4 == 1 else (0.0, 1.0, -1.0, 1.0)
    ok = True
    for prob in ("ot", "kh"):
        cfg = mhd_cfg(nx, ny, dom=dom)
        s = lsc.MusclNew(nx, ny, *dom, problem=prob)
        ok &= np.array_equal(bits(s.initial_condition()), bits(mhd_ic(cfg, prob)))
        s.close()
    for prob in ("ot", "quadrants", "cylinder"):
        cfg = euler_cfg(nx, ny, dom, bc="transmissive" if prob != "ot" else "periodic")
        s = lsc.SlicEuler2d(nx, ny, *dom, bc="transmissive" if prob != "ot" else "periodic", problem=prob)
        ok &= np.array_equal(bits(s.initial_condition()), bits(euler_ic(cfg, prob)))
        s.close()
    print(f"ic    seed {seed:5d} {nx:3d}x{ny:<3d} domain {dom}  {'ok' if ok else 'MISMATCH'}", flush=True)
    return ok


if __name__ == "__main__":
    which = sys.argv[1] if len(sys.argv) > 1 else "both"
    first = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    count = int(sys.argv[3]) if len(sys.argv) > 3 else 40
    bad = 0
    for seed in range(first, first + count):
        if which in ("mhd", "both"):
            bad += not mhd_case(seed)
        if which in ("euler", "both"):
            bad += not euler_case(seed)
        if which == "slabs":
            bad += not mhd_case(seed, slabs=True)
            bad += not euler_case(seed, slabs=True)
        if which == "params":
            bad += not mhd_case(seed, params=True)
            bad += not euler_case(seed, params=True)
        if which == "tabeos":
            bad += not tabeos_case(seed)
        if which == "ic":
            bad += not ic_case(seed)
        if which == "run":
            bad += not run_case(seed)
        if which == "reuse":      # few distinct shapes, so that contexts really are reused
            REUSE = {} if REUSE is None else REUSE
            sizes = lambda rng, sd, small: (12 + 8 * (sd % 3), 20 + 4 * (sd % 2))  # noqa: E731
            bad += not mhd_case(seed)
            bad += not euler_case(seed)
        if which == "long":
            SCHEDULE = (19,)
            bad += not mhd_case(seed)
            bad += not euler_case(seed)
    print(f"{bad} mismatches")
    sys.exit(1 if bad else 0)
