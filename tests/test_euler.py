"""Euler SLIC (Euler_codes/SLIC_euler_2d.cpp, generaliseSLIC.cpp): CPU tests of the oracle against the fixtures
made by the real reference, and GPU parity tests (bitwise) of the CUDA path through the C ABI."""
import glob
import hashlib
import json
import os

import numpy as np
import pytest

from oracle_util import OracleEuler, euler_cfg, euler_ic, have_ref, run_euler_ref

GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "euler_*.npz")))
IDS = [os.path.basename(p)[6:-4] for p in GOLDEN]
DOM = (0.0, 1.0, 0.0, 1.0)


def bits(a):
    return np.ascontiguousarray(a).view(np.uint64)


# ------------------------------------------------------------------ CPU: oracle vs the real reference
@pytest.mark.parametrize("path", GOLDEN, ids=IDS)
def test_oracle_matches_golden(path):
    z = np.load(path)
    m = json.loads(str(z["meta"]))
    cfg = euler_cfg(m["nx"], m["ny"], DOM, bc=m["bc"])
    o = OracleEuler(cfg, euler_ic(cfg, m["ic"]))
    log = o.step(m["nsteps"])
    assert np.array_equal(bits(log), bits(z["log"]))
    assert np.array_equal(bits(o.state()[0]), bits(z["u"]))


def test_golden_present():
    assert len(GOLDEN) >= 4


@pytest.mark.skipif(not have_ref("slic_ref"), reason="oracle/_ref/slic_ref not built")
@pytest.mark.parametrize("exe,bc,ic,code,nx,ny,n", [
    ("slic_ref", "periodic", "ot", 0, 24, 28, 9),
    ("gslic_ref", "transmissive", "quadrants", 0, 30, 22, 11),
    ("gslic_ref", "transmissive", "cylinder", 1, 32, 32, 14),
])
def test_oracle_matches_live_reference(exe, bc, ic, code, nx, ny, n):
    cfg = euler_cfg(nx, ny, DOM, bc=bc)
    o = OracleEuler(cfg, euler_ic(cfg, ic))
    log = o.step(n)
    ur, logr, _ = run_euler_ref(exe, nx, ny, DOM, 1e9, 0.8, 1.4, n, ic=code)
    assert np.array_equal(bits(log), bits(logr))
    assert np.array_equal(bits(o.state()[0]), bits(ur))


# ------------------------------------------------------------------ GPU parity
gpu = pytest.mark.gpu


@gpu
@pytest.mark.parametrize("path", GOLDEN, ids=IDS)
def test_cuda_matches_reference_golden(path):
    import lsc_internship_b200 as lsc
    z = np.load(path)
    m = json.loads(str(z["meta"]))
    s = lsc.SlicEuler2d(m["nx"], m["ny"], bc=m["bc"], problem=m["ic"])
    u0 = s.initial_condition()
    s.upload(u0)
    log = []
    for _ in range(m["nsteps"]):
        s.step(1)
        t, dt, _n = s.scalars()
        log.append((t, dt))
    u = s.download()
    assert s.ieee_fallback_count() == 0
    s.close()
    assert np.array_equal(bits(np.array(log)), bits(z["log"])), "(t, dt) log differs from the reference"
    assert np.array_equal(bits(u), bits(z["u"])), "state differs from the reference"


@gpu
@pytest.mark.parametrize("bc,ic,nx,ny,n", [
    ("periodic", "ot", 256, 256, 10),
    ("transmissive", "cylinder", 200, 136, 10),       # BASELINE config 3 at a size the oracle finishes quickly
    ("transmissive", "quadrants", 131, 67, 8),        # ragged
    ("periodic", "ot", 126, 126, 5),                  # exactly one y block
    ("periodic", "ot", 4, 4, 3),                      # minimum
    ("transmissive", "quadrants", 5, 9, 4),
])
def test_cuda_matches_oracle(bc, ic, nx, ny, n):
    import lsc_internship_b200 as lsc
    cfg = euler_cfg(nx, ny, DOM, bc=bc)
    u0 = euler_ic(cfg, ic)
    s = lsc.SlicEuler2d(nx, ny, bc=bc, problem=ic)
    assert np.array_equal(bits(s.initial_condition()), bits(u0)), "library IC differs from the reference formulas"
    s.upload(u0)
    assert np.array_equal(bits(s.download()), bits(u0))
    o = OracleEuler(cfg, u0)
    for k in range(n):
        lo = o.step(1)[0]
        s.step(1)
        t, dt, _ = s.scalars()
        assert (t, dt) == tuple(lo), f"step {k}"
    u, ur = s.download(), o.state()[0]
    s.close()
    # BASELINE tolerance: max relative L-inf <= 1e-12 per conserved variable; enforced gate: every bit
    for k in range(4):
        den = np.max(np.abs(ur[2:-2, 2:-2, k]))
        assert np.max(np.abs(u[2:-2, 2:-2, k] - ur[2:-2, 2:-2, k])) <= 1e-12 * max(den, 1e-300)
    assert np.array_equal(bits(u), bits(ur))


@gpu
@pytest.mark.parametrize("bc,ic,nx,ny,P,n", [
    ("periodic", "ot", 96, 64, 2, 6),                 # two slabs on a periodic ring: each is both neighbours
    ("periodic", "ot", 101, 40, 3, 5),                # ragged slabs
    ("transmissive", "cylinder", 120, 72, 4, 6),      # end slabs carry the G:203-208 images
    ("transmissive", "quadrants", 33, 20, 2, 5),
])
def test_slab_decomposition_does_not_change_a_bit(bc, ic, nx, ny, P, n):
    """SURVEY 8e (S/G row): P slabs along x, two halo columns per neighbour and one max per step.  The slab
    group (peer-copy halos, the same kernels and slab-edge path the NCCL ranks run) must reproduce the
    single-context state and (t, dt) exactly, and both must equal the oracle."""
    import lsc_internship_b200 as lsc
    cfg = euler_cfg(nx, ny, DOM, bc=bc)
    u0 = euler_ic(cfg, ic)
    one = lsc.SlicEuler2d(nx, ny, bc=bc, problem=ic)
    one.upload(u0)
    one.step(n)
    grp = lsc.SlicEuler2dGroup(P, nxCells=nx, nyCells=ny, bc=bc, problem=ic)
    grp.upload(u0)
    grp.step(n)
    o = OracleEuler(cfg, u0)
    o.step(n)
    assert grp.scalars() == one.scalars()
    a, b = grp.download(), one.download()
    assert np.array_equal(bits(a), bits(b))
    assert np.array_equal(bits(a), bits(o.state()[0]))
    # library-generated slabs (no whole-grid array on the host) give the same run
    grp2 = lsc.SlicEuler2dGroup(P, nxCells=nx, nyCells=ny, bc=bc, problem=ic)
    grp2.init_problem()
    grp2.step(n)
    assert np.array_equal(bits(grp2.download()), bits(a))
    for x in (one, grp, grp2):
        x.close()


@gpu
@pytest.mark.parametrize("nx,ny,n", [(40, 40, 6), (33, 21, 4), (130, 70, 3)])
def test_tabeos_cuda_matches_oracle_and_reference(nx, ny, n, tmp_path):
    """SURVEY 8f N2 / BASELINE config 5: the sweeps with T's table EoS (ConservativeToPrimative T:248-327 in every
    flux, SoundSpeed T:330-404 in the time step) against oracle/euler_oracle.c and, where it is built, the REAL
    TabEoS.cpp (oracle/_ref/tabeos_ref), on the synthetic Plasma19-format table: table parsing, IC, every (t, dt)
    and the full state, bit for bit."""
    import lsc_internship_b200 as lsc
    from oracle_util import parsed_eos_table, run_tabeos_ref, tabeos_ic, write_eos_table
    path = str(tmp_path / "Plasma19_EoS.txt")
    write_eos_table(path)
    table = lsc.SlicEuler2d.read_eos_table(path)
    for a, b in zip(table, parsed_eos_table()):
        assert np.array_equal(bits(a), bits(b)), "library table parser differs from strtod of the same file"
    cfg = euler_cfg(nx, ny, DOM, bc="transmissive")
    u0 = tabeos_ic(cfg, table)
    s = lsc.SlicEuler2d(nx, ny, bc="transmissive", problem="tabeos", tStop=0.012 / 10 ** 2.5)
    s.set_eos_table(*table)
    assert np.array_equal(bits(s.tabeos_initial_condition()), bits(u0))
    s.init_problem()
    assert np.array_equal(bits(s.download()), bits(u0))
    o = OracleEuler(cfg, u0)
    o.set_table(*table)
    for k in range(n):
        lo = o.step(1)[0]
        s.step(1)
        t, dt, _ = s.scalars()
        assert (t, dt) == tuple(lo), f"step {k}"
    u = s.download()
    assert np.array_equal(bits(u), bits(o.state()[0]))
    if have_ref("tabeos_ref"):
        ur, lg = run_tabeos_ref(nx, ny, n)
        assert np.array_equal(bits(u), bits(ur))
        assert (t, dt) == tuple(lg[-1])
    # the searches were warm-started (the synthetic table passes the monotonicity check); the reference's plain
    # bisection must give the same bits
    assert s.eos_search_is_warm
    cold = lsc.SlicEuler2d(nx, ny, bc="transmissive", problem="tabeos")
    cold.set_eos_table(*table)
    cold.set_eos_search(False)
    assert not cold.eos_search_is_warm
    cold.init_problem()
    cold.step(n)
    assert cold.scalars()[:2] == (t, dt)
    assert np.array_equal(bits(cold.download()), bits(u))
    cold.close()
    # slabs: the table path through the decomposition
    grp = lsc.SlicEuler2dGroup(2, nxCells=nx, nyCells=ny, bc="transmissive", problem="tabeos")
    for part in grp.parts:
        part.set_eos_table(*table)
    grp.init_problem()
    grp.step(n)
    assert np.array_equal(bits(grp.download()), bits(u))
    grp.close()
    s.close()


@gpu
def test_tabeos_non_monotone_table_falls_back_to_the_bisection(tmp_path):
    """A table whose energy columns are not ascending (two pressure rows swapped) fails the library's check: the
    searches are then the reference's bisections, whose result depends on the path -- compare with the oracle, which
    runs T's loops literally, on that same table."""
    import lsc_internship_b200 as lsc
    from oracle_util import tabeos_ic, write_eos_table
    path = str(tmp_path / "Plasma19_EoS.txt")
    write_eos_table(path)
    d, p, ss, en = lsc.SlicEuler2d.read_eos_table(path)
    en = en.copy()
    for row in range(40, 460, 7):  # inversions all over the pressure range
        en[[row, row + 1]] = en[[row + 1, row]]
    table = (d, p, ss, en)
    nx, ny, n = 48, 36, 4
    cfg = euler_cfg(nx, ny, DOM, bc="transmissive")
    u0 = tabeos_ic(cfg, table)
    s = lsc.SlicEuler2d(nx, ny, bc="transmissive", problem="tabeos")
    s.set_eos_table(*table)
    assert not s.eos_search_is_warm
    s.upload(u0)
    o = OracleEuler(cfg, u0)
    o.set_table(*table)
    for k in range(n):
        lo = o.step(1)[0]
        s.step(1)
        assert s.scalars()[:2] == tuple(lo), f"step {k}"
    assert np.array_equal(bits(s.download()), bits(o.state()[0]))
    s.close()


@gpu
@pytest.mark.parametrize("name", ["40x40_s6", "33x21_s4", "130x70_s3"])
def test_tabeos_cuda_matches_reference_golden(name, tmp_path):
    """tests/golden/tabeos_*.npz: states and (t, dt) logs written by the real TabEoS.cpp on the synthetic table"""
    import lsc_internship_b200 as lsc
    from oracle_util import write_eos_table
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", f"tabeos_{name}.npz"))
    nx, ny, n = int(z["nx"]), int(z["ny"]), int(z["nsteps"])
    path = str(tmp_path / "Plasma19_EoS.txt")
    write_eos_table(path)
    s = lsc.SlicEuler2d(nx, ny, bc="transmissive", problem="tabeos")
    s.set_eos_table(*lsc.SlicEuler2d.read_eos_table(path))
    s.init_problem()
    for k in range(n):
        s.step(1)
        t, dt, _ = s.scalars()
        assert (t, dt) == tuple(z["log"][k]), f"step {k}"
    assert np.array_equal(bits(s.download()), bits(z["u"]))
    s.close()


@gpu
def test_tab_eos_driver_is_a_drop_in(tmp_path):
    """drivers/tab_eos run like the reference program (table in the working directory): the same 't = .. dt = ..'
    lines and a byte-identical SLIC2d.dat as the real TabEoS.cpp produced (tests/golden/tabeos_program_48x40.npz)."""
    import subprocess
    from oracle_util import write_eos_table
    exe = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "drivers", "_bin", "tab_eos")
    if not os.path.exists(exe):
        pytest.skip("drivers/_bin/tab_eos not built")
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "tabeos_program_48x40.npz"))
    write_eos_table(str(tmp_path / "Plasma19_EoS.txt"))
    r = subprocess.run([exe, "--nx", "48", "--ny", "40"], cwd=tmp_path, capture_output=True, text=True, check=True)
    ref_lines = bytes(z["stdout"]).decode().splitlines()
    mine = r.stdout.splitlines()
    assert mine[0] == "finished loop!"
    assert mine[1:] == [ln for ln in ref_lines if ln.startswith("t = ")]
    assert (tmp_path / "SLIC2d.dat").read_bytes() == bytes(z["dat"])


@gpu
def test_random_state_parity():
    import lsc_internship_b200 as lsc
    nx, ny = 90, 70
    cfg = euler_cfg(nx, ny, DOM, bc="periodic")
    rng = np.random.default_rng(77)
    rho = rng.uniform(0.5, 2.0, (nx + 4, ny + 4))
    vx, vy = rng.uniform(-1, 1, (2, nx + 4, ny + 4))
    p = rng.uniform(0.5, 2.0, (nx + 4, ny + 4))
    u0 = np.stack([rho, rho * vx, rho * vy, p / 0.4 + 0.5 * rho * (vx * vx + vy * vy)], axis=-1)
    o = OracleEuler(cfg, u0)
    s = lsc.SlicEuler2d(nx, ny, bc="periodic")
    s.upload(u0)          # ghosts as given: the first dt sees row/column n+2 of the uploaded array (S:173-174)
    o.step(4)
    s.step(4)
    assert np.array_equal(bits(s.download()), bits(o.state()[0]))
    s.close()



@pytest.mark.gpu
def test_euler_cfl_upper_bound():
    """The Euler y sweep skips the exact signal speed for warps whose fp32 upper bound of speed/run_max is < 1
    (euler_physics.cuh: IdealEos::cfl_speed_over_max_f32).  bound*run_max must be >= the exact speed (up to 1e-12
    of run_max, the fp32 underflow floor) wherever the bound is finite, and a state that could raise the maximum is
    never skipped.  Regimes: solver range, pressure = tiny remainder of E, at rest, scales around the fp32 range."""
    import lsc_internship_b200 as lsc
    rng = np.random.default_rng(77)
    n = 300000
    g = 1.4
    rho = 10.0 ** rng.uniform(-3, 3, n)
    v = rng.uniform(-3, 3, (n, 2)) * 10.0 ** rng.uniform(-6, 2, (n, 1))
    p = 10.0 ** rng.uniform(-9, 3, n)
    q = n // 4
    v[:q] = 0.0
    scale = np.ones(n)
    scale[q:2 * q] = 10.0 ** rng.uniform(-30, 30, q)
    U = np.empty((n, 4))
    U[:, 0] = rho * scale
    U[:, 1:3] = rho[:, None] * v * scale[:, None]
    U[:, 3] = (p / (g - 1) + 0.5 * rho * (v * v).sum(1)) * scale
    med = float(np.median(np.sqrt(g * p / rho)))
    for run_max in (med, 1e-3 * med, 1e3 * med, 1e-20, 1e20, 0.0):
        exact, bound = lsc.euler_kat_cfl_bound(U, g, run_max)
        trusted = np.isfinite(bound)
        ok = ~np.isfinite(exact) | ~trusted | (bound * run_max >= exact - 1e-12 * run_max)
        assert ok.all(), f"run_max={run_max}: {(~ok).sum()} states with bound < exact"
        assert not np.any((bound < 1.0) & (exact >= run_max))
        if run_max == 0.0:
            assert not trusted.any()
    exact, bound = lsc.euler_kat_cfl_bound(U, g, med)
    ordinary = np.isfinite(bound) & np.isfinite(exact) & (scale == 1.0) & (p > 1e-3 * U[:, 3])
    assert ordinary.sum() > n // 4
    assert np.median(bound[ordinary] * med / exact[ordinary]) < 1.01


@gpu
def test_out_of_window_operands_take_the_ieee_path():
    import lsc_internship_b200 as lsc
    nx, ny = 40, 36
    cfg = euler_cfg(nx, ny, DOM, bc="periodic")
    u0 = euler_ic(cfg, "ot")
    u0[5:9, 6:20, 1] = 1e-210
    o = OracleEuler(cfg, u0)
    s = lsc.SlicEuler2d(nx, ny)
    s.upload(u0)
    o.step(3)
    s.step(3)
    assert s.ieee_fallback_count() >= 1
    assert np.array_equal(bits(s.download()), bits(o.state()[0]))
    s.close()


@gpu
@pytest.mark.parametrize("scale_exp", [-1074, -575, -568, -560])
def test_subnormal_scalings_match_the_reference_literally(scale_exp):
    """0.5*xi*Delta issued as one fma double-rounds once xi*Delta is subnormal.  Momentum patches (both components,
    so that rho*vx*vy underflows and the tangential momentum keeps its tiny detail through the sweep) of monotone
    ramps through exact zeros on a gas at rest: at 2^-1074 the momenta themselves are subnormal (and xi*Delta
    underflows to zero); at 2^-575 .. 2^-560 they are ~1e-160 and xi*Delta ~ dM*Delta*1e8 is a non-zero subnormal
    (the `+ 1e-8` of S:266 makes the limiter ratio tiny) next to u0 == 0 -- the build before the literal forms
    fails these three (profiles/r2/subnormal_kat_previous_build.log).  Those blocks are flagged and recomputed with
    the reference's literal expressions: bit equality with the oracle, which evaluates S:281-282 as written."""
    import lsc_internship_b200 as lsc
    nx, ny = 56, 48
    cfg = euler_cfg(nx, ny, DOM, bc="periodic")
    u0 = np.zeros((nx + 4, ny + 4, 4))      # gas at rest: the flux differences of rho, E and p vanish exactly, so
    u0[..., 0] = 1.0                        # the tiny detail of the momenta survives the step
    u0[..., 3] = 1.0 / 0.4
    rng = np.random.default_rng(500 - scale_exp)

    def ramps(shape):
        ri = np.cumsum(rng.integers(1, 1 << 30, shape[0]))
        rj = np.cumsum(rng.integers(1, 1 << 30, shape[1]))
        k = ri[:, None] + rj[None, :]
        k = k - k[shape[0] // 2, shape[1] // 2]
        k[shape[0] // 2 - 1, :] = 0
        return np.ldexp(k.astype(np.float64), scale_exp)

    u0[6:26, 5:25, 1] = ramps((20, 20))
    u0[6:26, 5:25, 2] = -ramps((20, 20))
    u0[30:50, 20:40, 2] = ramps((20, 20))
    o = OracleEuler(cfg, u0)
    s = lsc.SlicEuler2d(nx, ny)
    s.upload(u0)
    for n in (1, 1, 2):
        o.step(n)
        s.step(n)
        ref = o.state()[0]
        assert np.array_equal(bits(s.download()), bits(ref))
        assert np.count_nonzero((np.abs(ref) < 2.0 ** -500) & (ref != 0)) > 300
    assert s.ieee_fallback_count() >= 1
    s.close()


@gpu
@pytest.mark.parametrize("bc,ic,nx,ny", [
    ("periodic", "ot", 96, 64),
    ("transmissive", "cylinder", 70, 130),
    ("transmissive", "quadrants", 8, 40),             # narrowest slab that is split (4 + 4 columns)
])
def test_split_y_sweep_with_overlapped_ghost_fill(bc, ic, nx, ny):
    """Halo/compute overlap order on one rank (lsc_euler_set_overlap(ctx, 2)): slab-edge columns of the y sweep
    first, ghost fill on the second stream while the interior is swept; edge and interior blocks on the IEEE fix
    lists.  Must not change a bit."""
    import lsc_internship_b200 as lsc
    cfg = euler_cfg(nx, ny, DOM, bc=bc)
    u0 = euler_ic(cfg, ic)
    u0[2:4, 6:20, 1] = 1e-210
    u0[nx // 2, 5:9, 2] = 1e-210
    u0[nx + 1, 12:20, 1] = 1e-210
    o = OracleEuler(cfg, u0)
    s = lsc.SlicEuler2d(nx, ny, bc=bc, problem=ic)
    s.set_overlap(2)
    s.upload(u0)
    log = o.step(6)
    s.step(6)
    t, dt, _ = s.scalars()
    assert (t, dt) == tuple(log[-1])
    assert s.ieee_fallback_count() >= 3
    assert np.array_equal(bits(s.download()), bits(o.state()[0]))
    s.close()


@gpu
def test_committed_program_reproduces_reference_dat_files(tmp_path, golden_dir):
    """SlicEuler2d().run() = main() of SLIC_euler_2d.cpp as committed: all twenty euler_<k>.dat must be
    byte-identical to the files committed in the reference repo (MHD_plot_data/euler_*.dat)."""
    import lsc_internship_b200 as lsc
    sha = json.load(open(os.path.join(golden_dir, "dat_sha256.json")))
    s = lsc.SlicEuler2d()
    s.init_problem()
    times = []
    steps = s.run(outdir=str(tmp_path), stdout=times)
    s.close()
    assert steps == 152       # the reference takes 152 steps (oracle/_ref/slic_ref)
    for k in range(1, 21):
        h = hashlib.sha256(open(tmp_path / f"euler_{k}.dat", "rb").read()).hexdigest()
        assert h == sha[f"euler_{k}.dat"], f"euler_{k}.dat differs from the reference's committed file"


@gpu
def test_generalise_slic_program(tmp_path, golden_dir):
    """generaliseSLIC.cpp as committed (transmissive, four quadrants): SLIC2d.dat byte-identical to what the real
    program writes, and the 't = .. dt = ..' log identical."""
    import lsc_internship_b200 as lsc
    sha = json.load(open(os.path.join(golden_dir, "dat_sha256.json")))
    s = lsc.SlicEuler2d(bc="transmissive", problem="quadrants")
    s.init_problem()
    lines = []
    n = s.run_generalise(outdir=str(tmp_path), log=lambda k, t, dt: lines.append(f"t = {t:.6g} dt = {dt:.6g}"))
    s.close()
    ref = open(os.path.join(golden_dir, "generalise_slic_stdout.txt")).read().splitlines()
    assert n == len(ref) == 141
    assert lines == ref
    h = hashlib.sha256(open(tmp_path / "SLIC2d.dat", "rb").read()).hexdigest()
    assert h == sha["SLIC2d.dat@generaliseSLIC"]


@gpu
def test_cpp_drivers_are_drop_ins(tmp_path, golden_dir):
    """drivers/slic_euler_2d with no arguments = SLIC_euler_2d.cpp; with --generalise = generaliseSLIC.cpp: stdout and
    output files identical to the real programs'."""
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = os.path.join(root, "drivers", "_bin", "slic_euler_2d")
    if not os.path.exists(exe):
        subprocess.run(["make", "-C", os.path.join(root, "drivers")], check=True)
    sha = json.load(open(os.path.join(golden_dir, "dat_sha256.json")))
    a = tmp_path / "s"
    a.mkdir()
    r = subprocess.run([exe], cwd=a, check=True, capture_output=True, text=True)
    assert r.stdout == open(os.path.join(golden_dir, "slic_euler_2d_stdout.txt")).read()
    for k in (1, 7, 20):
        assert hashlib.sha256(open(a / f"euler_{k}.dat", "rb").read()).hexdigest() == sha[f"euler_{k}.dat"]
    b = tmp_path / "g"
    b.mkdir()
    r = subprocess.run([exe, "--generalise"], cwd=b, check=True, capture_output=True, text=True)
    assert r.stdout == open(os.path.join(golden_dir, "generalise_slic_stdout.txt")).read()
    assert hashlib.sha256(open(b / "SLIC2d.dat", "rb").read()).hexdigest() == sha["SLIC2d.dat@generaliseSLIC"]


@gpu
def test_large_grid_conservation():
    """4096^2 (BASELINE config 3 size) with periodic BCs: mass, momenta and energy are conserved to rounding."""
    import lsc_internship_b200 as lsc
    n = 4096
    s = lsc.SlicEuler2d(n, n, bc="periodic", problem="ot")
    s.init_problem()
    u0 = s.download()[2:-2, 2:-2]
    s.step(5)
    u1 = s.download()[2:-2, 2:-2]
    s.close()
    assert np.isfinite(u1).all()
    for k in (0, 3):
        a, b = u0[..., k].sum(dtype=np.longdouble), u1[..., k].sum(dtype=np.longdouble)
        assert abs(a - b) / abs(a) < 1e-12


@gpu
def test_euler_time_loop_is_device_resident():
    """dt, t and the sweeps' scalars are formed on the device (euler_prepare_kernel); a call that advances many steps
    synchronises with the host once, run_until's log shows every step's (t, dt) as the reference prints them."""
    import lsc_internship_b200 as lsc
    nx, ny, n = 72, 60, 43
    cfg = euler_cfg(nx, ny, DOM, bc="transmissive")
    u0 = euler_ic(cfg, "cylinder")
    o = OracleEuler(cfg, u0)
    log = o.step(n)
    s = lsc.SlicEuler2d(nx, ny, bc="transmissive", problem="cylinder")
    s.upload(u0)
    a, _ = s.loop_stats()
    s.step(n)
    b, _ = s.loop_stats()
    assert b - a == 1
    t, dt, k = s.scalars()
    assert (t, dt) == tuple(log[-1]) and k == n
    assert np.array_equal(bits(s.download()), bits(o.state()[0]))
    s.upload(u0)
    o2 = OracleEuler(cfg, u0)
    lines = []
    nrun = s.run_until(0.05, log=lambda k, t, dt: lines.append((t, dt)))
    ref = o2.step(nrun)
    assert np.array_equal(bits(np.array(lines)), bits(ref))
    assert ref[-1][0] >= 0.05 > ref[-2][0]
    c, _ = s.loop_stats()
    assert c - b <= 3
    assert np.array_equal(bits(s.download()), bits(o2.state()[0]))
    s.close()


@gpu
def test_cpp_drivers_with_slabs_are_the_same_programs(tmp_path, golden_dir):
    """drivers/slic_euler_2d and drivers/tab_eos cut into slabs inside ONE process (--gpus N on a multi-GPU box, --slabs
    3 on one GPU): stdout and files identical to the single-context runs, i.e. to the reference programs'."""
    import subprocess
    import lsc_internship_b200 as lsc
    from oracle_util import write_eos_table
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    ndev = lsc.lib().lsc_device_count()
    opt = ["--gpus", str(min(ndev, 4))] if ndev >= 2 else ["--slabs", "3"]
    exe = os.path.join(root, "drivers", "_bin", "slic_euler_2d")
    sha = json.load(open(os.path.join(golden_dir, "dat_sha256.json")))
    a = tmp_path / "s"
    a.mkdir()
    r = subprocess.run([exe] + opt, cwd=a, check=True, capture_output=True, text=True)
    assert r.stdout == open(os.path.join(golden_dir, "slic_euler_2d_stdout.txt")).read()
    for k in (1, 7, 20):
        assert hashlib.sha256(open(a / f"euler_{k}.dat", "rb").read()).hexdigest() == sha[f"euler_{k}.dat"]
    b = tmp_path / "g"
    b.mkdir()
    r = subprocess.run([exe, "--generalise"] + opt, cwd=b, check=True, capture_output=True, text=True)
    assert r.stdout == open(os.path.join(golden_dir, "generalise_slic_stdout.txt")).read()
    assert hashlib.sha256(open(b / "SLIC2d.dat", "rb").read()).hexdigest() == sha["SLIC2d.dat@generaliseSLIC"]
    # tab_eos: one context against the slabs, byte for byte
    exe = os.path.join(root, "drivers", "_bin", "tab_eos")
    outs = []
    for name, extra in (("t1", []), ("tn", opt)):
        d = tmp_path / name
        d.mkdir()
        write_eos_table(str(d / "Plasma19_EoS.txt"))
        r = subprocess.run([exe, "--nx", "48", "--ny", "40"] + extra, cwd=d, check=True, capture_output=True, text=True)
        outs.append((r.stdout, open(d / "SLIC2d.dat", "rb").read()))
    assert outs[0] == outs[1]
