"""ctypes access to the CPU oracle (oracle/liblsc_oracle.so) and to the real
reference binaries in oracle/_ref/.  Test infrastructure only."""
import ctypes
import os
import subprocess
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
REF_DIR = os.path.join(ORACLE_DIR, "_ref")

BCX = {"periodic": 0, "transmissive": 1}
BCY = {"reflective": 0, "periodic": 1, "transmissive": 2}
PROB = {"kh": 0, "ot": 1}


class OracleMhdCfg(ctypes.Structure):
    _fields_ = [("nx", ctypes.c_int), ("ny", ctypes.c_int),
                ("x0", ctypes.c_double), ("x1", ctypes.c_double),
                ("y0", ctypes.c_double), ("y1", ctypes.c_double),
                ("gamma", ctypes.c_double), ("cfl", ctypes.c_double),
                ("bcx", ctypes.c_int), ("bcy", ctypes.c_int)]


_lib = None


def oracle_lib():
    global _lib
    if _lib is None:
        path = os.path.join(ORACLE_DIR, "liblsc_oracle.so")
        if not os.path.exists(path):
            subprocess.run(["make", "-C", ORACLE_DIR, "liblsc_oracle.so"], check=True)
        _lib = ctypes.CDLL(path)
        _lib.lsc_oracle_mhd_create.restype = ctypes.c_void_p
        _lib.lsc_oracle_mhd_run_until.restype = ctypes.c_long
    return _lib


def _dp(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_double))


def mhd_cfg(nx, ny, dom=(0.0, 1.0, 0.0, 1.0), gamma=5.0 / 3.0, cfl=0.8, bcx="periodic", bcy="periodic"):
    return OracleMhdCfg(nx, ny, dom[0], dom[1], dom[2], dom[3], gamma, cfl, BCX[bcx], BCY[bcy])


def mhd_ic(cfg, prob):
    u = np.zeros((cfg.nx + 4, cfg.ny + 4, 9), dtype=np.float64)
    oracle_lib().lsc_oracle_mhd_ic(ctypes.byref(cfg), PROB[prob], _dp(u))
    return u


class OracleMhd:
    def __init__(self, cfg, u0):
        self.cfg = cfg
        self.lib = oracle_lib()
        u0 = np.ascontiguousarray(u0, dtype=np.float64)
        assert u0.shape == (cfg.nx + 4, cfg.ny + 4, 9)
        self.h = ctypes.c_void_p(self.lib.lsc_oracle_mhd_create(ctypes.byref(cfg), _dp(u0)))

    def step(self, n=1):
        log = np.zeros((n, 3), dtype=np.float64)
        self.lib.lsc_oracle_mhd_step(self.h, ctypes.c_long(n), _dp(log))
        return log

    def run_until(self, t_stop):
        return self.lib.lsc_oracle_mhd_run_until(self.h, ctypes.c_double(t_stop))

    def state(self):
        u = np.zeros((self.cfg.nx + 4, self.cfg.ny + 4, 9), dtype=np.float64)
        t = ctypes.c_double()
        dt = ctypes.c_double()
        ch = ctypes.c_double()
        self.lib.lsc_oracle_mhd_get(self.h, _dp(u), ctypes.byref(t), ctypes.byref(dt), ctypes.byref(ch))
        return u, t.value, dt.value, ch.value

    def close(self):
        if self.h:
            self.lib.lsc_oracle_mhd_destroy(self.h)
            self.h = None

    def __del__(self):
        self.close()


def have_ref(name="muscl_ref"):
    return os.path.exists(os.path.join(REF_DIR, name))


def run_muscl_ref(prob, nx, ny, dom, t_stop, cfl, bcx, bcy, nsteps, cwd=None, writedat=False):
    """Run the real reference (patched only for configuration) and return
    (state AoS incl. ghosts, steplog[nsteps,3] of (t,dt,ch), stderr text)."""
    with tempfile.TemporaryDirectory() as td:
        st = os.path.join(td, "state.bin")
        lg = os.path.join(td, "log.bin")
        cmd = [os.path.join(REF_DIR, "muscl_ref"), prob, str(nx), str(ny)] + [repr(float(d)) for d in dom] + \
              [repr(float(t_stop)), repr(float(cfl)), bcx, bcy, str(nsteps), "1" if writedat else "0", st, lg]
        r = subprocess.run(cmd, cwd=cwd or td, check=True, capture_output=True, text=True)
        u = np.fromfile(st, dtype=np.float64).reshape(nx + 4, ny + 4, 9)
        log = np.fromfile(lg, dtype=np.float64).reshape(-1, 3)
    return u, log, r.stderr


# ----------------------------------------------------------------------------------------------
# Euler SLIC (Euler_codes/SLIC_euler_2d.cpp, generaliseSLIC.cpp)
# ----------------------------------------------------------------------------------------------
EULER_BC = {"periodic": 0, "transmissive": 1}
EULER_IC = {"ot": 0, "quadrants": 1, "cylinder": 2}


class OracleEulerCfg(ctypes.Structure):
    _fields_ = [("nx", ctypes.c_int), ("ny", ctypes.c_int),
                ("x0", ctypes.c_double), ("x1", ctypes.c_double),
                ("y0", ctypes.c_double), ("y1", ctypes.c_double),
                ("gamma", ctypes.c_double), ("cfl", ctypes.c_double), ("omega", ctypes.c_double),
                ("bc", ctypes.c_int)]


def euler_cfg(nx, ny, dom=(0.0, 1.0, 0.0, 1.0), gamma=1.4, cfl=0.8, omega=0.0, bc="periodic"):
    return OracleEulerCfg(nx, ny, dom[0], dom[1], dom[2], dom[3], gamma, cfl, omega, EULER_BC[bc])


def euler_ic(cfg, ic):
    u = np.zeros((cfg.nx + 4, cfg.ny + 4, 4), dtype=np.float64)
    oracle_lib().lsc_oracle_euler_ic(ctypes.byref(cfg), EULER_IC[ic], _dp(u))
    return u


class OracleEuler:
    def __init__(self, cfg, u0):
        self.cfg = cfg
        self.lib = oracle_lib()
        self.lib.lsc_oracle_euler_create.restype = ctypes.c_void_p
        u0 = np.ascontiguousarray(u0, dtype=np.float64)
        assert u0.shape == (cfg.nx + 4, cfg.ny + 4, 4)
        self.h = ctypes.c_void_p(self.lib.lsc_oracle_euler_create(ctypes.byref(cfg), _dp(u0)))

    def set_table(self, dens, pres, sound, energies):
        """switch this instance to TabEoS.cpp's equation of state (T:147-438)"""
        args = [np.ascontiguousarray(a, dtype=np.float64) for a in (dens, pres, sound, energies)]
        self.lib.lsc_oracle_euler_set_table(self.h, *[_dp(a) for a in args])

    def step(self, n=1):
        log = np.zeros((n, 2), dtype=np.float64)
        self.lib.lsc_oracle_euler_step(self.h, ctypes.c_long(n), _dp(log))
        return log

    def state(self):
        u = np.zeros((self.cfg.nx + 4, self.cfg.ny + 4, 4), dtype=np.float64)
        t, dt = ctypes.c_double(), ctypes.c_double()
        self.lib.lsc_oracle_euler_get(self.h, _dp(u), ctypes.byref(t), ctypes.byref(dt))
        return u, t.value, dt.value

    def close(self):
        if self.h:
            self.lib.lsc_oracle_euler_destroy(self.h)
            self.h = None

    def __del__(self):
        self.close()


def run_euler_ref(which, nx, ny, dom, t_stop, cfl, gamma, nsteps, ic=0, cwd=None, writedat=False):
    """Run oracle/_ref/slic_ref (periodic, S) or gslic_ref (transmissive, G; ic 0 = its quadrants,
    1 = cylindrical explosion).  Returns (state incl. ghosts, log[n,2] of (t,dt), stderr)."""
    with tempfile.TemporaryDirectory() as td:
        st, lg = os.path.join(td, "state.bin"), os.path.join(td, "log.bin")
        cmd = [os.path.join(REF_DIR, which), str(nx), str(ny)] + [repr(float(d)) for d in dom] + \
              [repr(float(t_stop)), repr(float(cfl)), repr(float(gamma)), str(nsteps), "1" if writedat else "0",
               str(ic), st, lg]
        r = subprocess.run(cmd, cwd=cwd or td, check=True, capture_output=True, text=True)
        u = np.fromfile(st, dtype=np.float64).reshape(nx + 4, ny + 4, 4)
        log = np.fromfile(lg, dtype=np.float64).reshape(-1, 2)
    return u, log, r.stderr


# ---------------------------------------------------------------------------------------------------------
# Tabulated EoS (MHD_codes/TabEoS.cpp = T).  The real Plasma19_EoS.txt is not in the reference tree (stripped
# large blob), so every checker and the CUDA path run on a SYNTHETIC table in the same file format (T:36-128):
# line 10 = 500 densities, line 12 = 500 pressures, lines 14-513 sound speed, 515-1014 specific internal energy,
# 1016-1515 temperature, 1517-2016 electrical and 2018-2517 thermal conductivity; rows = pressure index, columns
# = density index.  A gas whose adiabatic index drifts from 1.6 (low p) to 1.2 (high p): not an ideal gas (so the
# lookup is really exercised), e(p) strictly increasing at fixed rho (T's reverse search needs that).
# ---------------------------------------------------------------------------------------------------------
import sys  # noqa: E402
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
from lsc_internship_b200.synthetic_eos import eos_table, write_eos_table  # noqa: E402,F401  (the synthetic table)


def run_tabeos_ref(nx, ny, nsteps, dom=(0.0, 1.0, 0.0, 1.0), t_stop=1e9, cfl=0.8, workdir=None):
    """Run oracle/_ref/tabeos_ref (the real TabEoS.cpp) on the synthetic table; returns (u, [(t, dt)...])."""
    import tempfile
    exe = os.path.join(REF_DIR, "tabeos_ref")
    with tempfile.TemporaryDirectory() as td:
        write_eos_table(os.path.join(td, "Plasma19_EoS.txt"))
        dump, log = os.path.join(td, "u.bin"), os.path.join(td, "log.bin")
        subprocess.run([exe, str(nx), str(ny)] + [repr(float(v)) for v in dom] + [repr(float(t_stop)), repr(float(cfl)),
                       str(nsteps), "0", dump, log], cwd=td, check=True, capture_output=True)
        u = np.fromfile(dump, dtype=np.float64).reshape(nx + 4, ny + 4, 4)
        lg = np.fromfile(log, dtype=np.float64).reshape(-1, 2)
    return u, lg


def tabeos_ic(cfg, table):
    """IC of TabEoS.cpp main() (T:918-941 + boundary fill T:954) through the oracle's table P2C."""
    u = np.zeros((cfg.nx + 4, cfg.ny + 4, 4), dtype=np.float64)
    args = [np.ascontiguousarray(a, dtype=np.float64) for a in table]
    oracle_lib().lsc_oracle_tabeos_ic(ctypes.byref(cfg), *[_dp(a) for a in args], _dp(u))
    return u


_TABLE = None


def parsed_eos_table():
    """the synthetic table as the reference / the library parse it back from the text file (cached)"""
    global _TABLE
    if _TABLE is None:
        with tempfile.TemporaryDirectory() as td:
            path = os.path.join(td, "Plasma19_EoS.txt")
            write_eos_table(path)
            rows = open(path).read().split("\n")
        num = lambda r: np.array([float(x) for x in r.split()])  # noqa: E731
        dens, pres = num(rows[9]), num(rows[11])
        sound = np.array([num(r) for r in rows[13:513]])
        en = np.array([num(r) for r in rows[514:1014]])
        _TABLE = (dens, pres, sound, en)
    return _TABLE


# ---------------------------------------------------------------------------------------------------------
# First-order HLLC + GLM program (MHD_codes/HLLC_MHD_2D.cpp): the real program built by
# oracle/build_ref_hllc2d.py.  Arrays carry ONE ghost layer: [(nx+2)][(ny+2)][9].
# ---------------------------------------------------------------------------------------------------------
def run_hllc2d_ref(nx, ny, dom, t_stop, cfl, nsteps, cwd=None, writedat=False):
    """Returns (state after the program's final boundary fill, log[n,3] of (t, dt, ch), stdout, stderr)."""
    with tempfile.TemporaryDirectory() as td:
        st, lg = os.path.join(td, "state.bin"), os.path.join(td, "log.bin")
        cmd = [os.path.join(REF_DIR, "hllc2d_ref"), str(nx), str(ny)] + [repr(float(d)) for d in dom] + \
              [repr(float(t_stop)), repr(float(cfl)), str(nsteps), "1" if writedat else "0", st, lg]
        r = subprocess.run(cmd, cwd=cwd or td, check=True, capture_output=True, text=True)
        u = np.fromfile(st, dtype=np.float64).reshape(nx + 2, ny + 2, 9)
        log = np.fromfile(lg, dtype=np.float64).reshape(-1, 3)
    return u, log, r.stdout, r.stderr
