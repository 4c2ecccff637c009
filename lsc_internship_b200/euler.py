"""Host-side mirror of the reference programs Euler_codes/SLIC_euler_2d.cpp (S) and
Euler_codes/generaliseSLIC.cpp (G): 2-D Euler, SLIC (limited MUSCL-Hancock + FORCE flux), averaged
XthenY / YthenX.  As for MHD, the reference has no callable API; `SlicEuler2d` exposes the constants of
main() (S:529-540), the initial-condition block, the boundary loops, the stdout line and the .dat files,
and drives the CUDA library through its C ABI.  No numerical code lives here."""
import ctypes
import os

import numpy as np

from . import _capi
from ._capi import EulerConfig, check, lib

_DP = ctypes.POINTER(ctypes.c_double)


def _dp(a):
    return a.ctypes.data_as(_DP)


class SlicEuler2d:
    """Defaults = the committed configuration of SLIC_euler_2d.cpp: 100x100 on [0,1]^2, periodic, Euler
    "Orszag-Tang" initial data, C = 0.8, gamma = 1.4, omega = 0, tStop = 0.5.
    bc="transmissive" + problem="quadrants" gives generaliseSLIC.cpp's committed configuration."""

    NVAR = 4

    def __init__(self, nxCells=100, nyCells=100, x0=0.0, x1=1.0, y0=0.0, y1=1.0, tStop=0.5, C=0.8, gamma=1.4,
                 omega=0.0, bc="periodic", problem="ot", device=0, rank=0, nranks=1):
        self.tStop = float(tStop)
        self.problem = _capi.EULER_PROBLEMS[problem] if isinstance(problem, str) else int(problem)
        bcv = {"periodic": _capi.BC_PERIODIC, "transmissive": _capi.BC_TRANSMISSIVE_E}[bc]
        self.cfg = EulerConfig(int(nxCells), int(nyCells), float(x0), float(x1), float(y0), float(y1), float(gamma),
                               float(C), float(omega), bcv, int(device), int(rank), int(nranks))
        self._h = ctypes.c_void_p()
        check(lib().lsc_euler_create(ctypes.byref(self.cfg), ctypes.byref(self._h)))
        # slab of this rank: nxl interior columns starting at global interior column i0
        self.nxl = lib().lsc_euler_local_nx(self._h)
        self.i0 = lib().lsc_euler_local_i0(self._h)

    # -- tabulated equation of state (MHD_codes/TabEoS.cpp = T) -----------------------------------------
    @staticmethod
    def read_eos_table(path):
        """Plasma19() of T:18-133: (densities[500], pressures[500], sound_speeds[500,500], energies[500,500])."""
        dens, pres = np.empty(500), np.empty(500)
        sound, en = np.empty((500, 500)), np.empty((500, 500))
        check(lib().lsc_eos_table_read(os.fsencode(path), _dp(dens), _dp(pres), _dp(sound), _dp(en)))
        return dens, pres, sound, en

    def set_eos_table(self, densities, pressures, sound_speeds, energies):
        """Every pressure / sound speed of the sweeps now comes from the table (T:248-404); gamma is unused."""
        self.table = tuple(np.ascontiguousarray(a, dtype=np.float64) for a in
                           (densities, pressures, sound_speeds, energies))
        assert self.table[0].shape == (500,) and self.table[1].shape == (500,)
        assert self.table[2].shape == (500, 500) and self.table[3].shape == (500, 500)
        check(lib().lsc_euler_set_eos_table(self._h, *[_dp(a) for a in self.table]))

    def set_eos_search(self, warm_start):
        """warm_start=False: every table search is the reference's bisection; True (default): warm-started searches
        when the table passed the library's monotonicity check (bit-identical results either way)."""
        check(lib().lsc_euler_set_eos_search(self._h, 1 if warm_start else 0))

    @property
    def eos_search_is_warm(self):
        return bool(lib().lsc_euler_eos_search_is_warm(self._h))

    def tabeos_initial_condition(self):
        """IC loop of T's main() (T:912-954) as a host array."""
        u = np.empty(self.shape, dtype=np.float64)
        d, p, _, e = self.table
        check(lib().lsc_euler_tabeos_initial_condition(ctypes.byref(self.cfg), _dp(d), _dp(p), _dp(e), _dp(u)))
        return u

    def write_tabeos_dat(self, u, path):
        """T:974-999 (SLIC2d.dat through the table's ConservativeToPrimative)."""
        u = np.ascontiguousarray(u, dtype=np.float64)
        d, p, _, e = self.table
        check(lib().lsc_euler_tabeos_write_dat(ctypes.byref(self.cfg), _dp(u), os.fsencode(path), _dp(d), _dp(p), _dp(e)))

    def run_tabeos(self, outdir=None, log=None):
        """main() of TabEoS.cpp from T:959: do{}while(t < tStop), then SLIC2d.dat (T:974-999)."""
        n = self.run_until(self.tStop, log)
        if outdir:
            self.write_tabeos_dat(self.download(), os.path.join(outdir, "SLIC2d.dat"))
        return n

    def comm_init(self, unique_id):
        """Join the NCCL communicator of the slab decomposition (id from lsc_comm_unique_id on rank 0)."""
        buf = ctypes.create_string_buffer(bytes(unique_id), 128)
        check(lib().lsc_euler_comm_init(self._h, ctypes.cast(buf, ctypes.c_void_p)))

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            lib().lsc_euler_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def shape(self):
        """host array of this context: the whole padded grid, or this rank's padded slab"""
        return (self.nxl + 4, self.cfg.ny + 4, self.NVAR)

    def _prob(self, problem):
        if problem is None:
            return self.problem
        return _capi.EULER_PROBLEMS[problem] if isinstance(problem, str) else int(problem)

    def initial_condition(self, problem=None):
        """IC loop + the boundary fill after it (S:552-596), as a host array."""
        u = np.empty(self.shape, dtype=np.float64)
        check(lib().lsc_euler_initial_condition(ctypes.byref(self.cfg), self._prob(problem), _dp(u)))
        return u

    def upload(self, u):
        u = np.ascontiguousarray(u, dtype=np.float64)
        if u.shape != self.shape:
            raise ValueError(f"state must have shape {self.shape}, got {u.shape}")
        check(lib().lsc_euler_upload(self._h, _dp(u)))

    def init_problem(self, problem=None):
        check(lib().lsc_euler_init_problem(self._h, self._prob(problem)))

    def download(self, out=None):
        u = np.empty(self.shape, dtype=np.float64) if out is None else out
        if not (isinstance(u, np.ndarray) and u.dtype == np.float64 and u.shape == self.shape
                and u.flags["C_CONTIGUOUS"] and u.flags["WRITEABLE"]):
            raise ValueError(f"out must be a writeable C-contiguous float64 array of shape {self.shape}")
        check(lib().lsc_euler_download(self._h, _dp(u)))
        return u

    def state_hash(self):
        """this slab's word of the decomposition-independent interior fingerprint (lsc_euler_state_hash)"""
        h = ctypes.c_ulonglong()
        check(lib().lsc_euler_state_hash(self._h, ctypes.byref(h)))
        return h.value

    def step(self, nsteps=1):
        check(lib().lsc_euler_step(self._h, int(nsteps)))

    def step_timed(self, nsteps=1):
        ms = ctypes.c_float()
        check(lib().lsc_euler_step_timed(self._h, int(nsteps), ctypes.byref(ms)))
        return ms.value

    def set_overlap(self, mode=1):
        """0: y sweep, then halo exchange; 1 (default): exchange overlapped with the interior of the y sweep when
        the slab has NCCL neighbours; 2: split the y sweep on a single rank too (tests)."""
        check(lib().lsc_euler_set_overlap(self._h, int(mode)))

    def set_profiling(self, enabled=True):
        check(lib().lsc_euler_set_profiling(self._h, 1 if enabled else 0))

    def kernel_times(self):
        ms = (ctypes.c_double * 2)()
        n = (ctypes.c_long * 2)()
        check(lib().lsc_euler_kernel_times(self._h, ms, n))
        return ms[0], ms[1], n[0], n[1]

    def scalars(self):
        t, dt = ctypes.c_double(), ctypes.c_double()
        n = ctypes.c_long()
        check(lib().lsc_euler_get_scalars(self._h, ctypes.byref(t), ctypes.byref(dt), ctypes.byref(n)))
        return t.value, dt.value, n.value

    def ieee_fallback_count(self):
        return lib().lsc_euler_ieee_fallback_count(self._h)

    def loop_stats(self):
        """(host synchronisations of the time loop so far, 0): the Euler loop has no exp table"""
        a = ctypes.c_long()
        check(lib().lsc_euler_loop_stats(self._h, ctypes.byref(a)))
        return a.value, 0

    def run_until(self, t_stop, log=None):
        """do { step } while (t < t_stop) -- S:600-617 / G:627-646.  Returns the number of steps."""
        err = []

        def on_log(n, t, dt, _u):  # an exception in a ctypes callback would be swallowed: keep it, stop the loop
            try:
                log(n, t, dt)
                return 0
            except BaseException as e:  # noqa: BLE001
                err.append(e)
                return 1

        cb = _capi.EULER_LOG_CB(on_log) if log else ctypes.cast(None, _capi.EULER_LOG_CB)
        n = ctypes.c_long()
        rc = lib().lsc_euler_run_until(self._h, float(t_stop), cb, None, ctypes.byref(n))
        if err:
            raise err[0]
        check(rc)
        return n.value

    def write_dat(self, u, path, blank_line_per_row=False):
        u = np.ascontiguousarray(u, dtype=np.float64)
        check(lib().lsc_euler_write_dat(ctypes.byref(self.cfg), _dp(u), os.fsencode(path), int(blank_line_per_row)))

    def run(self, outdir=None, stdout=None):
        """main() of SLIC_euler_2d.cpp from S:598: twenty segments, segment k runs do{}while(t < tStop/(21-k)) and
        then writes euler_<k>.dat and prints the segment time (S:618-642).  Returns total steps."""
        total = 0
        buf = np.empty(self.shape, dtype=np.float64) if outdir else None
        for counter in range(1, 21):
            seg = self.tStop / (21 - counter)
            total += self.run_until(seg)
            if stdout is not None:
                stdout.append(seg)
            if outdir:
                self.download(buf)
                self.write_dat(buf, os.path.join(outdir, f"euler_{counter}.dat"))
        return total

    def run_generalise(self, outdir=None, log=None):
        """main() of generaliseSLIC.cpp from G:625: one do{}while(t < tStop), then SLIC2d.dat (G:645-667)."""
        n = self.run_until(self.tStop, log)
        if outdir:
            u = self.download()
            self.write_dat(u, os.path.join(outdir, "SLIC2d.dat"), blank_line_per_row=True)
        return n


def euler_kat_cfl_bound(U, gamma, run_max, device=0):
    """(exact fp64 signal speed S:175-184, fp32 upper bound of speed/run_max used by the Euler y sweep's warp
    filter) per state; U is (n, 4) conserved states."""
    U = np.ascontiguousarray(U, dtype=np.float64)
    out = np.empty((U.shape[0], 2), dtype=np.float64)
    check(lib().lsc_euler_kat_cfl_bound(device, U.shape[0], _dp(U), float(gamma), float(run_max), _dp(out)))
    return out[:, 0], out[:, 1]


class SlicEuler2dGroup:
    """The n slabs of one grid advanced in lock step inside ONE process (lsc_euler_group_step): halos by peer
    copies, the CFL maximum on the host.  Same kernels and slab-edge code path as the NCCL mode; used to check
    that the decomposition does not change a single bit."""

    def __init__(self, nslabs, devices=None, **kw):
        ndev = lib().lsc_device_count()
        devices = devices if devices is not None else [r % max(ndev, 1) for r in range(nslabs)]
        self.parts = [SlicEuler2d(device=devices[r], rank=r, nranks=nslabs, **kw) for r in range(nslabs)]
        self._arr = (ctypes.c_void_p * nslabs)(*[p._h for p in self.parts])
        self.nx, self.ny = self.parts[0].cfg.nx, self.parts[0].cfg.ny

    def init_problem(self, problem=None):
        for p in self.parts:
            p.init_problem(problem)

    def upload(self, u):
        """u: the whole padded grid [(nx+4)][(ny+4)][4]; every slab takes its columns incl. halos"""
        for p in self.parts:
            p.upload(u[p.i0:p.i0 + p.nxl + 4])

    def step(self, nsteps=1):
        check(lib().lsc_euler_group_step(self._arr, len(self.parts), int(nsteps)))

    def download(self):
        u = np.empty((self.nx + 4, self.ny + 4, 4), dtype=np.float64)
        for r, p in enumerate(self.parts):
            a = p.download()
            lo = 0 if r == 0 else 2
            hi = p.nxl + 4 if r == len(self.parts) - 1 else p.nxl + 2
            u[p.i0 + lo:p.i0 + hi] = a[lo:hi]
        return u

    def scalars(self):
        return self.parts[0].scalars()

    def close(self):
        for p in self.parts:
            p.close()
