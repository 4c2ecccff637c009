"""Host-side mirror of the reference program MHD_codes/MUSCL_NEW.cpp.

The reference has no callable API: its "interface" is the block of constants at
the top of main() (MUSCL_NEW.cpp:659-671), the initial-condition block that is
commented in (:726-819), the boundary loops that are commented in (:607-655),
the per-step log line (:836) and the MHD<k>.dat files (:989-1001).  `MusclNew`
exposes exactly those knobs with the committed values as defaults and drives the
CUDA library through its C ABI; it holds no numerical code of its own.
"""
import ctypes
import os

import numpy as np

from . import _capi
from ._capi import MhdConfig, check, lib

_DP = ctypes.POINTER(ctypes.c_double)


def _dp(a):
    return a.ctypes.data_as(_DP)


class MusclNew:
    """2-D ideal MHD, MUSCL-Hancock + HLLC + mixed GLM, on one B200 (or one slab of a
    multi-GPU run).  Defaults are the committed configuration of MUSCL_NEW.cpp:
    Kelvin-Helmholtz, 128x256 on [0,1]x[-1,1], periodic-x / reflective-y, C=0.8,
    gamma=5/3, tStop=12."""

    NVAR = 9

    def __init__(self, nxCells=128, nyCells=256, x0=0.0, x1=1.0, y0=-1.0, y1=1.0, tStop=12.0, C=0.8,
                 gamma=5.0 / 3.0, bc_x="periodic", bc_y="reflective", problem="kh", device=0, rank=0, nranks=1,
                 arithmetic="reference", slab=None):
        self.tStop = float(tStop)
        self.problem = _capi.MHD_PROBLEMS[problem] if isinstance(problem, str) else int(problem)
        # slab = (first interior column, width): explicit bounds of this rank's slab (default: the even split)
        slab_i0, slab_nx = (0, 0) if slab is None else (int(slab[0]), int(slab[1]))
        self.cfg = MhdConfig(int(nxCells), int(nyCells), float(x0), float(x1), float(y0), float(y1),
                             float(gamma), float(C), _capi.BC_NAMES[bc_x], _capi.BC_NAMES[bc_y],
                             int(device), int(rank), int(nranks), slab_i0, slab_nx)
        self._h = ctypes.c_void_p()
        check(lib().lsc_mhd_create(ctypes.byref(self.cfg), ctypes.byref(self._h)))
        self.nxl = lib().lsc_mhd_local_nx(self._h)
        self.i0 = lib().lsc_mhd_local_i0(self._h)
        # "reference": bit-for-bit the reference's IEEE arithmetic (default, the parity path);
        # "relaxed": FMA-contracted sweeps with 1-multiply quotients (include/lsc_b200.h: LSC_ARITH_RELAXED)
        self.arithmetic = arithmetic
        check(lib().lsc_mhd_set_arithmetic(self._h, {"reference": 0, "relaxed": 1}[arithmetic]))

    # -- lifetime ---------------------------------------------------------
    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            lib().lsc_mhd_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # -- state ------------------------------------------------------------
    @property
    def shape(self):
        """Host AoS shape of this rank's slab, ghosts included (MUSCL_NEW.cpp:673-674)."""
        return (self.nxl + 4, self.cfg.ny + 4, self.NVAR)

    def initial_condition(self, problem=None):
        """The IC loop MUSCL_NEW.cpp:719-825 for this slab, as a host array."""
        u = np.empty(self.shape, dtype=np.float64)
        check(lib().lsc_mhd_initial_condition(self._h, self.problem if problem is None else
                                              (_capi.MHD_PROBLEMS[problem] if isinstance(problem, str) else problem),
                                              _dp(u)))
        return u

    def upload(self, u):
        u = np.ascontiguousarray(u, dtype=np.float64)
        if u.shape != self.shape:
            raise ValueError(f"state must have shape {self.shape}, got {u.shape}")
        check(lib().lsc_mhd_upload(self._h, _dp(u)))

    def init_problem(self, problem=None):
        p = self.problem if problem is None else (_capi.MHD_PROBLEMS[problem] if isinstance(problem, str) else problem)
        check(lib().lsc_mhd_init_problem(self._h, p))

    def download(self, out=None):
        u = np.empty(self.shape, dtype=np.float64) if out is None else out
        if not (isinstance(u, np.ndarray) and u.dtype == np.float64 and u.shape == self.shape
                and u.flags["C_CONTIGUOUS"] and u.flags["WRITEABLE"]):
            raise ValueError(f"out must be a writeable C-contiguous float64 array of shape {self.shape}")
        check(lib().lsc_mhd_download(self._h, _dp(u)))
        return u

    def state_hash(self):
        """This slab's word of the decomposition-independent interior fingerprint (lsc_mhd_state_hash); the
        fingerprint of the whole grid is the sum of the slabs' words modulo 2^64."""
        h = ctypes.c_ulonglong()
        check(lib().lsc_mhd_state_hash(self._h, ctypes.byref(h)))
        return h.value

    def state_hash_noalloc(self):
        """state_hash() without any device allocation on the call path (lsc_mhd_state_hash_noalloc: scratch in a
        static device array, ny <= 1048576, calls on one device must not overlap)."""
        h = ctypes.c_ulonglong()
        check(lib().lsc_mhd_state_hash_noalloc(self._h, ctypes.byref(h)))
        return h.value

    def comm_init(self, unique_id: bytes):
        buf = ctypes.create_string_buffer(unique_id, 128)
        check(lib().lsc_mhd_comm_init(self._h, ctypes.cast(buf, ctypes.c_void_p)))

    # -- time loop --------------------------------------------------------
    def step(self, nsteps=1):
        check(lib().lsc_mhd_step(self._h, int(nsteps)))

    def step_timed(self, nsteps=1):
        """step(nsteps) bracketed by CUDA events on the solver's stream; returns device milliseconds."""
        ms = ctypes.c_float()
        check(lib().lsc_mhd_step_timed(self._h, int(nsteps), ctypes.byref(ms)))
        return ms.value

    def set_overlap(self, mode=1):
        """0: y sweep, then halo exchange; 1 (default): exchange overlapped with the interior of the y sweep when
        the slab has NCCL neighbours; 2: split the y sweep on a single rank too (tests)."""
        check(lib().lsc_mhd_set_overlap(self._h, int(mode)))

    def set_profiling(self, enabled=True):
        check(lib().lsc_mhd_set_profiling(self._h, 1 if enabled else 0))

    def kernel_times(self):
        """(ms_x_sweeps, ms_y_sweeps, n_x, n_y) accumulated since the last call (profiling mode)."""
        ms = (ctypes.c_double * 2)()
        n = (ctypes.c_long * 2)()
        check(lib().lsc_mhd_kernel_times(self._h, ms, n))
        return ms[0], ms[1], n[0], n[1]

    def comm_times(self):
        """(halo exchange ms, allreduce ms, halo calls, allreduce calls) since the last call (profiling, nranks > 1)"""
        ms = (ctypes.c_double * 2)()
        n = (ctypes.c_long * 2)()
        check(lib().lsc_mhd_comm_times(self._h, ms, n))
        return ms[0], ms[1], n[0], n[1]

    def ieee_fallback_count(self):
        """Steps recomputed with IEEE division because a fast division flagged an operand."""
        return lib().lsc_mhd_ieee_fallback_count(self._h)

    def scalars(self):
        t, dt, ch = ctypes.c_double(), ctypes.c_double(), ctypes.c_double()
        n = ctypes.c_long()
        check(lib().lsc_mhd_get_scalars(self._h, ctypes.byref(t), ctypes.byref(dt), ctypes.byref(ch),
                                        ctypes.byref(n)))
        return t.value, dt.value, ch.value, n.value

    def loop_stats(self):
        """(host synchronisations of the time loop so far, psi-damping exp-table misses) -- lsc_mhd_loop_stats"""
        a, b = ctypes.c_long(), ctypes.c_long()
        check(lib().lsc_mhd_loop_stats(self._h, ctypes.byref(a), ctypes.byref(b)))
        return a.value, b.value

    def _run_until(self, tStop, dump_cb, log_cb, n):
        return lib().lsc_mhd_run_until(self._h, tStop, dump_cb, log_cb, None, ctypes.byref(n))

    def run(self, tStop=None, outdir=None, log=None):
        """The do{}while loop MUSCL_NEW.cpp:831-1003 followed by the final dump :1006-1032.
        Writes MHD<counter>.dat into `outdir` at the reference's cadence when given; `log`
        receives (step, t, dt, ch) per step.  Returns the number of steps taken."""
        tStop = self.tStop if tStop is None else float(tStop)
        state = {"counter": 0}
        buf = np.empty(self.shape, dtype=np.float64) if outdir else None

        # an exception inside a ctypes callback would be printed and swallowed: keep it, make the C loop stop
        # (non-zero return from the dump callback aborts lsc_mhd_run_until), re-raise once the call has returned
        def on_dump(_ctx, counter, t, dt, _user):
            if "error" in state:
                return -1
            try:
                if outdir:
                    self.download(buf)
                    self.write_dat(buf, os.path.join(outdir, f"MHD{counter}.dat"))
                state["counter"] = counter + 1
                return 0
            except BaseException as e:  # noqa: BLE001
                state["error"] = e
                return -1

        def on_log(step, t, dt, ch, _user):
            if "error" in state:
                return 1
            try:
                log(step, t, dt, ch)
                return 0
            except BaseException as e:  # noqa: BLE001
                state["error"] = e
                return 1

        dump_cb = _capi.DUMP_CB(on_dump)
        log_cb = _capi.LOG_CB(on_log) if log else ctypes.cast(None, _capi.LOG_CB)
        n = ctypes.c_long()
        rc = self._run_until(tStop, dump_cb, log_cb, n)
        if "error" in state:
            raise state["error"]
        check(rc)
        if outdir:
            self.download(buf)
            self.write_dat(buf, os.path.join(outdir, f"MHD{state['counter']}.dat"))
        return n.value

    def write_dat(self, u, path):
        """MUSCL_NEW.cpp:989-1001 -- byte-compatible .dat writer (full-grid arrays only)."""
        if self.cfg.nranks != 1:
            raise ValueError("write_dat needs the assembled full-grid array")
        u = np.ascontiguousarray(u, dtype=np.float64)
        check(lib().lsc_mhd_write_dat(ctypes.byref(self.cfg), _dp(u), os.fsencode(path)))


def slab_range(nx, rank, nranks):
    """First global interior column and width of rank's slab (same arithmetic as lsc_mhd_create)."""
    b0 = rank * nx // nranks
    b1 = (rank + 1) * nx // nranks
    return b0, b1 - b0


def neighbours(rank, nranks, bc_x):
    """(left, right) neighbour ranks of a slab; -1 at a non-periodic physical boundary."""
    ring = bc_x in ("periodic", _capi.BC_PERIODIC)
    left = rank - 1 if rank > 0 else (nranks - 1 if ring else -1)
    right = rank + 1 if rank < nranks - 1 else (0 if ring else -1)
    return left, right


class MusclNewGroup:
    """P slabs of one grid driven in lock step from this process (lsc_mhd_group_step): on one GPU for
    decomposition-invariance tests, or one slab per GPU with peer copies over NVLink."""

    def __init__(self, nslabs, devices=None, _parts=None, **kw):
        ndev = lib().lsc_device_count()
        devices = devices if devices is not None else [r % max(ndev, 1) for r in range(nslabs)]
        self.parts = _parts if _parts is not None else \
            [MusclNew(device=devices[r], rank=r, nranks=nslabs, **kw) for r in range(nslabs)]
        self.cfg = MhdConfig.from_buffer_copy(self.parts[0].cfg)
        self.cfg.rank, self.cfg.nranks, self.cfg.slab_i0, self.cfg.slab_nx = 0, 1, 0, 0
        self._arr = (ctypes.c_void_p * nslabs)(*[p._h for p in self.parts])

    @classmethod
    def from_parts(cls, parts):
        """a team of contexts the caller created itself (e.g. with explicit, uneven slab bounds)"""
        return cls(len(parts), _parts=list(parts))

    @property
    def shape(self):
        return (self.cfg.nx + 4, self.cfg.ny + 4, 9)

    def upload(self, u):
        u = np.ascontiguousarray(u, dtype=np.float64)
        if u.shape != self.shape:
            raise ValueError(f"state must have shape {self.shape}")
        for p in self.parts:
            p.upload(u[p.i0:p.i0 + p.nxl + 4])

    def init_problem(self, problem=None):
        for p in self.parts:
            p.init_problem(problem)

    def step(self, nsteps=1):
        check(lib().lsc_mhd_group_step(self._arr, len(self.parts), int(nsteps)))

    def step_timed(self, nsteps=1):
        ms = ctypes.c_float()
        check(lib().lsc_mhd_group_step_timed(self._arr, len(self.parts), int(nsteps), ctypes.byref(ms)))
        return ms.value

    def scalars(self):
        return self.parts[0].scalars()

    def loop_stats(self):
        return self.parts[0].loop_stats()

    def ieee_fallback_count(self):
        return sum(p.ieee_fallback_count() for p in self.parts)

    def run(self, tStop=12.0, outdir=None, log=None):
        """MusclNew.run for the slabs of this process: the loop M:831-1003 through lsc_mhd_group_run_until, the
        MHD<counter>.dat files from the assembled grid."""
        state = {"counter": 0}

        def on_dump(_ctx, counter, t, dt, _user):
            if "error" in state:
                return -1
            try:
                if outdir:
                    self.write_dat(self.download(), os.path.join(outdir, f"MHD{counter}.dat"))
                state["counter"] = counter + 1
                return 0
            except BaseException as e:  # noqa: BLE001
                state["error"] = e
                return -1

        def on_log(step, t, dt, ch, _user):
            if "error" in state:
                return 1
            try:
                log(step, t, dt, ch)
                return 0
            except BaseException as e:  # noqa: BLE001
                state["error"] = e
                return 1

        dump_cb = _capi.DUMP_CB(on_dump)
        log_cb = _capi.LOG_CB(on_log) if log else ctypes.cast(None, _capi.LOG_CB)
        n = ctypes.c_long()
        rc = lib().lsc_mhd_group_run_until(self._arr, len(self.parts), float(tStop), dump_cb, log_cb, None,
                                           ctypes.byref(n))
        if "error" in state:
            raise state["error"]
        check(rc)
        if outdir:
            self.write_dat(self.download(), os.path.join(outdir, f"MHD{state['counter']}.dat"))
        return n.value

    def write_dat(self, u, path):
        u = np.ascontiguousarray(u, dtype=np.float64)
        check(lib().lsc_mhd_write_dat(ctypes.byref(self.cfg), _dp(u), os.fsencode(path)))

    def state_hash(self):
        return sum(p.state_hash() for p in self.parts) & ((1 << 64) - 1)

    def download(self, out=None):
        """The full-grid array in the reference layout, ghost ring rebuilt as the reference holds it
        (lsc_mhd_group_download)."""
        u = np.zeros(self.shape, dtype=np.float64) if out is None else out
        check(lib().lsc_mhd_group_download(self._arr, len(self.parts), _dp(u)))
        return u

    def close(self):
        for p in self.parts:
            p.close()


def kat_hllc(L, R, gamma, ch, direction, device=0):
    """Device HLLC flux of n interface state pairs (test hook)."""
    L = np.ascontiguousarray(L, dtype=np.float64)
    R = np.ascontiguousarray(R, dtype=np.float64)
    F = np.empty_like(L)
    check(lib().lsc_mhd_kat_hllc(device, int(direction), L.shape[0], _dp(L), _dp(R), float(gamma), float(ch), _dp(F)))
    return F


def kat_cfl_bound(U, gamma, run_max, device=0):
    """(exact fp64 signal speed, fp32 upper bound of speed/run_max used by the y sweep's warp filter) per state."""
    U = np.ascontiguousarray(U, dtype=np.float64)
    out = np.empty((U.shape[0], 2), dtype=np.float64)
    check(lib().lsc_mhd_kat_cfl_bound(device, U.shape[0], _dp(U), float(gamma), float(run_max), _dp(out)))
    return out[:, 0], out[:, 1]


def kat_fluxdef(U, gamma, ch, direction, device=0):
    U = np.ascontiguousarray(U, dtype=np.float64)
    F = np.empty_like(U)
    check(lib().lsc_mhd_kat_fluxdef(device, int(direction), U.shape[0], _dp(U), float(gamma), float(ch), _dp(F)))
    return F
