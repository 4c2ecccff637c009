"""Host-side mirror of the reference program MHD_codes/HLLC_MHD_2D.cpp (H): first-order Godunov + HLLC + mixed GLM
on the 9-component MHD system, ONE ghost cell, periodic, Orszag-Tang.  Like the other
programs of the reference it has no callable API; `HllcMhd2d` exposes the constants of its main() (H:616-627) with the
committed values as defaults, the log line (H:738) and `MHD.dat` (H:793-814), and drives the CUDA library through
its C ABI (lsc_hllc2d_*); it holds no numerical code of its own."""
import ctypes
import os

import numpy as np

from . import _capi
from ._capi import Hllc2dConfig, check, lib

_DP = ctypes.POINTER(ctypes.c_double)


def _dp(a):
    return a.ctypes.data_as(_DP)


class HllcMhd2d:
    NVAR = 9

    def __init__(self, nxCells=800, nyCells=800, x0=0.0, x1=1.0, y0=0.0, y1=1.0, tStop=0.5, C=0.75, gamma=5.0 / 3.0,
                 device=0):
        self.tStop = float(tStop)
        self.cfg = Hllc2dConfig(int(nxCells), int(nyCells), float(x0), float(x1), float(y0), float(y1), float(gamma),
                                float(C), int(device))
        self._h = ctypes.c_void_p()
        check(lib().lsc_hllc2d_create(ctypes.byref(self.cfg), ctypes.byref(self._h)))

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            lib().lsc_hllc2d_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def shape(self):
        """host array with its ONE ghost layer (H:630-631)"""
        return (self.cfg.nx + 2, self.cfg.ny + 2, self.NVAR)

    def initial_condition(self):
        u = np.empty(self.shape, dtype=np.float64)
        check(lib().lsc_hllc2d_initial_condition(ctypes.byref(self.cfg), _dp(u)))
        return u

    def upload(self, u):
        u = np.ascontiguousarray(u, dtype=np.float64)
        if u.shape != self.shape:
            raise ValueError(f"state must have shape {self.shape}, got {u.shape}")
        check(lib().lsc_hllc2d_upload(self._h, _dp(u)))

    def init_problem(self):
        check(lib().lsc_hllc2d_init_problem(self._h))

    def step(self, nsteps=1):
        check(lib().lsc_hllc2d_step(self._h, int(nsteps)))

    def scalars(self):
        t, dt, ch = ctypes.c_double(), ctypes.c_double(), ctypes.c_double()
        n = ctypes.c_long()
        check(lib().lsc_hllc2d_get_scalars(self._h, ctypes.byref(t), ctypes.byref(dt), ctypes.byref(ch), ctypes.byref(n)))
        return t.value, dt.value, ch.value, n.value

    def download(self, out=None):
        u = np.empty(self.shape, dtype=np.float64) if out is None else out
        if not (isinstance(u, np.ndarray) and u.dtype == np.float64 and u.shape == self.shape
                and u.flags["C_CONTIGUOUS"] and u.flags["WRITEABLE"]):
            raise ValueError(f"out must be a writeable C-contiguous float64 array of shape {self.shape}")
        check(lib().lsc_hllc2d_download(self._h, _dp(u)))
        return u

    def run(self, tStop=None, outdir=None, log=None):
        """main() of H from H:727: do{}while(t < tStop), then MHD.dat.  Returns the number of steps."""
        tStop = self.tStop if tStop is None else float(tStop)
        err = []

        def on_log(step, t, dt, ch, _u):
            try:
                log(step, t, dt, ch)
                return 0
            except BaseException as e:  # noqa: BLE001
                err.append(e)
                return 1

        cb = _capi.LOG_CB(on_log) if log else ctypes.cast(None, _capi.LOG_CB)
        n = ctypes.c_long()
        rc = lib().lsc_hllc2d_run_until(self._h, tStop, cb, None, ctypes.byref(n))
        if err:
            raise err[0]
        check(rc)
        if outdir:
            self.write_dat(self.download(), os.path.join(outdir, "MHD.dat"))
        return n.value

    def write_dat(self, u, path):
        u = np.ascontiguousarray(u, dtype=np.float64)
        check(lib().lsc_hllc2d_write_dat(ctypes.byref(self.cfg), _dp(u), os.fsencode(path)))
