#!/usr/bin/env python3
"""Fixtures for the BASELINE.json configurations at the sizes / end times they name, from the REAL reference
(oracle/_ref/*, built from /root/reference by oracle/build_ref*.py).  Run in the build container.

The states are too large to commit (76 MB ... 540 MB), so a fixture holds
  sha256   of the raw fp64 state the reference dumps (ghost ring included), and of its interior alone,
  log      the per-step (t, dt, ch) / (t, dt) records,
  rho      (OT 256^2, t = 0.5 only) the interior density field, for the L1(rho) <= 1e-9 gate of BASELINE.json.
tests/test_named_sizes_gpu.py compares the CUDA path with these and, when oracle/_ref is present on the box,
with the reference run live.

  C1  Orszag-Tang 256^2 periodic to t = 0.5 (407 steps)                      muscl_ref
  C2  Kelvin-Helmholtz 1024^2, [0,1]x[-1,1], periodic-x / reflective-y, 10 steps   muscl_ref
  C4' Orszag-Tang 2048^2 periodic, 3 steps (stand-in for 16384^2)  muscl_ref
  C3  Euler cylindrical explosion 4096^2 transmissive, 2 steps               gslic_ref
"""
import hashlib
import json
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from oracle_util import run_euler_ref, run_muscl_ref  # noqa: E402

MHD_LARGE = {
    # name: (prob, nx, ny, domain, bcx, bcy, tStop, nsteps (-1 = until tStop))
    "c1_ot256_t05": ("ot", 256, 256, (0.0, 1.0, 0.0, 1.0), "periodic", "periodic", 0.5, -1),
    "c2_kh1024_s10": ("kh", 1024, 1024, (0.0, 1.0, -1.0, 1.0), "periodic", "reflective", 1e9, 10),
    "c4_ot2048_s3": ("ot", 2048, 2048, (0.0, 1.0, 0.0, 1.0), "periodic", "periodic", 1e9, 3),
}
EULER_LARGE = {
    # name: (binary, bc, ic name, ic code, nx, ny, nsteps)
    "c3_cylinder4096_s2": ("gslic_ref", "transmissive", "cylinder", 1, 4096, 4096, 2),
}


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def main(only=None):
    for name, (prob, nx, ny, dom, bcx, bcy, tstop, n) in MHD_LARGE.items():
        if only and name not in only:
            continue
        t0 = time.time()
        u, log, err = run_muscl_ref(prob, nx, ny, dom, tstop, 0.8, bcx, bcy, n)
        meta = dict(prob=prob, nx=nx, ny=ny, dom=dom, bcx=bcx, bcy=bcy, tstop=tstop, nsteps=int(log.shape[0]), cfl=0.8,
                    gamma="5/3", sha256_state=sha(u), sha256_interior=sha(u[2:-2, 2:-2]), ref_stderr=err.strip())
        extra = {}
        if name.startswith("c1_"):
            extra["rho"] = u[2:-2, 2:-2, 0].copy()
        np.savez_compressed(os.path.join(HERE, f"large_mhd_{name}.npz"), log=log, meta=json.dumps(meta), **extra)
        print(name, u.shape, log.shape, log[-1], f"{time.time() - t0:.1f}s", flush=True)
    for name, (exe, bc, ic, code, nx, ny, n) in EULER_LARGE.items():
        if only and name not in only:
            continue
        t0 = time.time()
        u, log, err = run_euler_ref(exe, nx, ny, (0.0, 1.0, 0.0, 1.0), 1e9, 0.8, 1.4, n, ic=code)
        meta = dict(exe=exe, bc=bc, ic=ic, nx=nx, ny=ny, nsteps=n, cfl=0.8, gamma=1.4, sha256_state=sha(u),
                    sha256_interior=sha(u[2:-2, 2:-2]), ref_stderr=err.strip())
        np.savez_compressed(os.path.join(HERE, f"large_euler_{name}.npz"), log=log, meta=json.dumps(meta))
        print(name, u.shape, log[-1], f"{time.time() - t0:.1f}s", flush=True)


if __name__ == "__main__":
    main(set(sys.argv[1:]) or None)
