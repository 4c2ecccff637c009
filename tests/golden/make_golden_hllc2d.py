#!/usr/bin/env python3
"""Fixtures of the first-order sibling scheme (MHD_codes/HLLC_MHD_2D.cpp) from the REAL program
(oracle/_ref/hllc2d_ref, oracle/build_ref_hllc2d.py).  Run in the build container.
  tests/golden/hllc2d_<case>.npz     state after the program's final boundary fill + per-step (t, dt, ch) log
  tests/golden/hllc2d_program_64.npz the program run like the reference (64 x 64 to t = 0.1): stdout, sha256 of MHD.dat
"""
import hashlib
import json
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from oracle_util import run_hllc2d_ref  # noqa: E402

DOM = (0.0, 1.0, 0.0, 1.0)
CASES = {"ot32_s1": (32, 32, 1), "ot32_s12": (32, 32, 12), "ot50x38_s25": (50, 38, 25), "ot96x64_s40": (96, 64, 40),
         "ot3x5_s4": (3, 5, 4)}


def main():
    for name, (nx, ny, n) in CASES.items():
        u, log, _, _ = run_hllc2d_ref(nx, ny, DOM, 1e9, 0.75, n)
        np.savez_compressed(os.path.join(HERE, f"hllc2d_{name}.npz"), u=u, log=log,
                            meta=json.dumps(dict(nx=nx, ny=ny, dom=DOM, nsteps=n, cfl=0.75, gamma="5/3")))
        print(name, u.shape, log[-1])
    with tempfile.TemporaryDirectory() as td:
        u, log, out, _ = run_hllc2d_ref(64, 64, DOM, 0.1, 0.75, -1, cwd=td, writedat=True)
        with open(os.path.join(td, "MHD.dat"), "rb") as f:
            sha = hashlib.sha256(f.read()).hexdigest()
    np.savez_compressed(os.path.join(HERE, "hllc2d_program_64.npz"), u=u, log=log, stdout=out, dat_sha256=sha)
    print("program 64x64 to t=0.1:", len(log), "steps", sha[:16])


if __name__ == "__main__":
    main()
