#!/usr/bin/env python3
"""Generate the committed golden fixtures from the REAL reference.

Run in the build container (needs /root/reference and oracle/_ref/muscl_ref built by
oracle/build_ref.py).  Writes
  tests/golden/mhd_<case>.npz      final conserved state incl. ghost ring after N steps of the
                                   reference program + its per-step (t, dt, ch) log
  tests/golden/dat_sha256.json     sha256 of the reference's own committed output files
                                   MHD_plot_data/MHD{0..121}.dat and euler_{1..20}.dat
The GPU box has no /root/reference; the tests there read only these fixtures.
"""
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from oracle_util import run_euler_ref, run_muscl_ref  # noqa: E402

REF = "/root/reference"

MHD_CASES = {
    # name: (prob, nx, ny, domain, bcx, bcy, nsteps)
    "ot32_s1": ("ot", 32, 32, (0.0, 1.0, 0.0, 1.0), "periodic", "periodic", 1),
    "ot32_s10": ("ot", 32, 32, (0.0, 1.0, 0.0, 1.0), "periodic", "periodic", 10),
    "ot64_s40": ("ot", 64, 64, (0.0, 1.0, 0.0, 1.0), "periodic", "periodic", 40),
    "kh32x64_s10": ("kh", 32, 64, (0.0, 1.0, -1.0, 1.0), "periodic", "reflective", 10),
    "kh48x40_s25": ("kh", 48, 40, (0.0, 1.0, -1.0, 1.0), "periodic", "reflective", 25),
    "ot40x24_trans_s12": ("ot", 40, 24, (0.0, 1.0, 0.0, 1.0), "transmissive", "transmissive", 12),
    "kh36x52_trans_per_s8": ("kh", 36, 52, (0.0, 1.0, -1.0, 1.0), "transmissive", "periodic", 8),
}


EULER_CASES = {
    # name: (binary, bc, ic name, ic code for the binary, nx, ny, nsteps)
    "slic_ot40x32_s12": ("slic_ref", "periodic", "ot", 0, 40, 32, 12),
    "slic_ot100_s30": ("slic_ref", "periodic", "ot", 0, 100, 100, 30),
    "gslic_quadrants36x44_s15": ("gslic_ref", "transmissive", "quadrants", 0, 36, 44, 15),
    "gslic_cylinder64_s25": ("gslic_ref", "transmissive", "cylinder", 1, 64, 64, 25),
}


def main():
    for name, (exe, bc, ic, code, nx, ny, n) in EULER_CASES.items():
        u, log, _ = run_euler_ref(exe, nx, ny, (0.0, 1.0, 0.0, 1.0), 1e9, 0.8, 1.4, n, ic=code)
        np.savez_compressed(os.path.join(HERE, f"euler_{name}.npz"), u=u, log=log,
                            meta=json.dumps(dict(exe=exe, bc=bc, ic=ic, nx=nx, ny=ny, nsteps=n, cfl=0.8, gamma=1.4)))
        print(name, u.shape, log[-1])
    for name, (prob, nx, ny, dom, bcx, bcy, n) in MHD_CASES.items():
        u, log, _ = run_muscl_ref(prob, nx, ny, dom, 1e9, 0.8, bcx, bcy, n)
        np.savez_compressed(os.path.join(HERE, f"mhd_{name}.npz"), u=u, log=log,
                            meta=json.dumps(dict(prob=prob, nx=nx, ny=ny, dom=dom, bcx=bcx, bcy=bcy, nsteps=n,
                                                 cfl=0.8, gamma="5/3")))
        print(name, u.shape, log[-1])
    sha = {}
    for f in [f"MHD{k}.dat" for k in range(122)] + [f"euler_{k}.dat" for k in range(1, 21)]:
        with open(os.path.join(REF, "MHD_plot_data", f), "rb") as fh:
            sha[f] = hashlib.sha256(fh.read()).hexdigest()
    # the reference program itself (committed config) run to tStop = 0.25: its stdout log and final dump
    import subprocess
    import tempfile
    from oracle_util import REF_DIR
    with tempfile.TemporaryDirectory() as td:
        r = subprocess.run([os.path.join(REF_DIR, "muscl_ref"), "kh", "128", "256", "0", "1", "-1", "1", "0.25", "0.8",
                            "periodic", "reflective", "-1", "1"], cwd=td, check=True, capture_output=True, text=True)
        with open(os.path.join(HERE, "muscl_new_stdout_t025.txt"), "w") as fh:
            fh.write(r.stdout)
        with open(os.path.join(td, "MHD3.dat"), "rb") as fh:
            sha["MHD3.dat@tStop=0.25"] = hashlib.sha256(fh.read()).hexdigest()
    with tempfile.TemporaryDirectory() as td:
        r = subprocess.run([os.path.join(REF_DIR, "slic_ref"), "100", "100", "0", "1", "0", "1", "0.5", "0.8", "1.4", "-1",
                            "1"], cwd=td, check=True, capture_output=True, text=True)
        with open(os.path.join(HERE, "slic_euler_2d_stdout.txt"), "w") as fh:
            fh.write(r.stdout)
    # generaliseSLIC.cpp as committed (no golden output exists in the reference repo for it: Euler_codes/SLIC2d.dat
    # was written by TabEoS.cpp) -- record what the real program writes here
    with tempfile.TemporaryDirectory() as td:
        r = subprocess.run([os.path.join(REF_DIR, "gslic_ref"), "100", "100", "0", "1", "0", "1", "0.5", "0.8", "1.4",
                            "-1", "1"], cwd=td, check=True, capture_output=True, text=True)
        with open(os.path.join(HERE, "generalise_slic_stdout.txt"), "w") as fh:
            fh.write(r.stdout)
        with open(os.path.join(td, "SLIC2d.dat"), "rb") as fh:
            sha["SLIC2d.dat@generaliseSLIC"] = hashlib.sha256(fh.read()).hexdigest()
    with open(os.path.join(HERE, "dat_sha256.json"), "w") as fh:
        json.dump(sha, fh, indent=1, sort_keys=True)
    print(sha)


if __name__ == "__main__":
    main()
