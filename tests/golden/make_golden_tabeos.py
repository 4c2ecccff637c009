#!/usr/bin/env python3
"""Generate the tabulated-EoS golden fixtures from the REAL reference program MHD_codes/TabEoS.cpp.

Run in the build container (needs /root/reference and oracle/_ref/tabeos_ref built by
oracle/build_ref_tabeos.py).  The real Plasma19_EoS.txt is not in the reference tree, so the program runs on the
synthetic table of tests/oracle_util.py::write_eos_table (written next to it as ./Plasma19_EoS.txt).  Writes

  tests/golden/tabeos_<case>.npz   conserved state incl. ghost ring after N steps + the per-step (t, dt) log
  tests/golden/tabeos_program_48x40.npz
                                   the complete program at 48x40 (reference tStop): its "t = .. dt = .." stdout
                                   lines and its SLIC2d.dat, for the drop-in test of drivers/tab_eos
The GPU box has no /root/reference; the tests there read only these fixtures.
"""
import os
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from oracle_util import REF_DIR, run_tabeos_ref, write_eos_table  # noqa: E402

CASES = {"40x40_s6": (40, 40, 6), "33x21_s4": (33, 21, 4), "130x70_s3": (130, 70, 3)}

if __name__ == "__main__":
    for name, (nx, ny, n) in CASES.items():
        u, lg = run_tabeos_ref(nx, ny, n)
        np.savez_compressed(os.path.join(HERE, f"tabeos_{name}.npz"), u=u, log=lg, nx=nx, ny=ny, nsteps=n)
        print("wrote", name, lg[-1])
    with tempfile.TemporaryDirectory() as td:
        write_eos_table(os.path.join(td, "Plasma19_EoS.txt"))
        t_stop = repr(0.012 / 10 ** 2.5)
        r = subprocess.run([os.path.join(REF_DIR, "tabeos_ref"), "48", "40", "0.0", "1.0", "0.0", "1.0", t_stop, "0.8",
                            "-1", "1"], cwd=td, check=True, capture_output=True, text=True)
        dat = open(os.path.join(td, "SLIC2d.dat"), "rb").read()
    np.savez_compressed(os.path.join(HERE, "tabeos_program_48x40.npz"), stdout=np.frombuffer(r.stdout.encode(), dtype=np.uint8),
                        dat=np.frombuffer(dat, dtype=np.uint8))
    print("program: %d stdout lines, SLIC2d.dat %d bytes" % (r.stdout.count("\n"), len(dat)))
