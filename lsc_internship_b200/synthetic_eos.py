"""A synthetic equation-of-state table in the file layout MHD_codes/TabEoS.cpp parses (T:36-128).

The reference's own `Plasma19_EoS.txt` is a stripped large blob (not in the reference tree), so every run of the
tabulated-EoS path -- the parity tests, the drop-in driver's fixtures, bench.py's C5 workload -- uses this stand-in:
line 10 = 500 densities, line 12 = 500 pressures, lines 14-513 sound speed, 515-1014 specific internal energy,
1016-1515 temperature, 1517-2016 electrical and 2018-2517 thermal conductivity; rows = pressure index, columns =
density index.  A gas whose adiabatic index drifts from 1.6 (low p) to 1.2 (high p): not an ideal gas (so the lookups
are really exercised), e(p) strictly increasing at fixed rho (T's reverse search needs that)."""
import numpy as np


def eos_table():
    dens = np.logspace(-3.0, 2.0, 500)
    pres = np.logspace(1.0, 10.0, 500)
    g = 1.4 - 0.2 * np.tanh(np.log10(pres / 1.0e6))
    P, R = pres[:, None], dens[None, :]
    energies = P / ((g[:, None] - 1.0) * R)
    sound = np.sqrt(g[:, None] * P / R)
    return dens, pres, sound, energies


def write_eos_table(path):
    """Write the synthetic table as a Plasma19_EoS.txt look-alike; returns the arrays that were written."""
    dens, pres, sound, energies = eos_table()
    temp = energies / 718.0
    row = lambda v: " ".join(repr(float(x)) for x in v)  # noqa: E731
    lines = ["# synthetic EoS table in the layout TabEoS.cpp parses (T:36-128)"] * 9
    lines += [row(dens), "# pressures", row(pres), "# sound speed [p][rho]"]
    lines += [row(r) for r in sound] + ["# specific internal energy [p][rho]"]
    lines += [row(r) for r in energies] + ["# temperature"]
    lines += [row(r) for r in temp] + ["# electrical conductivity"]
    lines += [row(r) for r in np.ones_like(temp)] + ["# thermal conductivity"]
    lines += [row(r) for r in np.ones_like(temp)]
    assert len(lines) == 2517
    with open(path, "w") as f:
        f.write("\n".join(lines) + "\n")
    return dens, pres, sound, energies
