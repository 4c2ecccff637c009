#!/usr/bin/env python3
"""Build the REAL reference programs from /root/reference into oracle/_ref/.

TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is on the product path; only
tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may execute what this script produces.

The reference (JosephineMellor/LSC_internship) is configured by editing source:
grid, initial condition, boundary condition and end time are constants and
commented-in/out blocks inside main().  This script reads the reference
sources *where they lie* under /root/reference, applies a handful of exact
string substitutions IN MEMORY (each asserted to match exactly once), pipes the
patched translation unit to g++ on stdin, and keeps only the resulting binary
in oracle/_ref/ (git-ignored).  No reference source text is written to disk.

The substitutions do not touch any arithmetic.  They only
  (1) read grid/domain/tStop/C from argv instead of literals   (MUSCL_NEW.cpp:659-667)
  (2) select the IC block at run time (KH :800-808 | OT :811-819)
  (3) select the BC loops at run time (:609-614 | :617-622 ; :625-638 | :641-646 | :649-654)
  (4) optionally bound the loop by a step count instead of t<tStop (:1003)
  (5) optionally suppress the .dat writers (:975, :1023-1032)
  (6) append a raw fp64 dump of the conserved state and a per-step (t,dt,ch) log
  (7) time the loop with steady_clock and print it on stderr.

Build flags: g++ -std=c++17 -O3 (-O0/-O2/-O3 agree bit for bit), no -march (FMA contraction changes the
answer).
"""
import os
import subprocess
import sys

REF = os.environ.get("LSC_REFERENCE_ROOT", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")


def sub1(src, old, new, count=1):
    n = src.count(old)
    assert n == count, f"expected {count} match(es), found {n}: {old[:60]!r}"
    return src.replace(old, new)


# --------------------------------------------------------------------------
# MHD_codes/MUSCL_NEW.cpp
# --------------------------------------------------------------------------
MHD_PROLOGUE = r"""
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
static int g_prob = 0;      /* 0 = Kelvin-Helmholtz (committed), 1 = Orszag-Tang */
static int g_bcx = 0;       /* 0 periodic (committed), 1 transmissive */
static int g_bcy = 0;       /* 0 reflective (committed), 1 periodic, 2 transmissive */
static long g_nsteps = -1;  /* <0: run until t>=tStop as the reference does */
static long g_stepcount = 0;
static int g_writedat = 1;
static FILE* g_logf = nullptr;
"""

MHD_USAGE = ("usage: muscl_ref prob(kh|ot) nx ny x0 x1 y0 y1 tStop C bcx(periodic|transmissive) "
             "bcy(reflective|periodic|transmissive) nsteps(-1=until tStop) writedat(0|1) "
             "[state_dump.bin [steplog.bin]]")


def patch_mhd(src):
    src = sub1(src, "#include <tuple>\n", "#include <tuple>\n" + MHD_PROLOGUE)
    # ---- (3) boundary conditions, run-time selectable ----
    src = sub1(src,
               "    for (int j = 0; j < nyCells + 2; ++j) {\n        u[0][j] = u[nxCells][j]; \n",
               "    if (g_bcx == 0) for (int j = 0; j < nyCells + 2; ++j) {\n        u[0][j] = u[nxCells][j]; \n")
    src = sub1(src,
               "    // for (int j = 0; j < nyCells + 2; ++j) {\n"
               "    //     u[0][j] = u[2][j]; \n"
               "    //     u[1][j] = u[3][j];      // Left boundary\n"
               "    //     u[nxCells + 2][j] = u[nxCells][j]; \n"
               "    //     u[nxCells + 3][j] = u[nxCells + 1][j];  // Right boundary\n"
               "    // }\n",
               "    if (g_bcx == 1) for (int j = 0; j < nyCells + 2; ++j) {\n"
               "        u[0][j] = u[2][j]; \n"
               "        u[1][j] = u[3][j];      // Left boundary\n"
               "        u[nxCells + 2][j] = u[nxCells][j]; \n"
               "        u[nxCells + 3][j] = u[nxCells + 1][j];  // Right boundary\n"
               "    }\n")
    src = sub1(src,
               "    for (int i = 0; i < nxCells + 2; ++i) {//reflect in u_y and B_y\n",
               "    if (g_bcy == 0) for (int i = 0; i < nxCells + 2; ++i) {//reflect in u_y and B_y\n")
    src = sub1(src,
               "    // for (int i = 0; i < nxCells + 2; ++i) {\n"
               "    //     u[i][0] = u[i][nyCells];      // Bottom boundary\n"
               "    //     u[i][1] = u[i][nyCells+1];    \n"
               "    //     u[i][nyCells + 2] = u[i][2];  // Top boundary\n"
               "    //     u[i][nyCells + 3] = u[i][3];  \n"
               "    // }\n",
               "    if (g_bcy == 1) for (int i = 0; i < nxCells + 2; ++i) {\n"
               "        u[i][0] = u[i][nyCells];      // Bottom boundary\n"
               "        u[i][1] = u[i][nyCells+1];    \n"
               "        u[i][nyCells + 2] = u[i][2];  // Top boundary\n"
               "        u[i][nyCells + 3] = u[i][3];  \n"
               "    }\n")
    src = sub1(src,
               "    // for (int i = 0; i < nxCells + 2; ++i) {\n"
               "    //     u[i][0] = u[i][2];      // Bottom boundary\n"
               "    //     u[i][1] = u[i][3];    \n"
               "    //     u[i][nyCells + 2] = u[i][nyCells];  // Top boundary\n"
               "    //     u[i][nyCells + 3] = u[i][nyCells + 1];  \n"
               "    // }\n",
               "    if (g_bcy == 2) for (int i = 0; i < nxCells + 2; ++i) {\n"
               "        u[i][0] = u[i][2];      // Bottom boundary\n"
               "        u[i][1] = u[i][3];    \n"
               "        u[i][nyCells + 2] = u[i][nyCells];  // Top boundary\n"
               "        u[i][nyCells + 3] = u[i][nyCells + 1];  \n"
               "    }\n")
    # ---- (1) constants from argv ----
    src = sub1(src, "int main() { \n    int nxCells = 128; \n    int nyCells = 256;\n",
               "int main(int argc, char** argv) { \n"
               "    if (argc < 14) { std::fprintf(stderr, \"%s\\n\", \"" + MHD_USAGE + "\"); return 2; }\n"
               "    g_prob = std::strcmp(argv[1], \"ot\") == 0 ? 1 : 0;\n"
               "    g_bcx = std::strcmp(argv[10], \"transmissive\") == 0 ? 1 : 0;\n"
               "    g_bcy = std::strcmp(argv[11], \"periodic\") == 0 ? 1 : (std::strcmp(argv[11], \"transmissive\") == 0 ? 2 : 0);\n"
               "    g_nsteps = std::atol(argv[12]);\n"
               "    g_writedat = std::atoi(argv[13]);\n"
               "    if (argc > 15) g_logf = std::fopen(argv[15], \"wb\");\n"
               "    int nxCells = std::atoi(argv[2]); \n    int nyCells = std::atoi(argv[3]);\n")
    src = sub1(src, "    double x0 = 0.0;\n    double x1 = 1.0;\n    double y0 = -1.0;\n    double y1 = 1.0;\n",
               "    double x0 = std::strtod(argv[4], nullptr);\n    double x1 = std::strtod(argv[5], nullptr);\n"
               "    double y0 = std::strtod(argv[6], nullptr);\n    double y1 = std::strtod(argv[7], nullptr);\n")
    src = sub1(src, "    double tStop = 12.0;\n    double C = 0.8;\n",
               "    double tStop = std::strtod(argv[8], nullptr);\n    double C = std::strtod(argv[9], nullptr);\n")
    # ---- (2) IC selection ----
    src = sub1(src, "            //Kelvin-Helmhotz\n            prim[0] = 1.0; // Density\n",
               "            //Kelvin-Helmhotz\n            if (g_prob == 0) {\n            prim[0] = 1.0; // Density\n")
    src = sub1(src,
               "            prim[8] = 0;\n\n"
               "            //Orszag-Tang\n"
               "            // prim[0] = gamma*gamma; // Density\n"
               "            // prim[1] = -std::sin(2.0*pi*y); // Velocity\n"
               "            // prim[2] = std::sin(2.0*pi*x);\n"
               "            // prim[3] = 0.0;\n"
               "            // prim[4] = gamma; // pressure\n"
               "            // prim[5] = -std::sin(2.0*pi*y); // magnetic field\n"
               "            // prim[6] = std::sin(4.0*pi*x);\n"
               "            // prim[7] = 0.0; \n"
               "            // prim[8] = 0;\n",
               "            prim[8] = 0;\n            } else {\n"
               "            //Orszag-Tang\n"
               "            prim[0] = gamma*gamma; // Density\n"
               "            prim[1] = -std::sin(2.0*pi*y); // Velocity\n"
               "            prim[2] = std::sin(2.0*pi*x);\n"
               "            prim[3] = 0.0;\n"
               "            prim[4] = gamma; // pressure\n"
               "            prim[5] = -std::sin(2.0*pi*y); // magnetic field\n"
               "            prim[6] = std::sin(4.0*pi*x);\n"
               "            prim[7] = 0.0; \n"
               "            prim[8] = 0;\n            }\n")
    # ---- (6) per-step log + (7) timing ----
    src = sub1(src, "    double t = tStart;\n    int counter =0;\n    do {\n",
               "    double t = tStart;\n    int counter =0;\n"
               "    auto g_t0 = std::chrono::steady_clock::now();\n    do {\n")
    src = sub1(src, "        std::cout << \"t= \" << t << \" dt= \" << dt << \" ch= \"<< maxSpeed<<std::endl;\n",
               "        if (g_writedat) std::cout << \"t= \" << t << \" dt= \" << dt << \" ch= \"<< maxSpeed<<std::endl;\n"
               "        if (g_logf) { double rec[3] = {t, dt, maxSpeed}; std::fwrite(rec, 8, 3, g_logf); }\n")
    # ---- (5) .dat writers optional ----
    src = sub1(src, "        if(t > 0.1*counter && t-dt <= 0.1*counter){\n",
               "        if(g_writedat && t > 0.1*counter && t-dt <= 0.1*counter){\n")
    # ---- (4) loop bound + final dump ----
    src = sub1(src,
               "    } while (t < tStop );\n\n\n    applyBoundaryConditions(u , nxCells , nyCells);\n",
               "        ++g_stepcount;\n"
               "    } while (g_nsteps < 0 ? (t < tStop) : (g_stepcount < g_nsteps));\n"
               "    {\n"
               "        double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - g_t0).count();\n"
               "        std::fprintf(stderr, \"steps=%ld elapsed_s=%.6f t=%.17g\\n\", g_stepcount, secs, t);\n"
               "    }\n"
               "    if (g_logf) std::fclose(g_logf);\n\n"
               "    applyBoundaryConditions(u , nxCells , nyCells);\n"
               "    if (argc > 14 && std::strlen(argv[14]) > 0) {\n"
               "        FILE* f = std::fopen(argv[14], \"wb\");\n"
               "        for (size_t i = 0; i < u.size(); ++i) for (size_t j = 0; j < u[0].size(); ++j) std::fwrite(u[i][j].data(), 8, 9, f);\n"
               "        std::fclose(f);\n"
               "    }\n"
               "    if (!g_writedat) return 0;\n")
    return src


def compile_stdin(src, out, opt="-O3"):
    os.makedirs(os.path.dirname(out), exist_ok=True)
    cmd = ["g++", "-std=c++17", opt, "-w", "-x", "c++", "-", "-o", out]
    subprocess.run(cmd, input=src.encode(), check=True)


def build_mhd(opt="-O3"):
    with open(os.path.join(REF, "MHD_codes", "MUSCL_NEW.cpp")) as f:
        src = f.read()
    compile_stdin(patch_mhd(src), os.path.join(OUT, "muscl_ref"), opt)


def build_all():
    if not os.path.isdir(REF):
        print(f"[oracle/build_ref] {REF} absent - keeping prebuilt oracle/_ref", file=sys.stderr)
        return False
    build_mhd()
    try:
        from build_ref_euler import build_euler  # noqa
        build_euler()
    except ImportError:
        pass
    try:
        from build_ref_tabeos import build_tabeos  # noqa
        build_tabeos()
    except ImportError:
        pass
    try:
        from build_ref_hllc2d import build_hllc2d  # noqa
        build_hllc2d()
    except ImportError:
        pass
    return True


if __name__ == "__main__":
    sys.path.insert(0, HERE)
    ok = build_all()
    print("built" if ok else "skipped", OUT)
