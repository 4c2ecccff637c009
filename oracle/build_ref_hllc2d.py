#!/usr/bin/env python3
"""Build the REAL first-order program into oracle/_ref/hllc2d_ref (see build_ref.py for the rules).

  hllc2d_ref  <- MHD_codes/HLLC_MHD_2D.cpp  (first-order Godunov + HLLC + mixed GLM, ONE ghost cell, periodic in
                                              x and y, Orszag-Tang)

TEST INFRASTRUCTURE ONLY.  The substitutions touch no arithmetic: grid / domain / tStop / C from argv, an optional
step bound, optional suppression of the text output, a raw fp64 dump of the conserved state AFTER the program's
final applyBoundaryConditions (HLLC_MHD_2D.cpp:787), a per-step (t, dt, ch) log and a loop timer.
"""
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from build_ref import OUT, REF, compile_stdin, sub1  # noqa: E402

PROLOGUE = r"""
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
static long g_nsteps = -1;  /* <0: run until t>=tStop as the reference does */
static long g_stepcount = 0;
static int g_writedat = 1;
static FILE* g_logf = nullptr;
"""

USAGE = "usage: hllc2d_ref nx ny x0 x1 y0 y1 tStop C nsteps(-1=until tStop) writedat(0|1) [state_dump.bin [steplog.bin]]"


def patch_hllc2d(src):
    src = sub1(src, "#include <tuple>\n", "#include <tuple>\n" + PROLOGUE)
    src = sub1(src, "int main() { \n    int nxCells = 800; \n    int nyCells = 800;\n",
               "int main(int argc, char** argv) { \n"
               "    if (argc < 11) { std::fprintf(stderr, \"%s\\n\", \"" + USAGE + "\"); return 2; }\n"
               "    g_nsteps = std::atol(argv[9]);\n"
               "    g_writedat = std::atoi(argv[10]);\n"
               "    if (argc > 12) g_logf = std::fopen(argv[12], \"wb\");\n"
               "    int nxCells = std::atoi(argv[1]); \n    int nyCells = std::atoi(argv[2]);\n")
    src = sub1(src, "    double x0 = 0.0;\n    double x1 = 1.0;\n    double y0 = 0.0;\n    double y1 = 1.0;\n",
               "    double x0 = std::strtod(argv[3], nullptr);\n    double x1 = std::strtod(argv[4], nullptr);\n"
               "    double y0 = std::strtod(argv[5], nullptr);\n    double y1 = std::strtod(argv[6], nullptr);\n")
    src = sub1(src, "    double tStop = 0.5;\n    double C = 0.75;\n",
               "    double tStop = std::strtod(argv[7], nullptr);\n    double C = std::strtod(argv[8], nullptr);\n")
    src = sub1(src, "    double t = tStart;\n    do {\n",
               "    double t = tStart;\n    auto g_t0 = std::chrono::steady_clock::now();\n    do {\n")
    src = sub1(src, "        std::cout << \"t= \" << t << \" dt= \"<< dt << \" ch = \"<< maxSpeed << std::endl;\n",
               "        if (g_writedat) std::cout << \"t= \" << t << \" dt= \"<< dt << \" ch = \"<< maxSpeed << std::endl;\n"
               "        if (g_logf) { double rec[3] = {t, dt, maxSpeed}; std::fwrite(rec, 8, 3, g_logf); }\n")
    src = sub1(src, "        u = uPlus1;\n    } while (t < tStop);\n\n    applyBoundaryConditions(u , nxCells , nyCells);\n",
               "        u = uPlus1;\n        ++g_stepcount;\n"
               "    } while (g_nsteps < 0 ? (t < tStop) : (g_stepcount < g_nsteps));\n"
               "    {\n"
               "        double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - g_t0).count();\n"
               "        std::fprintf(stderr, \"steps=%ld elapsed_s=%.6f t=%.17g\\n\", g_stepcount, secs, t);\n"
               "    }\n"
               "    if (g_logf) std::fclose(g_logf);\n\n"
               "    applyBoundaryConditions(u , nxCells , nyCells);\n"
               "    if (argc > 11 && std::strlen(argv[11]) > 0) {\n"
               "        FILE* f = std::fopen(argv[11], \"wb\");\n"
               "        for (size_t i = 0; i < u.size(); ++i) for (size_t j = 0; j < u[0].size(); ++j) std::fwrite(u[i][j].data(), 8, 9, f);\n"
               "        std::fclose(f);\n"
               "    }\n"
               "    if (!g_writedat) return 0;\n")
    return src


def build_hllc2d(opt="-O3"):
    with open(os.path.join(REF, "MHD_codes", "HLLC_MHD_2D.cpp")) as f:
        compile_stdin(patch_hllc2d(f.read()), os.path.join(OUT, "hllc2d_ref"), opt)


if __name__ == "__main__":
    build_hllc2d()
    print("built", os.path.join(OUT, "hllc2d_ref"))
