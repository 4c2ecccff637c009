#!/usr/bin/env python3
"""Build the REAL Euler reference programs into oracle/_ref/ (see build_ref.py for the rules).

  slic_ref   <- Euler_codes/SLIC_euler_2d.cpp   (periodic, Euler "Orszag-Tang", 20 dump segments)
  gslic_ref  <- Euler_codes/generaliseSLIC.cpp  (transmissive, 4-quadrant Riemann IC, single dump)

TEST INFRASTRUCTURE ONLY.  The substitutions touch no arithmetic: constants from argv, an optional step
bound, optional suppression of the text output, a raw fp64 dump of the conserved state and a per-step
(t, dt) log.  gslic_ref additionally gets the cylindrical-explosion initial condition of BASELINE config 3,
which does not exist in the reference and is defined identically in
oracle/euler_oracle.c and csrc/euler_host.cpp.
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
static long g_nsteps = -1;
static long g_stepcount = 0;
static int g_writedat = 1;
static int g_ic = 0;
static FILE* g_logf = nullptr;
static const char* g_dumppath = nullptr;
static std::chrono::steady_clock::time_point g_t0;
static void g_finish(const std::vector<std::vector<std::array<double, 4> > >& u, double t) {
    double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - g_t0).count();
    std::fprintf(stderr, "steps=%ld elapsed_s=%.6f t=%.17g\n", g_stepcount, secs, t);
    if (g_logf) std::fclose(g_logf);
    if (g_dumppath && std::strlen(g_dumppath) > 0) {
        FILE* f = std::fopen(g_dumppath, "wb");
        for (size_t i = 0; i < u.size(); ++i) for (size_t j = 0; j < u[0].size(); ++j) std::fwrite(u[i][j].data(), 8, 4, f);
        std::fclose(f);
    }
}
"""

ARGV = (
    "int main(int argc, char** argv){\n"
    "    if (argc < 12) { std::fprintf(stderr, \"usage: nx ny x0 x1 y0 y1 tStop C gamma nsteps writedat [ic [dump [log]]]\\n\"); return 2; }\n"
    "    g_nsteps = std::atol(argv[10]);\n"
    "    g_writedat = std::atoi(argv[11]);\n"
    "    if (argc > 12) g_ic = std::atoi(argv[12]);\n"
    "    if (argc > 13) g_dumppath = argv[13];\n"
    "    if (argc > 14) g_logf = std::fopen(argv[14], \"wb\");\n"
    "    int nxCells = std::atoi(argv[1]);\n"
    "    int nyCells = std::atoi(argv[2]);\n")

CONSTS_OLD = ("    double x0 = 0.0;\n    double x1 = 1.0;\n    double y0 = 0.0;\n    double y1 = 1.0;\n"
              "    double tStart = 0.0;\n    double tStop = 0.5;\n    double C = 0.8;\n    double gamma = 1.4;\n")
CONSTS_NEW = ("    double x0 = std::strtod(argv[3], nullptr);\n    double x1 = std::strtod(argv[4], nullptr);\n"
              "    double y0 = std::strtod(argv[5], nullptr);\n    double y1 = std::strtod(argv[6], nullptr);\n"
              "    double tStart = 0.0;\n    double tStop = std::strtod(argv[7], nullptr);\n"
              "    double C = std::strtod(argv[8], nullptr);\n    double gamma = std::strtod(argv[9], nullptr);\n")

STEP_HOOK = ("g_stepcount++;\n"
             "            if (g_nsteps >= 0 && g_stepcount >= g_nsteps) { g_finish(u, t); return 0; }\n")


def patch_slic(src):
    src = sub1(src, "#include <filesystem>\n", "#include <filesystem>\n" + PROLOGUE)
    src = sub1(src, "int main(){\n    int nxCells = 100;\n    int nyCells = 100;\n", ARGV)
    src = sub1(src, CONSTS_OLD, CONSTS_NEW)
    src = sub1(src, "    double t = tStart;\n    for(int counter =1 ; counter<=20; ++counter){\n",
               "    double t = tStart;\n    g_t0 = std::chrono::steady_clock::now();\n"
               "    for(int counter =1 ; counter<=20; ++counter){\n")
    src = sub1(src, "            dt = ComputeTimeStep(u,C,dx,dy,gamma);\n            t +=dt;\n",
               "            dt = ComputeTimeStep(u,C,dx,dy,gamma);\n            t +=dt;\n"
               "            if (g_logf) { double rec[2] = {t, dt}; std::fwrite(rec, 8, 2, g_logf); }\n")
    src = sub1(src, "            applyBoundaryConditions(u , nxCells , nyCells);\n        }while(t<tStop/(21-counter));\n",
               "            applyBoundaryConditions(u , nxCells , nyCells);\n            " + STEP_HOOK +
               "        }while(t<tStop/(21-counter));\n")
    src = sub1(src, "        std::cout << time << std::endl;\n", "        if (g_writedat) std::cout << time << std::endl;\n")
    src = sub1(src, "        std::ofstream output(filename);\n",
               "        if (!g_writedat) filename = \"/dev/null\";\n        std::ofstream output(filename);\n")
    body = src.rstrip()
    assert body.endswith("}")
    return body[:-1] + "    g_finish(u, t);\n}\n"


CYL_IC = (
    "            if (g_ic == 1) {\n"
    "                if ((x-0.5)*(x-0.5) + (y-0.5)*(y-0.5) <= 0.2*0.2) {\n"
    "                    prim[0] = 1.0; prim[1] = 0; prim[2] = 0; prim[3] = 1.0;\n"
    "                } else {\n"
    "                    prim[0] = 0.125; prim[1] = 0; prim[2] = 0; prim[3] = 0.1;\n"
    "                }\n"
    "            }\n"
    "            else if ( x>0.5 && y>0.5) {\n")


def patch_gslic(src):
    src = sub1(src, "#include <filesystem>\n", "#include <filesystem>\n" + PROLOGUE)
    src = sub1(src, "int main(){\n    int nxCells = 100;\n    int nyCells = 100;\n", ARGV)
    src = sub1(src, CONSTS_OLD, CONSTS_NEW)
    src = sub1(src, "            if ( x>0.5 && y>0.5) {\n", CYL_IC)
    src = sub1(src, "    double t = tStart;\n\n    do{\n",
               "    double t = tStart;\n    g_t0 = std::chrono::steady_clock::now();\n\n    do{\n")
    src = sub1(src, "        std::cout << \"t = \"<< t<<\" dt = \"<< dt<< std::endl; \n",
               "        if (g_writedat) std::cout << \"t = \"<< t<<\" dt = \"<< dt<< std::endl; \n"
               "        if (g_logf) { double rec[2] = {t, dt}; std::fwrite(rec, 8, 2, g_logf); }\n")
    src = sub1(src, "        applyBoundaryConditions(u , nxCells , nyCells);\n    }while(t<tStop);\n",
               "        applyBoundaryConditions(u , nxCells , nyCells);\n        " + STEP_HOOK.replace("            ", "        ") +
               "    }while(t<tStop);\n")
    src = sub1(src, "    std::ofstream output(\"SLIC2d.dat\");\n",
               "    std::ofstream output(g_writedat ? \"SLIC2d.dat\" : \"/dev/null\");\n")
    body = src.rstrip()
    assert body.endswith("}")
    return body[:-1] + "    g_finish(u, t);\n}\n"


def build_euler(opt="-O3"):
    with open(os.path.join(REF, "Euler_codes", "SLIC_euler_2d.cpp")) as f:
        compile_stdin(patch_slic(f.read()), os.path.join(OUT, "slic_ref"), opt)
    with open(os.path.join(REF, "Euler_codes", "generaliseSLIC.cpp")) as f:
        compile_stdin(patch_gslic(f.read()), os.path.join(OUT, "gslic_ref"), opt)


if __name__ == "__main__":
    build_euler()
    print("built", OUT)
