#!/usr/bin/env python3
"""Build the REAL tabulated-EoS reference program into oracle/_ref/tabeos_ref (see build_ref.py for the rules).

  tabeos_ref  <- MHD_codes/TabEoS.cpp  (4-variable Euler SLIC, transmissive, EoS by lookup in a 500x500 table
                                        read from ./Plasma19_EoS.txt)

TEST INFRASTRUCTURE ONLY.  The substitutions touch no arithmetic: constants from argv, an optional step bound,
suppression of the text output, a raw fp64 dump of the conserved state and a per-step (t, dt) log.  The real
Plasma19_EoS.txt is a stripped large blob in the reference tree, so the program is always run in a directory
holding a synthetic table in the same file format (tests/oracle_util.py::write_eos_table): parity of the table
path is pinned against the reference's CODE on that table, not against the committed Euler_codes/SLIC2d.dat.
"""
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from build_ref import OUT, REF, compile_stdin, sub1  # noqa: E402
from build_ref_euler import PROLOGUE, STEP_HOOK  # noqa: E402

ARGV = (
    "int main(int argc, char** argv){\n"
    "    if (argc < 11) { std::fprintf(stderr, \"usage: nx ny x0 x1 y0 y1 tStop C nsteps writedat [dump [log]]\\n\"); return 2; }\n"
    "    g_nsteps = std::atol(argv[9]);\n"
    "    g_writedat = std::atoi(argv[10]);\n"
    "    if (argc > 11) g_dumppath = argv[11];\n"
    "    if (argc > 12) g_logf = std::fopen(argv[12], \"wb\");\n"
    "    int nxCells = std::atoi(argv[1]);\n"
    "    int nyCells = std::atoi(argv[2]);\n")

CONSTS_OLD = ("    double x0 = 0.0;\n    double x1 = 1.0;\n    double y0 = 0.0;\n    double y1 = 1.0;\n"
              "    double tStart = 0.0;\n    double tStop = 0.012/std::pow(10,2.5);\n    double C = 0.8;\n")
CONSTS_NEW = ("    double x0 = std::strtod(argv[3], nullptr);\n    double x1 = std::strtod(argv[4], nullptr);\n"
              "    double y0 = std::strtod(argv[5], nullptr);\n    double y1 = std::strtod(argv[6], nullptr);\n"
              "    double tStart = 0.0;\n    double tStop = std::strtod(argv[7], nullptr);\n"
              "    double C = std::strtod(argv[8], nullptr);\n")


def patch_tabeos(src):
    src = sub1(src, "#include <tuple>\n", "#include <tuple>\n" + PROLOGUE)
    src = sub1(src, "int main(){\n    int nxCells = 100;\n    int nyCells = 100;\n", ARGV)
    src = sub1(src, CONSTS_OLD, CONSTS_NEW)
    # the four progress prints per cell of the IC loop and the one after it
    src = sub1(src, "            std::cout<<\"i= \"<<i<<\" j= \"<<j<<std::endl;\n",
               "            if (g_writedat > 1) std::cout<<\"i= \"<<i<<\" j= \"<<j<<std::endl;\n")
    for var in ("u[i][j]", "v[i][j]"):
        line = ("            std::cout<<%s[0]<<\" \"<<%s[1]<<\" \"<<%s[2]<<\" \"<<%s[3]<<std::endl;\n" % ((var,) * 4))
        src = sub1(src, line, "            if (g_writedat > 1) " + line.lstrip())
    line = "            std::cout<<prim[0]<<\" \"<<prim[1]<<\" \"<<prim[2]<<\" \"<<prim[3]<<std::endl;\n"
    src = sub1(src, line, "            if (g_writedat > 1) " + line.lstrip())
    src = sub1(src, "    std::cout<<\"finished loop!\"<<std::endl;\n",
               "    if (g_writedat > 1) std::cout<<\"finished loop!\"<<std::endl;\n")
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


def build_tabeos(opt="-O3"):
    with open(os.path.join(REF, "MHD_codes", "TabEoS.cpp")) as f:
        compile_stdin(patch_tabeos(f.read()), os.path.join(OUT, "tabeos_ref"), opt)


if __name__ == "__main__":
    build_tabeos()
    print("built", os.path.join(OUT, "tabeos_ref"))
