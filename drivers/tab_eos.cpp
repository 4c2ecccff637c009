// drivers/tab_eos.cpp -- drop-in for the reference program MHD_codes/TabEoS.cpp (T).
//
// Like the reference it reads the equation-of-state table from ./Plasma19_EoS.txt (T:21; --table overrides the
// path), runs the committed configuration (100x100 on [0,1]^2, transmissive boundaries, four quadrants at rest
// with p = 0.01e5 / 1000e5, C = 0.8, tStop = 0.012/10^2.5; T:896-941), prints the log line "t = .. dt = .." every
// step (T:963) and writes SLIC2d.dat through the table's ConservativeToPrimative (T:974-999).  The per-cell
// progress lines of the reference's IC loop (T:943-949: four lines per cell, debugging output) are only printed
// with --verbose-ic.  The numerics run on the GPU behind include/lsc_b200.h; this file holds no arithmetic.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <string>
#include <chrono>
#include <vector>

#include "lsc_b200.h"

static int log_line(long, double t, double dt, void* quiet) {
    if (!*static_cast<bool*>(quiet)) std::cout << "t = " << t << " dt = " << dt << std::endl;  // T:963
    return 0;
}

int main(int argc, char** argv) {
    lsc_euler_config cfg = {100, 100, 0.0, 1.0, 0.0, 1.0, 1.4, 0.8, 0.0, LSC_BC_TRANSMISSIVE_E, 0};  // T:896-905
    double tStop = 0.012 / std::pow(10, 2.5);                                                         // T:903
    bool quiet = false, nodat = false, timing = false;
    int nslabs = 1, ngpus = 0;  // --gpus N: one process, N slabs along x, one GPU each (lsc_euler_group_*)
    std::string outdir, table = "Plasma19_EoS.txt";
    for (int a = 1; a < argc; ++a) {
        auto val = [&](const char* name) -> const char* {
            if (a + 1 >= argc) { std::fprintf(stderr, "%s needs a value\n", name); std::exit(2); }
            return argv[++a];
        };
        const std::string o = argv[a];
        if (o == "--nx") cfg.nx = std::atoi(val("--nx"));
        else if (o == "--ny") cfg.ny = std::atoi(val("--ny"));
        else if (o == "--x0") cfg.x0 = std::strtod(val("--x0"), nullptr);
        else if (o == "--x1") cfg.x1 = std::strtod(val("--x1"), nullptr);
        else if (o == "--y0") cfg.y0 = std::strtod(val("--y0"), nullptr);
        else if (o == "--y1") cfg.y1 = std::strtod(val("--y1"), nullptr);
        else if (o == "--tstop") tStop = std::strtod(val("--tstop"), nullptr);
        else if (o == "--cfl") cfg.cfl = std::strtod(val("--cfl"), nullptr);
        else if (o == "--omega") cfg.omega = std::strtod(val("--omega"), nullptr);
        else if (o == "--device") cfg.device = std::atoi(val("--device"));
        else if (o == "--outdir") outdir = std::string(val("--outdir")) + "/";
        else if (o == "--table") table = val("--table");
        else if (o == "--quiet") quiet = true;
        else if (o == "--no-dat") nodat = true;
        else if (o == "--gpus") { ngpus = std::atoi(val("--gpus")); nslabs = ngpus; }
        else if (o == "--slabs") nslabs = std::atoi(val("--slabs"));
        else if (o == "--time") timing = true;
        else {
            std::fprintf(stderr,
                         "usage: tab_eos [--table Plasma19_EoS.txt] [--nx N --ny N --x0 v --x1 v --y0 v --y1 v]\n"
                         "       [--tstop t] [--cfl C] [--omega w] [--outdir d] [--device k] [--gpus N | --slabs N] [--time]\n"
                         "       [--quiet] [--no-dat]\n");
            return 2;
        }
    }
    std::vector<double> dens(500), pres(500), sound(500 * 500), en(500 * 500);
    if (lsc_eos_table_read(table.c_str(), dens.data(), pres.data(), sound.data(), en.data())) {  // T:18-144
        std::fprintf(stderr, "%s\n", lsc_last_error());
        return 1;
    }
    const int ndev = lsc_device_count();
    if (nslabs < 1 || ngpus > ndev) { std::fprintf(stderr, "--gpus %d: %d CUDA device(s) visible\n", ngpus, ndev); return 1; }
    auto die = [] { std::fprintf(stderr, "%s\n", lsc_last_error()); return 1; };
    std::vector<lsc_euler_ctx*> slabs(nslabs, nullptr);
    for (int r = 0; r < nslabs; ++r) {
        lsc_euler_config c = cfg;
        c.rank = r;
        c.nranks = nslabs;
        c.device = ndev > 0 ? (cfg.device + r) % ndev : 0;
        if (lsc_euler_create(&c, &slabs[r]) ||
            lsc_euler_set_eos_table(slabs[r], dens.data(), pres.data(), sound.data(), en.data()) ||
            lsc_euler_init_problem(slabs[r], LSC_EULER_IC_TABEOS))
            return die();
    }
    if (!quiet) std::cout << "finished loop!" << std::endl;                                  // T:952
    if (lsc_euler_group_step(slabs.data(), nslabs, 0)) return die();  // one-time team setup (peer access) outside --time
    const auto t0 = std::chrono::steady_clock::now();
    long steps = 0;
    if (lsc_euler_group_run_until(slabs.data(), nslabs, tStop, log_line, &quiet, &steps)) return die();  // T:959-971
    if (timing) {
        const double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        std::fprintf(stderr, "steps=%ld elapsed_s=%.6f cell_updates_per_s=%.6e slabs=%d\n", steps, secs,
                     (double)cfg.nx * cfg.ny * steps / secs, nslabs);
    }
    if (!nodat) {
        std::vector<double> host((size_t)(cfg.nx + 4) * (cfg.ny + 4) * LSC_EULER_NVAR);
        if (lsc_euler_group_download(slabs.data(), nslabs, host.data()) ||
            lsc_euler_tabeos_write_dat(&cfg, host.data(), (outdir + "SLIC2d.dat").c_str(), dens.data(), pres.data(),
                                       en.data()))
            return die();
    }
    for (auto c : slabs) lsc_euler_destroy(c);
    return 0;
}
