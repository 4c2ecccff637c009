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
#include <vector>

#include "lsc_b200.h"

static int log_line(long, double t, double dt, void* quiet) {
    if (!*static_cast<bool*>(quiet)) std::cout << "t = " << t << " dt = " << dt << std::endl;  // T:963
    return 0;
}

int main(int argc, char** argv) {
    lsc_euler_config cfg = {100, 100, 0.0, 1.0, 0.0, 1.0, 1.4, 0.8, 0.0, LSC_BC_TRANSMISSIVE_E, 0};  // T:896-905
    double tStop = 0.012 / std::pow(10, 2.5);                                                         // T:903
    bool quiet = false, nodat = false;
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
        else {
            std::fprintf(stderr,
                         "usage: tab_eos [--table Plasma19_EoS.txt] [--nx N --ny N --x0 v --x1 v --y0 v --y1 v]\n"
                         "       [--tstop t] [--cfl C] [--omega w] [--outdir d] [--device k] [--quiet] [--no-dat]\n");
            return 2;
        }
    }
    std::vector<double> dens(500), pres(500), sound(500 * 500), en(500 * 500);
    if (lsc_eos_table_read(table.c_str(), dens.data(), pres.data(), sound.data(), en.data())) {  // T:18-144
        std::fprintf(stderr, "%s\n", lsc_last_error());
        return 1;
    }
    lsc_euler_ctx* ctx = nullptr;
    if (lsc_euler_create(&cfg, &ctx) || lsc_euler_set_eos_table(ctx, dens.data(), pres.data(), sound.data(), en.data()) ||
        lsc_euler_init_problem(ctx, LSC_EULER_IC_TABEOS)) {
        std::fprintf(stderr, "%s\n", lsc_last_error());
        return 1;
    }
    if (!quiet) std::cout << "finished loop!" << std::endl;                                  // T:952
    if (lsc_euler_run_until(ctx, tStop, log_line, &quiet, nullptr)) {                        // T:959-971
        std::fprintf(stderr, "%s\n", lsc_last_error());
        return 1;
    }
    std::vector<double> host((size_t)(cfg.nx + 4) * (cfg.ny + 4) * LSC_EULER_NVAR);
    if (!nodat && (lsc_euler_download(ctx, host.data()) ||
                   lsc_euler_tabeos_write_dat(&cfg, host.data(), (outdir + "SLIC2d.dat").c_str(), dens.data(),
                                              pres.data(), en.data()))) {
        std::fprintf(stderr, "%s\n", lsc_last_error());
        return 1;
    }
    lsc_euler_destroy(ctx);
    return 0;
}
