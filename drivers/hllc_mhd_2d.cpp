// drivers/hllc_mhd_2d.cpp -- drop-in for the reference program MHD_codes/HLLC_MHD_2D.cpp (first-order Godunov +
// HLLC + mixed GLM).  With no arguments it runs the committed configuration
// (Orszag-Tang 800 x 800 on the unit square, periodic, C = 0.75, gamma = 5/3, tStop = 0.5; HLLC_MHD_2D.cpp:616-627)
// and writes the same stdout log (:738) and the same MHD.dat (:793-814).  All numerics run on the GPU behind the C
// ABI of include/lsc_b200.h (lsc_hllc2d_*).
#include <cstdio>
#include <cstdlib>
#include <iostream>
#include <string>
#include <vector>

#include "lsc_b200.h"

static int log_line(long, double t, double dt, double ch, void* quiet) {
    if (!*static_cast<bool*>(quiet))
        std::cout << "t= " << t << " dt= " << dt << " ch = " << ch << std::endl;  // HLLC_MHD_2D.cpp:738
    return 0;
}

int main(int argc, char** argv) {
    lsc_hllc2d_config cfg = {800, 800, 0.0, 1.0, 0.0, 1.0, 5.0 / 3.0, 0.75, 0};  // :616-625
    double tStop = 0.5;
    bool quiet = false, nodat = false;
    std::string outdir;
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
        else if (o == "--gamma") cfg.gamma = std::strtod(val("--gamma"), nullptr);
        else if (o == "--device") cfg.device = std::atoi(val("--device"));
        else if (o == "--outdir") outdir = std::string(val("--outdir")) + "/";
        else if (o == "--quiet") quiet = true;
        else if (o == "--no-dat") nodat = true;
        else {
            std::fprintf(stderr, "usage: hllc_mhd_2d [--nx N --ny N --x0 v --x1 v --y0 v --y1 v] [--tstop t] [--cfl C]\n"
                                 "                   [--gamma g] [--outdir d] [--device k] [--quiet] [--no-dat]\n");
            return 2;
        }
    }
    auto die = [] { std::fprintf(stderr, "%s\n", lsc_last_error()); return 1; };
    lsc_hllc2d_ctx* ctx = nullptr;
    if (lsc_hllc2d_create(&cfg, &ctx) || lsc_hllc2d_init_problem(ctx)) return die();  // :630-721
    if (lsc_hllc2d_run_until(ctx, tStop, log_line, &quiet, nullptr)) return die();      // :727-785
    if (!nodat) {
        std::vector<double> host((size_t)(cfg.nx + 2) * (cfg.ny + 2) * LSC_MHD_NVAR);
        if (lsc_hllc2d_download(ctx, host.data()) || lsc_hllc2d_write_dat(&cfg, host.data(), (outdir + "MHD.dat").c_str()))
            return die();
    }
    lsc_hllc2d_destroy(ctx);
    return 0;
}
