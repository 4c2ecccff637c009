// drivers/muscl_new.cpp -- drop-in for the reference program MHD_codes/MUSCL_NEW.cpp.
//
// With no arguments it runs the committed configuration (Kelvin-Helmholtz 128x256, periodic-x /
// reflective-y, C = 0.8, gamma = 5/3, tStop = 12; MUSCL_NEW.cpp:659-671, :800-808) and writes the same
// stdout log (:836) and the same MHD<counter>.dat files (:975-1001, :1006-1032), byte for byte.  The
// constants the reference makes you edit in source are command-line options here.  All numerics run on
// the GPU behind the C ABI of include/lsc_b200.h; this file holds no arithmetic of its own.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <string>
#include <vector>

#include "lsc_b200.h"

namespace {

struct Run {
    lsc_mhd_config cfg;
    std::vector<double> host;
    std::string outdir;
    bool quiet = false;
    bool nodat = false;
    int last_counter = 0;
};

int on_log(long, double t, double dt, double ch, void* user) {
    if (!static_cast<Run*>(user)->quiet)
        std::cout << "t= " << t << " dt= " << dt << " ch= " << ch << std::endl;  // MUSCL_NEW.cpp:836
    return 0;
}

int on_dump(lsc_mhd_ctx* ctx, int counter, double, double, void* user) {
    Run* r = static_cast<Run*>(user);
    r->last_counter = counter + 1;
    if (r->nodat) return 0;
    if (int rc = lsc_mhd_download(ctx, r->host.data())) return rc;
    const std::string name = r->outdir + "MHD" + std::to_string(counter) + ".dat";  // :990
    return lsc_mhd_write_dat(&r->cfg, r->host.data(), name.c_str());
}

int parse_bc(const char* s) {
    if (!std::strcmp(s, "periodic")) return LSC_BC_PERIODIC;
    if (!std::strcmp(s, "reflective")) return LSC_BC_REFLECTIVE;
    if (!std::strcmp(s, "transmissive")) return LSC_BC_TRANSMISSIVE;
    std::fprintf(stderr, "unknown boundary type %s\n", s);
    std::exit(2);
}

}  // namespace

int main(int argc, char** argv) {
    Run run;
    // committed constants, MUSCL_NEW.cpp:659-668
    run.cfg = {128, 256, 0.0, 1.0, -1.0, 1.0, 5.0 / 3.0, 0.8, LSC_BC_PERIODIC, LSC_BC_REFLECTIVE, 0, 0, 1};
    double tStop = 12.0;
    int problem = LSC_MHD_IC_KELVIN_HELMHOLTZ;
    long nsteps = -1;
    for (int a = 1; a < argc; ++a) {
        auto val = [&](const char* name) -> const char* {
            if (a + 1 >= argc) { std::fprintf(stderr, "%s needs a value\n", name); std::exit(2); }
            return argv[++a];
        };
        const std::string o = argv[a];
        if (o == "--nx") run.cfg.nx = std::atoi(val("--nx"));
        else if (o == "--ny") run.cfg.ny = std::atoi(val("--ny"));
        else if (o == "--x0") run.cfg.x0 = std::strtod(val("--x0"), nullptr);
        else if (o == "--x1") run.cfg.x1 = std::strtod(val("--x1"), nullptr);
        else if (o == "--y0") run.cfg.y0 = std::strtod(val("--y0"), nullptr);
        else if (o == "--y1") run.cfg.y1 = std::strtod(val("--y1"), nullptr);
        else if (o == "--tstop") tStop = std::strtod(val("--tstop"), nullptr);
        else if (o == "--cfl") run.cfg.cfl = std::strtod(val("--cfl"), nullptr);
        else if (o == "--gamma") run.cfg.gamma = std::strtod(val("--gamma"), nullptr);
        else if (o == "--bcx") run.cfg.bc_x = parse_bc(val("--bcx"));
        else if (o == "--bcy") run.cfg.bc_y = parse_bc(val("--bcy"));
        else if (o == "--device") run.cfg.device = std::atoi(val("--device"));
        else if (o == "--steps") nsteps = std::atol(val("--steps"));
        else if (o == "--outdir") run.outdir = std::string(val("--outdir")) + "/";
        else if (o == "--quiet") run.quiet = true;
        else if (o == "--no-dat") run.nodat = true;
        else if (o == "--problem") {
            const std::string p = val("--problem");
            if (p == "kh") problem = LSC_MHD_IC_KELVIN_HELMHOLTZ;
            else if (p == "ot") problem = LSC_MHD_IC_ORSZAG_TANG;
            else { std::fprintf(stderr, "unknown problem %s (kh|ot)\n", p.c_str()); return 2; }
        } else {
            std::fprintf(stderr,
                         "usage: muscl_new [--problem kh|ot] [--nx N] [--ny N] [--x0 v --x1 v --y0 v --y1 v]\n"
                         "                 [--tstop t] [--cfl C] [--gamma g] [--bcx periodic|transmissive]\n"
                         "                 [--bcy reflective|periodic|transmissive] [--steps n] [--outdir d]\n"
                         "                 [--device k] [--quiet] [--no-dat]\n");
            return 2;
        }
    }

    lsc_mhd_ctx* ctx = nullptr;
    if (lsc_mhd_create(&run.cfg, &ctx)) { std::fprintf(stderr, "%s\n", lsc_last_error()); return 1; }
    run.host.resize((size_t)(run.cfg.nx + 4) * (run.cfg.ny + 4) * LSC_MHD_NVAR);
    // initial conditions: MUSCL_NEW.cpp:719-825, evaluated on the host with the same libm
    if (lsc_mhd_initial_condition(ctx, problem, run.host.data()) || lsc_mhd_upload(ctx, run.host.data())) {
        std::fprintf(stderr, "%s\n", lsc_last_error());
        return 1;
    }
    int rc;
    if (nsteps >= 0) {
        rc = 0;
        for (long s = 0; s < nsteps && !rc; ++s) {
            rc = lsc_mhd_step(ctx, 1);
            double t, dt, ch;
            lsc_mhd_get_scalars(ctx, &t, &dt, &ch, nullptr);
            on_log(s, t, dt, ch, &run);
        }
    } else {
        rc = lsc_mhd_run_until(ctx, tStop, on_dump, on_log, &run, nullptr);  // :831-1003
    }
    if (rc) { std::fprintf(stderr, "%s\n", lsc_last_error()); return 1; }
    if (!run.nodat) {  // final dump, :1006-1032
        if (lsc_mhd_download(ctx, run.host.data())) { std::fprintf(stderr, "%s\n", lsc_last_error()); return 1; }
        const std::string name = run.outdir + "MHD" + std::to_string(run.last_counter) + ".dat";
        if (lsc_mhd_write_dat(&run.cfg, run.host.data(), name.c_str())) {
            std::fprintf(stderr, "%s\n", lsc_last_error());
            return 1;
        }
    }
    lsc_mhd_destroy(ctx);
    return 0;
}
