// drivers/slic_euler_2d.cpp -- drop-in for the reference programs Euler_codes/SLIC_euler_2d.cpp (default) and
// Euler_codes/generaliseSLIC.cpp (--generalise).
//
// Default: the committed configuration of SLIC_euler_2d.cpp (100x100, periodic, Euler "Orszag-Tang", C = 0.8,
// gamma = 1.4, tStop = 0.5): twenty segments, after segment k it prints tStop/(21-k) (S:619) and writes
// euler_<k>.dat (S:632-642), byte for byte what the reference writes.
// --generalise: generaliseSLIC.cpp's main(): transmissive boundaries, four-quadrant initial data, the log line
// "t = .. dt = .." every step (G:631) and one SLIC2d.dat with a blank line after each row (G:659-667).
// The numerics run on the GPU behind include/lsc_b200.h; this file holds no arithmetic.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <string>
#include <chrono>
#include <vector>

#include "lsc_b200.h"

static int log_line(long, double t, double dt, void* quiet) {
    if (!*static_cast<bool*>(quiet)) std::cout << "t = " << t << " dt = " << dt << std::endl;  // G:631
    return 0;
}

int main(int argc, char** argv) {
    lsc_euler_config cfg = {100, 100, 0.0, 1.0, 0.0, 1.0, 1.4, 0.8, 0.0, LSC_BC_PERIODIC, 0};  // S:529-539
    double tStop = 0.5;
    int problem = -1;
    bool generalise = false, quiet = false, nodat = false, timing = false;
    int nslabs = 1, ngpus = 0;  // --gpus N: one process, N slabs along x, one GPU each (lsc_euler_group_*)
    std::string outdir;
    for (int a = 1; a < argc; ++a) {
        auto val = [&](const char* name) -> const char* {
            if (a + 1 >= argc) { std::fprintf(stderr, "%s needs a value\n", name); std::exit(2); }
            return argv[++a];
        };
        const std::string o = argv[a];
        if (o == "--generalise") generalise = true;
        else if (o == "--nx") cfg.nx = std::atoi(val("--nx"));
        else if (o == "--ny") cfg.ny = std::atoi(val("--ny"));
        else if (o == "--x0") cfg.x0 = std::strtod(val("--x0"), nullptr);
        else if (o == "--x1") cfg.x1 = std::strtod(val("--x1"), nullptr);
        else if (o == "--y0") cfg.y0 = std::strtod(val("--y0"), nullptr);
        else if (o == "--y1") cfg.y1 = std::strtod(val("--y1"), nullptr);
        else if (o == "--tstop") tStop = std::strtod(val("--tstop"), nullptr);
        else if (o == "--cfl") cfg.cfl = std::strtod(val("--cfl"), nullptr);
        else if (o == "--gamma") cfg.gamma = std::strtod(val("--gamma"), nullptr);
        else if (o == "--omega") cfg.omega = std::strtod(val("--omega"), nullptr);
        else if (o == "--device") cfg.device = std::atoi(val("--device"));
        else if (o == "--outdir") outdir = std::string(val("--outdir")) + "/";
        else if (o == "--quiet") quiet = true;
        else if (o == "--no-dat") nodat = true;
        else if (o == "--gpus") { ngpus = std::atoi(val("--gpus")); nslabs = ngpus; }
        else if (o == "--slabs") nslabs = std::atoi(val("--slabs"));
        else if (o == "--time") timing = true;
        else if (o == "--bc") {
            const std::string b = val("--bc");
            if (b == "periodic") cfg.bc = LSC_BC_PERIODIC;
            else if (b == "transmissive") cfg.bc = LSC_BC_TRANSMISSIVE_E;
            else { std::fprintf(stderr, "unknown boundary type %s\n", b.c_str()); return 2; }
        } else if (o == "--problem") {
            const std::string p = val("--problem");
            if (p == "ot") problem = LSC_EULER_IC_ORSZAG_TANG;
            else if (p == "quadrants") problem = LSC_EULER_IC_QUADRANTS;
            else if (p == "cylinder") problem = LSC_EULER_IC_CYLINDER;
            else { std::fprintf(stderr, "unknown problem %s (ot|quadrants|cylinder)\n", p.c_str()); return 2; }
        } else {
            std::fprintf(stderr,
                         "usage: slic_euler_2d [--generalise] [--problem ot|quadrants|cylinder] [--bc periodic|transmissive]\n"
                         "       [--nx N --ny N --x0 v --x1 v --y0 v --y1 v] [--tstop t] [--cfl C] [--gamma g] [--omega w]\n"
                         "       [--outdir d] [--device k] [--gpus N | --slabs N] [--time] [--quiet] [--no-dat]\n");
            return 2;
        }
    }
    if (generalise) {  // generaliseSLIC.cpp's committed choices unless overridden
        bool bc_given = false, prob_given = problem >= 0;
        for (int a = 1; a < argc; ++a) bc_given = bc_given || !std::strcmp(argv[a], "--bc");
        if (!bc_given) cfg.bc = LSC_BC_TRANSMISSIVE_E;
        if (!prob_given) problem = LSC_EULER_IC_QUADRANTS;
    }
    if (problem < 0) problem = LSC_EULER_IC_ORSZAG_TANG;

    const int ndev = lsc_device_count();
    if (nslabs < 1 || ngpus > ndev) { std::fprintf(stderr, "--gpus %d: %d CUDA device(s) visible\n", ngpus, ndev); return 1; }
    auto die = [] { std::fprintf(stderr, "%s\n", lsc_last_error()); return 1; };
    std::vector<lsc_euler_ctx*> slabs(nslabs, nullptr);
    for (int r = 0; r < nslabs; ++r) {
        lsc_euler_config c = cfg;
        c.rank = r;
        c.nranks = nslabs;
        c.device = ndev > 0 ? (cfg.device + r) % ndev : 0;
        if (lsc_euler_create(&c, &slabs[r]) || lsc_euler_init_problem(slabs[r], problem)) return die();
    }
    lsc_euler_ctx** g = slabs.data();
    std::vector<double> host(nodat ? 0 : (size_t)(cfg.nx + 4) * (cfg.ny + 4) * LSC_EULER_NVAR);
    if (lsc_euler_group_step(slabs.data(), nslabs, 0)) return die();  // one-time team setup (peer access) outside --time
    const auto t0 = std::chrono::steady_clock::now();
    long steps = 0, k = 0;
    if (generalise) {
        if (lsc_euler_group_run_until(g, nslabs, tStop, log_line, &quiet, &steps)) return die();  // G:627-646
        if (!nodat && (lsc_euler_group_download(g, nslabs, host.data()) ||
                       lsc_euler_write_dat(&cfg, host.data(), (outdir + "SLIC2d.dat").c_str(), 1)))
            return die();
    } else {
        for (int counter = 1; counter <= 20; ++counter) {  // S:599
            const double seg = tStop / (21 - counter);
            if (lsc_euler_group_run_until(g, nslabs, seg, nullptr, nullptr, &k)) return die();  // S:600-617
            steps += k;
            const double time = tStop / (21 - counter);  // S:618
            if (!quiet) std::cout << time << std::endl;  // S:619
            if (nodat) continue;
            const std::string name = outdir + "euler_" + std::to_string(counter) + ".dat";  // S:632
            if (lsc_euler_group_download(g, nslabs, host.data()) || lsc_euler_write_dat(&cfg, host.data(), name.c_str(), 0))
                return die();
        }
    }
    if (timing) {
        const double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        std::fprintf(stderr, "steps=%ld elapsed_s=%.6f cell_updates_per_s=%.6e slabs=%d\n", steps, secs,
                     (double)cfg.nx * cfg.ny * steps / secs, nslabs);
    }
    for (auto c : slabs) lsc_euler_destroy(c);
    return 0;
}
