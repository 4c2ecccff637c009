// drivers/muscl_new.cpp -- drop-in for the reference program MHD_codes/MUSCL_NEW.cpp.
//
// With no arguments it runs the committed configuration (Kelvin-Helmholtz 128x256, periodic-x /
// reflective-y, C = 0.8, gamma = 5/3, tStop = 12; MUSCL_NEW.cpp:659-671, :800-808) and writes the same
// stdout log (:836) and the same MHD<counter>.dat files (:975-1001, :1006-1032), byte for byte.  The
// constants the reference makes you edit in source are command-line options here.  All numerics run on
// the GPU behind the C ABI of include/lsc_b200.h; this file holds no arithmetic of its own.
//
// --gpus N: still ONE process, as the reference is (MUSCL_NEW.cpp:658-1034) -- the grid is cut into N slabs along x,
// one context per GPU (lsc_mhd_group_*: halos by peer copies over NVLink, device-side max reduction, one host
// synchronisation per batch of steps).  --slabs N with fewer GPUs puts several slabs on one device (tests).
// stdout and the .dat files do not depend on N.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <string>
#include <thread>
#include <vector>

#include "lsc_b200.h"

namespace {

struct Run {
    lsc_mhd_config cfg;
    std::vector<lsc_mhd_ctx*> slabs;
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

int on_dump(lsc_mhd_ctx*, int counter, double, double, void* user) {
    Run* r = static_cast<Run*>(user);
    r->last_counter = counter + 1;
    if (r->nodat) return 0;
    if (int rc = lsc_mhd_group_download(r->slabs.data(), (int)r->slabs.size(), r->host.data())) return rc;
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
    run.cfg = {128, 256, 0.0, 1.0, -1.0, 1.0, 5.0 / 3.0, 0.8, LSC_BC_PERIODIC, LSC_BC_REFLECTIVE, 0, 0, 1, 0, 0};
    double tStop = 12.0;
    int problem = LSC_MHD_IC_KELVIN_HELMHOLTZ;
    long nsteps = -1;
    int nslabs = 1, ngpus = 0;
    bool timing = false;
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
        else if (o == "--gpus") { ngpus = std::atoi(val("--gpus")); nslabs = ngpus; }
        else if (o == "--slabs") nslabs = std::atoi(val("--slabs"));
        else if (o == "--time") timing = true;
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
                         "                 [--device k] [--gpus N | --slabs N] [--time] [--quiet] [--no-dat]\n");
            return 2;
        }
    }

    if (nslabs < 1) { std::fprintf(stderr, "--gpus / --slabs must be >= 1\n"); return 2; }
    const int ndev = lsc_device_count();
    if (ngpus > ndev) { std::fprintf(stderr, "--gpus %d but only %d CUDA device(s) are visible\n", ngpus, ndev); return 1; }
    auto die = [] { std::fprintf(stderr, "%s\n", lsc_last_error()); return 1; };
    // one context per slab; slab r of N owns columns [r nx/N, (r+1) nx/N) and lives on GPU (device + r) mod #GPUs.
    // Initial conditions: MUSCL_NEW.cpp:719-825, evaluated on the host with the same libm, slab by slab -- one host
    // thread per slab (268 M cells of sin() at 16384^2 would otherwise take a minute before the first step).
    run.slabs.assign(nslabs, nullptr);
    std::vector<std::string> errors(nslabs);
    std::vector<std::thread> workers;
    for (int r = 0; r < nslabs; ++r)
        workers.emplace_back([&, r] {
            lsc_mhd_config c = run.cfg;
            c.rank = r;
            c.nranks = nslabs;
            c.device = ndev > 0 ? (run.cfg.device + r) % ndev : 0;
            if (lsc_mhd_create(&c, &run.slabs[r]) || lsc_mhd_init_problem(run.slabs[r], problem))
                errors[r] = lsc_last_error();
        });
    for (auto& w : workers) w.join();
    for (int r = 0; r < nslabs; ++r)
        if (!errors[r].empty()) { std::fprintf(stderr, "slab %d: %s\n", r, errors[r].c_str()); return 1; }
    lsc_mhd_ctx** g = run.slabs.data();
    if (!run.nodat) run.host.resize((size_t)(run.cfg.nx + 4) * (run.cfg.ny + 4) * LSC_MHD_NVAR);
    // one-time setup of the team (peer access between the GPUs) outside the timed span of --time
    if (lsc_mhd_group_step(g, nslabs, 0)) return die();
    const auto t0 = std::chrono::steady_clock::now();
    long steps_taken = 0;
    int rc;
    if (nsteps >= 0) {
        if (run.quiet) {
            rc = lsc_mhd_group_step(g, nslabs, nsteps);
        } else {
            rc = 0;
            for (long s = 0; s < nsteps && !rc; ++s) {
                rc = lsc_mhd_group_step(g, nslabs, 1);
                double t, dt, ch;
                lsc_mhd_get_scalars(g[0], &t, &dt, &ch, nullptr);
                on_log(s, t, dt, ch, &run);
            }
        }
        steps_taken = nsteps;
    } else {
        rc = lsc_mhd_group_run_until(g, nslabs, tStop, on_dump, run.quiet ? nullptr : on_log, &run, &steps_taken);  // :831-1003
    }
    if (rc) return die();
    if (timing) {
        const double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        long syncs = 0, misses = 0;
        lsc_mhd_loop_stats(g[0], &syncs, &misses);
        std::fprintf(stderr, "steps=%ld elapsed_s=%.6f cell_updates_per_s=%.6e slabs=%d host_syncs=%ld exp_table_misses=%ld\n",
                     steps_taken, secs, (double)run.cfg.nx * run.cfg.ny * steps_taken / secs, nslabs, syncs, misses);
    }
    if (!run.nodat) {  // final dump, :1006-1032
        if (lsc_mhd_group_download(g, nslabs, run.host.data())) return die();
        const std::string name = run.outdir + "MHD" + std::to_string(run.last_counter) + ".dat";
        if (lsc_mhd_write_dat(&run.cfg, run.host.data(), name.c_str())) return die();
    }
    for (auto c : run.slabs) lsc_mhd_destroy(c);
    return 0;
}
