// mhd_host.cpp -- host-side pieces of the MHD solver surface that must reproduce the
// reference bit-for-bit / byte-for-byte and therefore run on the host with the same libm
// and iostream formatting the reference uses: initial conditions (MUSCL_NEW.cpp:719-825),
// the ghost ring of a downloaded state (MUSCL_NEW.cpp:607-655) and the .dat writer
// (MUSCL_NEW.cpp:989-1001).  Compiled by g++ without FMA contraction.
#include "mhd_host.h"

#include <cmath>
#include <cstring>
#include <fstream>
#include <string>

namespace lsc {

namespace {
inline double* at(double* a, int ny, int i, int j) { return a + ((size_t)i * (size_t)(ny + 4) + (size_t)j) * 9; }
inline const double* at(const double* a, int ny, int i, int j) {
    return a + ((size_t)i * (size_t)(ny + 4) + (size_t)j) * 9;
}

// PrimitiveToConservative, MUSCL_NEW.cpp:9-21
inline void prim_to_cons(const double (&p)[9], double gamma, double* c) {
    c[0] = p[0];
    c[1] = p[0] * p[1];
    c[2] = p[0] * p[2];
    c[3] = p[0] * p[3];
    c[5] = p[5];
    c[6] = p[6];
    c[7] = p[7];
    c[4] = p[4] / (gamma - 1) + 0.5 * p[0] * (p[1] * p[1] + p[2] * p[2] + p[3] * p[3]) +
           0.5 * (p[5] * p[5] + p[6] * p[6] + p[7] * p[7]);
    c[8] = p[8];
}
}  // namespace

// Columns [il0, il0+ni) (slab-local indices, ghosts included) of the initial condition into
// `aos` (chunk-local: column il0 first).  i_global0 = global index of slab-local column 0.
void mhd_initial_condition(const lsc_mhd_config& c, int problem, int i_global0, int il0, int ni, double* aos) {
    const double pi = 3.14159265358979323846;  // MUSCL_NEW.cpp:671
    const double gamma = c.gamma;
    const double dx = (c.x1 - c.x0) / c.nx;
    const double dy = (c.y1 - c.y0) / c.ny;
    for (int q = 0; q < ni; ++q) {
        const int i = i_global0 + il0 + q;
        for (int j = 0; j < c.ny + 4; ++j) {
            const double x = c.x0 + (i - 0.5) * dx;  // MUSCL_NEW.cpp:722-723
            const double y = c.y0 + (j - 0.5) * dy;
            double prim[9];
            if (problem == LSC_MHD_IC_KELVIN_HELMHOLTZ) {  // MUSCL_NEW.cpp:800-808
                prim[0] = 1.0;
                prim[1] = 0.5 * std::tanh(20 * y);
                prim[2] = 0.01 * std::sin(2.0 * pi * x) * std::exp(-y * y / (0.01));
                prim[3] = 0.0;
                prim[4] = 1.0 / gamma;
                prim[5] = 0.1 * std::cos(pi / 3.0);
                prim[6] = 0.0;
                prim[7] = 0.1 * std::sin(pi / 3.0);
                prim[8] = 0;
            } else {  // Orszag-Tang, MUSCL_NEW.cpp:811-819
                prim[0] = gamma * gamma;
                prim[1] = -std::sin(2.0 * pi * y);
                prim[2] = std::sin(2.0 * pi * x);
                prim[3] = 0.0;
                prim[4] = gamma;
                prim[5] = -std::sin(2.0 * pi * y);
                prim[6] = std::sin(4.0 * pi * x);
                prim[7] = 0.0;
                prim[8] = 0;
            }
            prim_to_cons(prim, gamma, aos + ((size_t)q * (size_t)(c.ny + 4) + (size_t)j) * 9);
        }
    }
}

// One application of the reference's boundary function to a full-grid AoS array.
// Loop bounds stop two short (`< n+2`), x loop before y loop -- MUSCL_NEW.cpp:607-655.
static void apply_bc_x(const lsc_mhd_config& c, double* a) {
    const int nx = c.nx, ny = c.ny;
    const size_t cell = 9 * sizeof(double);
    for (int j = 0; j < ny + 2; ++j) {
        if (c.bc_x == LSC_BC_PERIODIC) {
            std::memcpy(at(a, ny, 0, j), at(a, ny, nx, j), cell);
            std::memcpy(at(a, ny, 1, j), at(a, ny, nx + 1, j), cell);
            std::memcpy(at(a, ny, nx + 2, j), at(a, ny, 2, j), cell);
            std::memcpy(at(a, ny, nx + 3, j), at(a, ny, 3, j), cell);
        } else {
            std::memcpy(at(a, ny, 0, j), at(a, ny, 2, j), cell);
            std::memcpy(at(a, ny, 1, j), at(a, ny, 3, j), cell);
            std::memcpy(at(a, ny, nx + 2, j), at(a, ny, nx, j), cell);
            std::memcpy(at(a, ny, nx + 3, j), at(a, ny, nx + 1, j), cell);
        }
    }
}

void mhd_apply_bc_y(const lsc_mhd_config& c, int ncols_minus2, double* a) {
    const int ny = c.ny;
    const size_t cell = 9 * sizeof(double);
    for (int i = 0; i < ncols_minus2; ++i) {
        if (c.bc_y == LSC_BC_PERIODIC) {
            std::memcpy(at(a, ny, i, 0), at(a, ny, i, ny), cell);
            std::memcpy(at(a, ny, i, 1), at(a, ny, i, ny + 1), cell);
            std::memcpy(at(a, ny, i, ny + 2), at(a, ny, i, 2), cell);
            std::memcpy(at(a, ny, i, ny + 3), at(a, ny, i, 3), cell);
        } else {
            std::memcpy(at(a, ny, i, 0), at(a, ny, i, 2), cell);
            std::memcpy(at(a, ny, i, 1), at(a, ny, i, 3), cell);
            std::memcpy(at(a, ny, i, ny + 2), at(a, ny, i, ny), cell);
            std::memcpy(at(a, ny, i, ny + 3), at(a, ny, i, ny + 1), cell);
            if (c.bc_y == LSC_BC_REFLECTIVE) {
                const int rows[4] = {0, 1, ny + 2, ny + 3};
                for (int r = 0; r < 4; ++r) {
                    double* g = at(a, ny, i, rows[r]);
                    g[2] = -g[2];
                    g[6] = -g[6];
                }
            }
        }
    }
}

// Ghost ring of the reference's `u` after MUSCL_NEW.cpp:971 given correct interior cells: the
// boundary function has by then been applied to the same interior twice (:958 and :971), which
// is its fixed point; the four cells [nx+2..nx+3][ny+2..ny+3] are never written by it and hold
// the zero-initialised uPlus1 values (:679-680).
void mhd_fill_ghost_ring(const lsc_mhd_config& c, double* a) {
    for (int pass = 0; pass < 2; ++pass) {
        apply_bc_x(c, a);
        mhd_apply_bc_y(c, c.nx + 2, a);
    }
    for (int i = c.nx + 2; i < c.nx + 4; ++i)
        for (int j = c.ny + 2; j < c.ny + 4; ++j) std::memset(at(a, c.ny, i, j), 0, 9 * sizeof(double));
}

// Writer of MUSCL_NEW.cpp:973-1001: ConservativeToPrimitive of every cell, rows j = 1..ny+3,
// columns i = 1..nx+3 (ghost ring included), default ostream formatting, blank line per row.
int mhd_write_dat(const lsc_mhd_config& c, const double* a, const char* path) {
    std::ofstream output(path);
    if (!output) return -1;
    const double dx = (c.x1 - c.x0) / c.nx;
    const double dy = (c.y1 - c.y0) / c.ny;
    const double gamma = c.gamma;
    for (int j = 1; j < c.ny + 4; j++) {
        for (int i = 1; i < c.nx + 4; i++) {
            const double* u = at(a, c.ny, i, j);
            double v[9];  // ConservativeToPrimitive, MUSCL_NEW.cpp:22-34
            v[0] = u[0];
            v[1] = u[1] / u[0];
            v[2] = u[2] / u[0];
            v[3] = u[3] / u[0];
            v[5] = u[5];
            v[6] = u[6];
            v[7] = u[7];
            v[4] = (u[4] - 0.5 * v[0] * (v[1] * v[1] + v[2] * v[2] + v[3] * v[3]) -
                    0.5 * (u[5] * u[5] + u[6] * u[6] + u[7] * u[7])) * (gamma - 1);
            v[8] = u[8];
            const double x = c.x0 + (i - 1) * dx;
            const double y = c.y0 + (j - 1) * dy;
            output << x << " " << y << " " << v[0] << " " << v[1] << " " << v[2] << " " << v[3] << " " << v[4]
                   << " " << std::sqrt(v[5] * v[5] + v[6] * v[6]) / v[7] << " " << v[6] << " " << v[7] << " "
                   << v[8] << "\n";
        }
        output << "\n";
    }
    return output.good() ? 0 : -1;
}

}  // namespace lsc
