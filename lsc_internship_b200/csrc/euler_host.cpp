// euler_host.cpp -- host-side pieces of the Euler SLIC solver surface that must match the reference
// bit-for-bit / byte-for-byte: initial conditions (SLIC_euler_2d.cpp:552-596, generaliseSLIC.cpp:586-622),
// the ghost ring of a downloaded state (S:198-226, G:199-215) and the .dat writers (S:621-642, G:645-667).
#include "euler_host.h"

#include <cmath>
#include <cstring>
#include <algorithm>
#include <cstdlib>
#include <fstream>
#include <string>
#include <vector>

namespace lsc {

namespace {
inline double* at(double* a, int ny, int i, int j) { return a + ((size_t)i * (size_t)(ny + 4) + (size_t)j) * 4; }
inline const double* at(const double* a, int ny, int i, int j) {
    return a + ((size_t)i * (size_t)(ny + 4) + (size_t)j) * 4;
}
}  // namespace

// Slab of rank r of P along x: interior columns [b0, b1) of the global grid.
void euler_slab_range(const lsc_euler_config& c, int& b0, int& b1) {
    const int P = c.nranks > 1 ? c.nranks : 1, r = c.nranks > 1 ? c.rank : 0;
    b0 = (int)((long long)c.nx * r / P);
    b1 = (int)((long long)c.nx * (r + 1) / P);
}

// One application of the reference's boundary function (full loop ranges, x loop then y loop) to a slab of
// nxl interior columns: the x images only where the slab ends at the physical boundary -- elsewhere the two
// columns on that side are a neighbour's interior cells and stay as they are.
void euler_fill_ghost_ring(const lsc_euler_config& c, double* a) {
    int b0, b1;
    euler_slab_range(c, b0, b1);
    const int nx = b1 - b0, ny = c.ny;
    const bool whole = (nx == c.nx);
    const size_t cell = 4 * sizeof(double);
    const bool per = (c.bc == LSC_BC_PERIODIC);
    for (int j = 0; j < ny + 4; ++j) {
        if (whole || (!per && b0 == 0)) {
            std::memcpy(at(a, ny, 0, j), at(a, ny, per ? nx : 2, j), cell);
            std::memcpy(at(a, ny, 1, j), at(a, ny, per ? nx + 1 : 2, j), cell);
        }
        if (whole || (!per && b1 == c.nx)) {
            std::memcpy(at(a, ny, nx + 2, j), at(a, ny, per ? 2 : nx + 1, j), cell);
            std::memcpy(at(a, ny, nx + 3, j), at(a, ny, per ? 3 : nx + 1, j), cell);
        }
    }
    for (int i = 0; i < nx + 4; ++i) {
        std::memcpy(at(a, ny, i, 0), at(a, ny, i, per ? ny : 2), cell);
        std::memcpy(at(a, ny, i, 1), at(a, ny, i, per ? ny + 1 : 2), cell);
        std::memcpy(at(a, ny, i, ny + 2), at(a, ny, i, per ? 2 : ny + 1), cell);
        std::memcpy(at(a, ny, i, ny + 3), at(a, ny, i, per ? 3 : ny + 1), cell);
    }
}

// IC loop + the boundary fill that follows it (S:552-596).  Cell centres x0 + (i-1.5)dx (S:556-557), i the
// GLOBAL column index.  For a slab the array holds columns b0 .. b1+3 of the global padded grid; its halo
// columns are the neighbour's interior cells, i.e. the same formula at their own coordinates, except at a
// periodic wrap where they are what the reference's boundary fill would have copied there (the formula at the
// coordinates of the source column).
void euler_initial_condition(const lsc_euler_config& c, int problem, double* aos) {
    const double pi = 3.14159265358979323846;  // S:540
    const double gamma = c.gamma;
    const double dx = (c.x1 - c.x0) / c.nx;
    const double dy = (c.y1 - c.y0) / c.ny;
    int b0, b1;
    euler_slab_range(c, b0, b1);
    const int nxl = b1 - b0;
    const bool slab = (nxl != c.nx);
    for (int il = 0; il < nxl + 4; ++il) {
        int i = b0 + il;  // global padded column
        if (slab && c.bc == LSC_BC_PERIODIC) {
            if (i < 2) i += c.nx;
            else if (i > c.nx + 1) i -= c.nx;
        }
        for (int j = 0; j < c.ny + 4; ++j) {
            const double x = c.x0 + (i - 1.5) * dx;
            const double y = c.y0 + (j - 1.5) * dy;
            double prim[4];
            if (problem == LSC_EULER_IC_ORSZAG_TANG) {  // S:585-589
                prim[0] = gamma * gamma;
                prim[1] = -std::sin(2.0 * pi * y);
                prim[2] = std::sin(2.0 * pi * x);
                prim[3] = gamma;
            } else if (problem == LSC_EULER_IC_QUADRANTS) {  // G:594-617
                if (x > 0.5 && y > 0.5) {
                    prim[0] = 1.5; prim[1] = 0; prim[2] = 0; prim[3] = 1.5;
                } else if (x <= 0.5 && y > 0.5) {
                    prim[0] = 0.5323; prim[1] = 1.206; prim[2] = 0; prim[3] = 0.3;
                } else if (x <= 0.5 && y <= 0.5) {
                    prim[0] = 0.138; prim[1] = 1.206; prim[2] = 1.206; prim[3] = 0.029;
                } else {
                    prim[0] = 0.5323; prim[1] = 0; prim[2] = 1.206; prim[3] = 0.3;
                }
            } else {  // cylindrical explosion of BASELINE config 3 (not in the reference)
                if ((x - 0.5) * (x - 0.5) + (y - 0.5) * (y - 0.5) <= 0.2 * 0.2) {
                    prim[0] = 1.0; prim[1] = 0; prim[2] = 0; prim[3] = 1.0;
                } else {
                    prim[0] = 0.125; prim[1] = 0; prim[2] = 0; prim[3] = 0.1;
                }
            }
            double* u = at(aos, c.ny, il, j);  // PrimativeToConservative, S:10-17
            u[0] = prim[0];
            u[1] = prim[0] * prim[1];
            u[2] = prim[0] * prim[2];
            u[3] = prim[3] / (gamma - 1) + (0.5 * prim[0] * (prim[1] * prim[1] + prim[2] * prim[2]));
        }
    }
    euler_fill_ghost_ring(c, aos);  // S:596
}

// S:621-642 (no blank lines) / G:645-667 (blank line after each j row): rows and columns 1..n+1.
int euler_write_dat(const lsc_euler_config& c, const double* a, const char* path, int blank_line_per_row) {
    std::ofstream output(path);
    if (!output) return -1;
    const double dx = (c.x1 - c.x0) / c.nx;
    const double dy = (c.y1 - c.y0) / c.ny;
    for (int j = 1; j < c.ny + 2; j++) {
        for (int i = 1; i < c.nx + 2; i++) {
            const double* u = at(a, c.ny, i, j);
            double prim[4];  // ConservativeToPrimative, S:19-26
            prim[0] = u[0];
            prim[1] = u[1] / u[0];
            prim[2] = u[2] / u[0];
            prim[3] = (c.gamma - 1) * (u[3] - 0.5 * prim[0] * (prim[1] * prim[1] + prim[2] * prim[2]));
            const double x = c.x0 + (i - 1) * dx;
            const double y = c.y0 + (j - 1) * dy;
            output << x << " " << y << " " << prim[0] << " " << prim[1] << " " << prim[2] << " " << prim[3] << "\n";
        }
        if (blank_line_per_row) output << "\n";
    }
    return output.good() ? 0 : -1;
}

// ---- tabulated EoS on the host (MHD_codes/TabEoS.cpp = T): table file, IC and .dat output -----------------
namespace {
std::vector<double> number_row(const std::string& line) {
    std::vector<double> v;
    const char* p = line.c_str();
    char* end = nullptr;
    for (;;) {
        const double x = std::strtod(p, &end);  // what `iss >> val` (T:42-45) parses: correctly rounded decimal
        if (end == p) break;
        v.push_back(x);
        p = end;
    }
    return v;
}

// the bisection of T:174-191 (point2 starts at 499)
void bracket499(const double* a, double v, int& lo, int& hi) {
    int point1 = 0, point2 = 499;
    while (std::abs(point1 - point2) > 1) {
        const int half = (point1 + point2) / 2;
        if (v > a[half]) point1 = half;
        else point2 = half;
    }
    lo = std::min(std::max(point1, 0), 499);
    hi = std::min(std::max(point2, 0), 499);
}
}  // namespace

// Plasma19(), T:18-133: line 10 = densities, 12 = pressures, 14-513 = sound speed rows, 515-1014 = specific
// internal energy rows (row = pressure index, column = density index); the temperature / conductivity blocks
// (T:89-125) are never used by the solver and are skipped.  Returns 0, or the 1-based number of the bad line.
int eos_table_read(const char* path, double* dens, double* pres, double* sound, double* en) {
    std::ifstream file(path);
    if (!file) return -1;
    std::string line;
    int row = 0, seen = 0;
    while (std::getline(file, line)) {
        row++;
        double* dst = nullptr;
        if (row == 10) dst = dens;
        else if (row == 12) dst = pres;
        else if (row >= 14 && row < 514) dst = sound + (size_t)(row - 14) * 500;
        else if (row >= 515 && row < 1015) dst = en + (size_t)(row - 515) * 500;
        if (!dst) continue;
        const std::vector<double> v = number_row(line);
        if (v.size() != 500) return row;  // the reference checks this for the energy rows (T:83-86)
        std::memcpy(dst, v.data(), 500 * sizeof(double));
        ++seen;
    }
    return seen == 1002 ? 0 : row + 1;
}

// PrimativeToConservative, T:166-246
void tabeos_p2c(const double* dens, const double* pres, const double* en, const double* prim, double* consv) {
    consv[0] = prim[0];
    consv[1] = prim[0] * prim[1];
    consv[2] = prim[0] * prim[2];
    const double rho = prim[0], p = prim[3];
    int db, dt, pb, pt;
    bracket499(dens, rho, db, dt);
    double density_ratio = (rho - dens[db]) / (dens[dt] - dens[db]);
    if (dens[db] == dens[dt]) density_ratio = 1;
    bracket499(pres, p, pb, pt);
    double pressure_ratio = (p - pres[pb]) / (pres[pt] - pres[pb]);
    if (pres[pb] == pres[pt]) pressure_ratio = 1;
    const double f_NE = en[pt * 500 + dt], f_SE = en[pb * 500 + dt], f_SW = en[pb * 500 + db], f_NW = en[pt * 500 + db];
    const double f_x_y1 = (1 - density_ratio) * f_SW + density_ratio * f_SE;  // T:147-153
    const double f_x_y2 = (1 - density_ratio) * f_NW + density_ratio * f_NE;
    const double energy = (1 - pressure_ratio) * f_x_y1 + pressure_ratio * f_x_y2;
    consv[3] = energy * rho + 0.5 * rho * (prim[1] * prim[1] + prim[2] * prim[2]);
}

// ConservativeToPrimative, T:248-327
void tabeos_c2p(const double* dens, const double* pres, const double* en, const double* consv, double* prim) {
    prim[0] = consv[0];
    prim[1] = consv[1] / consv[0];
    prim[2] = consv[2] / consv[0];
    const double rho = prim[0];
    int db, dt;
    bracket499(dens, rho, db, dt);
    double density_ratio = (rho - dens[db]) / (dens[dt] - dens[db]);
    if (dens[db] == dens[dt]) density_ratio = 1;
    const double e = (consv[3] - 0.5 * rho * (prim[1] * prim[1] + prim[2] * prim[2])) / rho;
    int point1 = 0, point2 = 499;
    while (std::abs(point1 - point2) > 1) {
        const int half = (point1 + point2) / 2;
        if (e > (en[half * 500 + db] + density_ratio * (en[half * 500 + dt] - en[half * 500 + db]))) point1 = half;
        else point2 = half;
    }
    const int pt = point2, pb = point1;
    const double f_NE = en[pt * 500 + dt], f_SE = en[pb * 500 + dt], f_SW = en[pb * 500 + db], f_NW = en[pt * 500 + db];
    const double f_x_y1 = (1 - density_ratio) * f_SW + density_ratio * f_SE;  // T:156-163
    const double f_x_y2 = (1 - density_ratio) * f_NW + density_ratio * f_NE;
    double y = (e - f_x_y1) / (f_x_y2 - f_x_y1);
    if (f_x_y2 == f_x_y1) y = 1;
    prim[3] = y * (pres[pt] - pres[pb]) + pres[pb];
}

// IC of T's main(), T:918-941 (four quadrants at rest, rho = 1, p = 0.01e5 for x > 0.5, 1000e5 otherwise) through
// the table's PrimativeToConservative, then the boundary fill T:954.  Slab-aware like euler_initial_condition.
void tabeos_initial_condition(const lsc_euler_config& c, const double* dens, const double* pres, const double* en,
                              double* aos) {
    const double dx = (c.x1 - c.x0) / c.nx;
    const double dy = (c.y1 - c.y0) / c.ny;
    int b0, b1;
    euler_slab_range(c, b0, b1);
    const int nxl = b1 - b0;
    for (int il = 0; il < nxl + 4; ++il) {
        const int i = b0 + il;
        for (int j = 0; j < c.ny + 4; ++j) {
            const double x = c.x0 + (i - 1.5) * dx;
            const double y = c.y0 + (j - 1.5) * dy;
            double prim[4] = {1.0, 0, 0, 0};
            if (x > 0.5 && y > 0.5) prim[3] = 0.01 * 1e5;
            else if (x <= 0.5 && y > 0.5) prim[3] = 1000 * 1e5;
            else if (x <= 0.5 && y <= 0.5) prim[3] = 1000 * 1e5;
            else prim[3] = 0.01 * 1e5;
            tabeos_p2c(dens, pres, en, prim, at(aos, c.ny, il, j));
        }
    }
    euler_fill_ghost_ring(c, aos);
}

// T:974-999: rows and columns 1..n+1 through the table's ConservativeToPrimative, std::endl after every record
// and a blank line after each j row.
int tabeos_write_dat(const lsc_euler_config& c, const double* a, const char* path, const double* dens,
                     const double* pres, const double* en) {
    std::ofstream output(path);
    if (!output) return -1;
    const double dx = (c.x1 - c.x0) / c.nx;
    const double dy = (c.y1 - c.y0) / c.ny;
    for (int j = 1; j < c.ny + 2; j++) {
        for (int i = 1; i < c.nx + 2; i++) {
            double prim[4];
            tabeos_c2p(dens, pres, en, at(a, c.ny, i, j), prim);
            const double x = c.x0 + (i - 1) * dx;
            const double y = c.y0 + (j - 1) * dy;
            output << x << " " << y << " " << prim[0] << " " << prim[1] << " " << prim[2] << " " << prim[3] << std::endl;
        }
        output << std::endl;
    }
    return output.good() ? 0 : -1;
}

bool eos_table_is_monotone(const double* en) {
    constexpr int N = 500;
    const double tol = std::ldexp(1.0, -49);  // 16 u: the three roundings of the interpolation are < 4 u * mag
    for (int p = 0; p + 1 < N; ++p) {
        for (int d = 0; d < N; ++d) {
            const int d2 = (d + 1 < N) ? d + 1 : d;
            const double a0 = en[p * N + d], a1 = en[(p + 1) * N + d];
            const double b0 = en[p * N + d2], b1 = en[(p + 1) * N + d2];
            const double gap = std::fmin(a1 - a0, b1 - b0);
            const double mag = std::fmax(std::fmax(std::fabs(a0), std::fabs(a1)), std::fmax(std::fabs(b0), std::fabs(b1))) +
                               2.0 * std::fmax(std::fabs(b0 - a0), std::fabs(b1 - a1));
            if (!(gap > tol * mag)) return false;  // also false for NaN
        }
    }
    return true;
}

}  // namespace lsc
