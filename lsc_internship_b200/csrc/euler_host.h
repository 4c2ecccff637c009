// euler_host.h -- host-side helpers of the Euler solver surface (see euler_host.cpp).
#pragma once
#include "../../include/lsc_b200.h"

namespace lsc {
void euler_slab_range(const lsc_euler_config& c, int& b0, int& b1);
void euler_initial_condition(const lsc_euler_config& c, int problem, double* aos);
void euler_fill_ghost_ring(const lsc_euler_config& c, double* a);
int euler_write_dat(const lsc_euler_config& c, const double* a, const char* path, int blank_line_per_row);
// tabulated EoS (MHD_codes/TabEoS.cpp): table file reader, IC, .dat writer
int eos_table_read(const char* path, double* dens, double* pres, double* sound, double* en);
// True when every energy column en[p][d] ascends in p by more than the rounding of the density interpolation
// fl(eb + fl(r*fl(et - eb))), 0 <= r <= 1, for every pair of adjacent density columns: then the comparand of the
// energy bisection (T:293-296) is monotone in the pressure index and the bracket the bisection ends in is unique
// (csrc/euler_physics.cuh: TableEos).
bool eos_table_is_monotone(const double* en);
void tabeos_initial_condition(const lsc_euler_config& c, const double* dens, const double* pres, const double* en,
                              double* aos);
int tabeos_write_dat(const lsc_euler_config& c, const double* a, const char* path, const double* dens,
                     const double* pres, const double* en);
}  // namespace lsc
