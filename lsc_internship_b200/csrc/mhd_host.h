// mhd_host.h -- host-side helpers of the MHD solver surface (see mhd_host.cpp).
#pragma once
#include "../../include/lsc_b200.h"

namespace lsc {
void mhd_initial_condition(const lsc_mhd_config& c, int problem, int i_global0, int il0, int ni, double* aos);
void mhd_apply_bc_y(const lsc_mhd_config& c, int ncols_minus2, double* a);
void mhd_fill_ghost_ring(const lsc_mhd_config& c, double* a);
int mhd_write_dat(const lsc_mhd_config& c, const double* a, const char* path);
}  // namespace lsc
