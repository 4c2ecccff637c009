// mhd_kernels_relaxed.cu -- the "relaxed arithmetic" instantiation of the two sweeps.
//
// Same source as mhd_kernels.cu, but this translation unit is compiled WITH FMA contraction (-fmad=true) and the
// sweeps are instantiated with RelaxedDiv (one multiplication per quotient, no operand guards, no IEEE fix
// kernels).  It is selected per context with lsc_mhd_set_arithmetic(ctx, LSC_ARITH_RELAXED); it is never the
// default and never enters a parity test except to have its deviation from the reference measured
// (tests/test_mhd_gpu.py::test_relaxed_arithmetic_deviation).
#define LSC_RELAXED_TU 1
#include "mhd_kernels.cu"
