// Instantiates the pair / full kernels for D = 8 (see dgmm_tables.cuh).
#include "dgmm_tables.cuh"
DGMM_DEFINE_DIM(8)
