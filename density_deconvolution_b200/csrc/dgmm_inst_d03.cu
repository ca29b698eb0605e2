// Instantiates the pair / full kernels for D = 3 (see dgmm_tables.cuh).
#include "dgmm_tables.cuh"
DGMM_DEFINE_DIM(3)
