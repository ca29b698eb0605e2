// Instantiates the pair / full kernels for D = 27 (see dgmm_tables.cuh).
#include "dgmm_tables.cuh"
DGMM_DEFINE_DIM(27)
