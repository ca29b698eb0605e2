// Instantiates the cooperative (team) kernels for D = 32 (see dgmm_team.cuh).
#include "dgmm_team.cuh"
DGMM_DEFINE_TEAM_DIM(32)
