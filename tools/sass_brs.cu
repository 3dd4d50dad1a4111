// Stand-alone translation unit with only the second-generation blind-rotation kernel, for fast
// SASS inspection (tools/sass_hist.py):  nvcc -gencode arch=compute_100a,code=sm_100a -cubin ...
#include "../cufhe_b200/csrc/k_blindrotate_s.cuh"
#ifndef SOLM
#define SOLM 0
#endif
#ifndef GATES
#define GATES 4
#endif
template __global__ void b2::k_blind_rotate_s<GATES, SOLM, false>(const b2::BrParams, const b2::TeamSTables*);
