// Stand-alone translation unit with only the second-generation blind-rotation kernel, for fast
// SASS inspection (tools/sass_hist.py, tools/sass_stalls.py):
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -cubin -o /tmp/brs.cubin tools/sass_brs.cu
#include "../cufhe_b200/csrc/k_blindrotate_s.cuh"
#ifndef GATES
#define GATES 4
#endif
#ifndef ABLM
#define ABLM 0
#endif
template __global__ void b2::k_blind_rotate_s<GATES, false, ABLM>(const b2::BrParams, const b2::TeamSTables*);
