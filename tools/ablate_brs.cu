// Timing ablations of k_blind_rotate_s on random data (no parity: most variants compute garbage on
// purpose).  Each line: what a component of the CMUX step costs on the real machine.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -o tools/ablate_brs tools/ablate_brs.cu
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <random>
#include "../cufhe_b200/csrc/k_blindrotate_s.cuh"
using namespace b2;
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e_), __LINE__); exit(1); } } while (0)

template <int G, int ABL, int KEYLD = 0>
static float run(const char* name, BrParams P, const TeamSTables* tabs, int reps = 2)
{
    auto kern = k_blind_rotate_s<G, false, ABL, KEYLD>;
    CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, brs_smem_bytes(G)));
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
    float best = 1e30f;
    const int grid = (P.count + G - 1) / G;
    for (int r = 0; r <= reps; r++) {
        CK(cudaEventRecord(a));
        kern<<<grid, 128 * G, brs_smem_bytes(G)>>>(P, tabs);
        CK(cudaGetLastError());
        CK(cudaEventRecord(b));
        CK(cudaEventSynchronize(b));
        float ms; CK(cudaEventElapsedTime(&ms, a, b));
        if (r && ms < best) best = ms;
    }
    printf("{\"probe\": \"ablate\", \"gates_per_cta\": %d, \"ablate_mask\": %d, \"what\": \"%s\", \"count\": %d, \"ms\": %.3f, \"gates_per_s\": %.0f}\n",
           G, ABL, name, P.count, best, P.count / best * 1e3);
    fflush(stdout);
    return best;
}

int main(int argc, char** argv)
{
    const int count = argc > 1 ? atoi(argv[1]) : 16384;
    static TeamSTables3 LS; static TeamSConst CS;
    make_tables16s(LS, CS);
    CK(cudaMemcpyToSymbol(c_s, &CS, sizeof CS));
    TeamSTables* tabs; CK(cudaMalloc(&tabs, sizeof LS)); CK(cudaMemcpy(tabs, &LS, sizeof LS, cudaMemcpyHostToDevice));   // the first two primes are a TeamSTables
    std::mt19937_64 rng(1);
    const size_t bkw = (size_t)kLwe * kRows * 2 * 2 * kN;
    std::vector<uint32_t> bk(bkw);
    for (size_t i = 0; i < bkw; i++) bk[i] = (uint32_t)(rng() % kP[(i / (2 * kN)) & 1]);
    std::vector<uint16_t> ct((size_t)count * kLvl0Words);
    for (auto& v : ct) v = (uint16_t)rng();
    uint32_t *dbk, *dout; uint16_t* dct;
    CK(cudaMalloc(&dbk, bkw * 4)); CK(cudaMemcpy(dbk, bk.data(), bkw * 4, cudaMemcpyHostToDevice));
    CK(cudaMalloc(&dct, ct.size() * 2)); CK(cudaMemcpy(dct, ct.data(), ct.size() * 2, cudaMemcpyHostToDevice));
    CK(cudaMalloc(&dout, (size_t)count * kLvl1Words * 4));
    BrParams P{};
    P.in0 = dct; P.out = dout; P.bk = dbk; P.count = count; P.ca = 1; P.cb = 0; P.off = 0; P.mu = kMu1;
#define R(A, name) run<4, A>(name, P, tabs)
    run<4, 0, 1>("key loads ld.global.cg", P, tabs);
    run<4, 0, 2>("key loads ld.global.cs", P, tabs);
    run<4, 0, 3>("key loads nc.L1::no_allocate", P, tabs);
    run<3, 0, 0>("3 gates per CTA (12 warps, more L1)", P, tabs);
    run<3, 0, 1>("3 gates per CTA, ld.global.cg", P, tabs);
    R(0, "full kernel");
    R(ABL_LUT, "conflict-free table look-ups");
    R(ABL_DIG, "no rotation/decomposition");
    R(ABL_KEY, "no key loads");
    R(ABL_BAR, "no barriers");
    R(ABL_TWD, "no twiddle loads");
    R(ABL_XCH, "no exchanges (still barriers)");
    R(ABL_LUT | ABL_DIG | ABL_KEY | ABL_BAR | ABL_XCH | ABL_TWD, "multiplies only (butterflies + key products + CRT)");
    R(ABL_LUT | ABL_KEY | ABL_TWD, "no conflicts, no key loads, no twiddle loads");
    return 0;
}
