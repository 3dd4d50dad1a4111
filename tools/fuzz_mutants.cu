// Does the schedule fuzzing (ABL_JIT builds of k_blind_rotate_s, tests/test_gpu_parity.py::
// test_schedule_jitter_is_bit_exact) notice a missing synchronisation?  Mutation test: for each of the ten
// synchronisation points of a CMUX step of k_blind_rotate_s (and the five in-loop ones of the cluster kernel
// k_blind_rotate_cl4s: __syncthreads x 3, mbarrier wait x 2), build the kernel WITHOUT it, with and without the random sleeps,
// run it on random ciphertexts and count the gates whose output differs from the unmodified kernel.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -o tools/fuzz_mutants tools/fuzz_mutants.cu
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <random>
#include "../cufhe_b200/csrc/k_blindrotate_cl4s.cuh"
using namespace b2;
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e_), __LINE__); exit(1); } } while (0)

static std::vector<uint32_t> g_ref;
static const char* kSite[11] = {"none", "fwd A->D: owners have read (lead)", "fwd A->D: stores visible", "fwd D->C: mates have read (__syncwarp)",
                                "fwd D->C: stores visible (__syncwarp)", "inv C->D: mates have read (__syncwarp)", "inv C->D: stores visible (__syncwarp)",
                                "inv D->A: owners have read", "inv D->A: stores visible", "hand-over: residues visible (128)", "step end: accumulator visible (128)"};

template <int ABL>
static void run(BrParams P, const TeamSTables* tabs, int reps)
{
    auto kern = k_blind_rotate_s<4, false, ABL>;
    CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, brs_smem_bytes(4)));
    std::vector<uint32_t> out((size_t)P.count * kLvl1Words);
    int worst = 0;
    for (int r = 0; r < reps; r++) {
        CK(cudaMemset(P.out, 0, out.size() * 4));
        kern<<<(P.count + 3) / 4, 512, brs_smem_bytes(4)>>>(P, tabs);
        CK(cudaGetLastError());
        CK(cudaDeviceSynchronize());
        CK(cudaMemcpy(out.data(), P.out, out.size() * 4, cudaMemcpyDeviceToHost));
        if (g_ref.empty()) { g_ref = out; continue; }
        int bad = 0;
        for (int g = 0; g < P.count; g++) bad += memcmp(&out[(size_t)g * kLvl1Words], &g_ref[(size_t)g * kLvl1Words], kLvl1Words * 4) != 0;
        if (bad > worst) worst = bad;
    }
    const int site = (ABL >> kMutShift) & 15;
    printf("{\"probe\": \"fuzz_mutant\", \"dropped_sync\": %d, \"site\": \"%s\", \"random_sleeps\": %s, \"gates\": %d, \"launches\": %d, \"max_gates_wrong\": %d}\n",
           site, kSite[site], (ABL & ABL_JIT) ? "true" : "false", P.count, reps, worst);
    fflush(stdout);
}

static std::vector<uint32_t> g_ref_cl;
static const char* kSiteCl[6] = {"none", "__syncthreads: transformed digits visible", "mbarrier wait: partial sum received", "__syncthreads: my inverse result visible",
                                 "mbarrier wait: other prime's inverse result received", "__syncthreads: accumulator visible (step end)"};
template <bool JIT, int MUT>
static void run_cl(BrParams P, const PrimeSTables* tabs, int reps)
{
    auto kern = k_blind_rotate_cl4s<JIT, MUT>;
    CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kCl4SmemBytes));
    std::vector<uint32_t> out((size_t)P.count * kLvl1Words);
    int worst = 0;
    for (int r = 0; r < reps; r++) {
        CK(cudaMemset(P.out, 0, out.size() * 4));
        kern<<<P.count * kCl4Ctas, kCl4Threads, kCl4SmemBytes>>>(P, tabs);
        CK(cudaGetLastError());
        CK(cudaDeviceSynchronize());
        CK(cudaMemcpy(out.data(), P.out, out.size() * 4, cudaMemcpyDeviceToHost));
        if (g_ref_cl.empty()) { g_ref_cl = out; continue; }
        int bad = 0;
        for (int g = 0; g < P.count; g++) bad += memcmp(&out[(size_t)g * kLvl1Words], &g_ref_cl[(size_t)g * kLvl1Words], kLvl1Words * 4) != 0;
        if (bad > worst) worst = bad;
    }
    printf("{\"probe\": \"fuzz_mutant_cl4s\", \"dropped_sync\": %d, \"site\": \"%s\", \"random_sleeps\": %s, \"gates\": %d, \"launches\": %d, \"max_gates_wrong\": %d}\n",
           MUT, kSiteCl[MUT], JIT ? "true" : "false", P.count, reps, worst);
    fflush(stdout);
}

int main(int argc, char** argv)
{
    const int count = argc > 1 ? atoi(argv[1]) : 1184;
    static TeamSTables3 LS; static TeamSConst CS;
    make_tables16s(LS, CS);
    CK(cudaMemcpyToSymbol(c_s, &CS, sizeof CS));
    TeamSTables* tabs; CK(cudaMalloc(&tabs, sizeof LS)); CK(cudaMemcpy(tabs, &LS, sizeof LS, cudaMemcpyHostToDevice));
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
    run<0>(P, tabs, 1);                 // reference bits
    run<0>(P, tabs, 2);                 // the unmodified kernel repeats them
    run<ABL_JIT>(P, tabs, 3);           // ... and so does its fuzzing build
#define MUT(S) run<(S << kMutShift)>(P, tabs, 3); run<ABL_JIT | (S << kMutShift)>(P, tabs, 3)
    MUT(1); MUT(2); MUT(3); MUT(4); MUT(5); MUT(6); MUT(7); MUT(8); MUT(9); MUT(10);
    // ---- the cluster kernel: 66 gates = two CTAs per SM on most SMs; the throughput kernel's bits are the reference ----
    P.count = 66;
    g_ref_cl.assign(g_ref.begin(), g_ref.begin() + (size_t)P.count * kLvl1Words);
    const PrimeSTables* ptabs = reinterpret_cast<const PrimeSTables*>(tabs);
    run_cl<false, 0>(P, ptabs, 2);
    run_cl<true, 0>(P, ptabs, 3);
#define MUTC(S) run_cl<false, S>(P, ptabs, 3); run_cl<true, S>(P, ptabs, 3)
    // the two mbarrier-wait mutants exist (MUT = 2, 4) but are not run: without a wait a receiver can re-arm an mbarrier whose
    // phase is still open, which is undefined and may hang the GPU -- not a risk to take on a shared box
    MUTC(1); MUTC(3); MUTC(5);
    return 0;
}
