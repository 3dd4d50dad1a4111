// k_keyswitch.cuh -- batched identity key switch lvl1 -> lvl0 with fused linear
// pre-add, plus the trivial elementwise gates.
//
// Replaces KeySwitchFromTLWE / KeySwitch / IdentityKeySwitchPreAdd (reference
// include/keyswitch_gpu.cuh:25-188): there one thread per OUTPUT column walks
// 1024 x 7 digits serially with scattered u16 loads.  Here one warp owns a gate,
// every selected key row is read as ten coalesced 128-byte lines, and two u16
// columns are accumulated per 32-bit add (exact: arithmetic is mod 2^16).
//
// Device key layout (built once by k_ksk_layout): T[j][d][3][320] u32, each word two
// u16 columns, rows padded to 640 columns and PRE-SIGNED so the kernel only adds:
//   slot 0: digit 0 (value -2) -> +ksk[j][d][1]
//   slot 1: digit 1 (value -1) -> +ksk[j][d][0]
//   slot 2: digit 3 (value +1) -> -ksk[j][d][0]        (digit 2 = value 0: skipped)
#pragma once
#include "dev_common.cuh"

namespace b2 {

constexpr int kKsRowWords = 320;                                   // 640 u16 columns
constexpr size_t kKskDevWords = (size_t)kN * kIksT * 3 * kKsRowWords;

__global__ void k_ksk_layout(const uint16_t* __restrict__ raw /*[1024][7][2][637]*/, uint32_t* __restrict__ T)
{
    const size_t total = kKskDevWords;
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
        const int w = (int)(idx % kKsRowWords);
        const size_t r = idx / kKsRowWords;
        const int slot = (int)(r % 3);
        const size_t jd = r / 3;
        const int src = slot == 0 ? 1 : 0;
        const uint16_t* row = raw + (jd * 2 + src) * kLvl0Words;
        uint32_t lo = 2 * w < kLvl0Words ? row[2 * w] : 0u;
        uint32_t hi = 2 * w + 1 < kLvl0Words ? row[2 * w + 1] : 0u;
        if (slot == 2) { lo = (0u - lo) & 0xFFFFu; hi = (0u - hi) & 0xFFFFu; }
        T[idx] = lo | (hi << 16);
    }
}

// key-switch work item: e = ca*A[s0] + cb*B[s1] + (0, off) over u32, out[slot] (637 u16)
struct KsItem {
    uint32_t s0, s1;
    int8_t ca, cb;
    int8_t pad[2];
    uint32_t off;
    uint32_t slot;
};
struct KsParams {
    const uint32_t* A;      // [*][1025]
    const uint32_t* B;      // may be null when cb == 0
    uint16_t* out;          // [*][637]
    const KsItem* items;    // null: uniform
    const uint32_t* T;      // device key
    int count;
    int ca, cb;
    uint32_t off;
    int trlwe;              // 1: A holds TRLWEs (2 x 1024 words); SampleExtract(0) is fused into the
                            //    read (SEIandKS, include/keyswitch_gpu.cuh:25-78); B must be null
};

// coefficient j of the lvl1 TLWE operand `x` (j = 1024: the body): plain TLWE or the
// on-the-fly SampleExtract(0) of a TRLWE [a(X)][b(X)]
__device__ __forceinline__ uint32_t ks_operand(const uint32_t* x, int j, int trlwe)
{
    if (!trlwe) return x[j];
    return j == 0 ? x[0] : j == kN ? x[kN] : 0u - x[kN - j];
}

// two u16 columns per word, each summed mod 2^16: ONE instruction (VIADD.16x2 on sm_100a) instead of a
// 32-bit add for the low column + a shifted add for the high one
__device__ __forceinline__ uint32_t add16x2(uint32_t a, uint32_t b)
{
    uint32_t r;
    asm("add.u16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
}

constexpr int kKsGates = 8;                       // warps (= gates) per CTA

__global__ void __launch_bounds__(32 * kKsGates) k_keyswitch(const KsParams P)
{
    __shared__ uint32_t s_t[kKsGates][kN];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int item = blockIdx.x * kKsGates + warp;
    if (item >= P.count) return;
    KsItem it;
    if (P.items) it = P.items[item];
    else { it.s0 = it.s1 = it.slot = (uint32_t)item; it.ca = (int8_t)P.ca; it.cb = (int8_t)P.cb; it.off = P.off; }
    const uint32_t* x0 = P.A + (size_t)it.s0 * (P.trlwe ? 2 * kN : kLvl1Words);
    const uint32_t* x1 = (it.cb != 0 && P.B) ? P.B + (size_t)it.s1 * kLvl1Words : nullptr;
    const uint32_t ca = (uint32_t)(int32_t)it.ca, cb = x1 ? (uint32_t)(int32_t)it.cb : 0u;
    // pre-add in u32 (keyswitch_gpu.cuh:158,168) + decomposition offset
    for (int j = lane; j < kN; j += 32) s_t[warp][j] = ca * ks_operand(x0, j, P.trlwe) + (x1 ? cb * x1[j] : 0u) + kIksAdd;
    const uint32_t eb = ca * ks_operand(x0, kN, P.trlwe) + (x1 ? cb * x1[kN] : 0u) + it.off;
    __syncwarp();

    uint32_t acc[10];
#pragma unroll
    for (int w = 0; w < 10; w++) acc[w] = 0u;
    const uint32_t* Tl = P.T + lane;
    for (int j = 0; j < kN; j++) {
        const uint32_t tj = s_t[warp][j];
#pragma unroll
        for (int d = 0; d < kIksT; d++) {
            const uint32_t dg = iks_digit(tj, d);
            if (dg != 2u) {
                const uint32_t* row = Tl + ((size_t)(j * kIksT + d) * 3 + min(dg, 2u)) * kKsRowWords;
#pragma unroll
                for (int w = 0; w < 10; w++) {
                    const uint32_t v = __ldg(row + 32 * w);
                    acc[w] = add16x2(acc[w], v);
                }
            }
        }
    }
    uint16_t* out = P.out + (size_t)it.slot * kLvl0Words;
#pragma unroll
    for (int w = 0; w < 10; w++) {
        const int c0 = 2 * (lane + 32 * w);
        if (c0 < kLwe) out[c0] = (uint16_t)acc[w];
        else if (c0 == kLwe) out[c0] = (uint16_t)(acc[w] + iks_round_b(eb));
        if (c0 + 1 < kLwe) out[c0 + 1] = (uint16_t)(acc[w] >> 16);
        else if (c0 + 1 == kLwe) out[c0 + 1] = (uint16_t)((acc[w] >> 16) + iks_round_b(eb));
    }
}

// Large-batch variant: W gates (warps) per CTA walk the 1024 mask coefficients IN STEP (a CTA
// barrier every SYNCJ coefficients), so the key rows of the current coefficients -- 21 candidate
// rows, 26.9 KB per coefficient, each wanted by W/4 of the warps on average -- are served to most
// of them by the SM's L1 instead of by L2.  Only bits 31..18 of the shifted coefficient select
// digits, so its upper half is staged (2 KB per gate).  Same sums as k_keyswitch.
template <int W, int SYNCJ>
__global__ void __launch_bounds__(32 * W) k_keyswitch_step(const KsParams P)
{
    extern __shared__ __align__(16) uint16_t s_th[];          // [W][1024]
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int item_raw = blockIdx.x * W + warp;
    const bool active = item_raw < P.count;
    const int item = active ? item_raw : P.count - 1;           // idle warps of the last CTA keep the barriers company
    KsItem it;
    if (P.items) it = P.items[item];
    else { it.s0 = it.s1 = it.slot = (uint32_t)item; it.ca = (int8_t)P.ca; it.cb = (int8_t)P.cb; it.off = P.off; }
    const uint32_t* x0 = P.A + (size_t)it.s0 * (P.trlwe ? 2 * kN : kLvl1Words);
    const uint32_t* x1 = (it.cb != 0 && P.B) ? P.B + (size_t)it.s1 * kLvl1Words : nullptr;
    const uint32_t ca = (uint32_t)(int32_t)it.ca, cb = x1 ? (uint32_t)(int32_t)it.cb : 0u;
    uint16_t* th = s_th + warp * kN;
    for (int j = lane; j < kN; j += 32)
        th[j] = (uint16_t)((ca * ks_operand(x0, j, P.trlwe) + (x1 ? cb * x1[j] : 0u) + kIksAdd) >> 16);
    const uint32_t eb = ca * ks_operand(x0, kN, P.trlwe) + (x1 ? cb * x1[kN] : 0u) + it.off;
    __syncthreads();

    uint32_t acc[10];
#pragma unroll
    for (int w = 0; w < 10; w++) acc[w] = 0u;
    const uint32_t* Tl = P.T + lane;
    for (int j0 = 0; j0 < kN; j0 += SYNCJ) {
        if (active) {
#pragma unroll
            for (int jj = 0; jj < SYNCJ; jj++) {
                const int j = j0 + jj;
                const uint32_t tj = (uint32_t)th[j] << 16;
#pragma unroll
                for (int d = 0; d < kIksT; d++) {
                    const uint32_t dg = iks_digit(tj, d);
                    if (dg != 2u) {
                        const uint32_t* row = Tl + ((size_t)(j * kIksT + d) * 3 + min(dg, 2u)) * kKsRowWords;
#pragma unroll
                        for (int w = 0; w < 10; w++) {
                            const uint32_t v = __ldg(row + 32 * w);
                            acc[w] = add16x2(acc[w], v);
                        }
                    }
                }
            }
        }
        __syncthreads();
    }
    if (!active) return;
    uint16_t* out = P.out + (size_t)it.slot * kLvl0Words;
#pragma unroll
    for (int w = 0; w < 10; w++) {
        const int c0 = 2 * (lane + 32 * w);
        if (c0 < kLwe) out[c0] = (uint16_t)acc[w];
        else if (c0 == kLwe) out[c0] = (uint16_t)(acc[w] + iks_round_b(eb));
        if (c0 + 1 < kLwe) out[c0 + 1] = (uint16_t)(acc[w] >> 16);
        else if (c0 + 1 == kLwe) out[c0 + 1] = (uint16_t)((acc[w] >> 16) + iks_round_b(eb));
    }
}

// Small-batch variant: one CTA of 16 warps per gate; warp w walks j in [64 w, 64 w + 64),
// then the 16 partial sums are added through shared memory.  Same arithmetic (mod 2^16
// per column), ~16x shorter critical path than one warp per gate.
// The walk is a chain of L2 round trips, so what matters is how many rows are in flight: the
// digit test is a PREDICATE on the loads (zero digits load nothing and add 0), not a branch, and
// the rows of DG digits (DG = 7: a whole coefficient, 70 words per thread) are requested before the
// first is added.  DG = 0 keeps the branch per digit (round-1 form, 10 loads in flight; A/B).
__device__ __forceinline__ uint32_t ldg_if(const uint32_t* p, uint32_t on)
{
    uint32_t v;
    asm("{ .reg .pred q; setp.ne.u32 q, %2, 0; mov.u32 %0, 0; @q ld.global.nc.u32 %0, [%1]; }" : "=r"(v) : "l"(p), "r"(on));
    return v;
}
constexpr int kKsLatWarps = 16;      // default; 32 for the smallest batches (half the chain of L2 round trips)
template <int DG, int WARPS = kKsLatWarps>
__global__ void __launch_bounds__(32 * WARPS) k_keyswitch_lat(const KsParams P)
{
    __shared__ uint32_t s_t[kN];
    __shared__ uint32_t s_red[WARPS][kKsRowWords];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int item = blockIdx.x;
    KsItem it;
    if (P.items) it = P.items[item];
    else { it.s0 = it.s1 = it.slot = (uint32_t)item; it.ca = (int8_t)P.ca; it.cb = (int8_t)P.cb; it.off = P.off; }
    const uint32_t* x0 = P.A + (size_t)it.s0 * (P.trlwe ? 2 * kN : kLvl1Words);
    const uint32_t* x1 = (it.cb != 0 && P.B) ? P.B + (size_t)it.s1 * kLvl1Words : nullptr;
    const uint32_t ca = (uint32_t)(int32_t)it.ca, cb = x1 ? (uint32_t)(int32_t)it.cb : 0u;
    for (int j = threadIdx.x; j < kN; j += blockDim.x) s_t[j] = ca * ks_operand(x0, j, P.trlwe) + (x1 ? cb * x1[j] : 0u) + kIksAdd;
    const uint32_t eb = ca * ks_operand(x0, kN, P.trlwe) + (x1 ? cb * x1[kN] : 0u) + it.off;
    __syncthreads();

    uint32_t acc[10];
#pragma unroll
    for (int w = 0; w < 10; w++) acc[w] = 0u;
    const uint32_t* Tl = P.T + lane;
    constexpr int per = kN / WARPS;
    for (int j = warp * per; j < (warp + 1) * per; j++) {
        const uint32_t tj = s_t[j];
        if (DG == 0) {
#pragma unroll
            for (int d = 0; d < kIksT; d++) {
                const uint32_t dg = iks_digit(tj, d);
                if (dg != 2u) {
                    const uint32_t* row = Tl + ((size_t)(j * kIksT + d) * 3 + min(dg, 2u)) * kKsRowWords;
#pragma unroll
                    for (int w = 0; w < 10; w++) acc[w] = add16x2(acc[w], __ldg(row + 32 * w));
                }
            }
        } else {
#pragma unroll
            for (int d0 = 0; d0 < kIksT; d0 += DG) {
                uint32_t v[DG ? DG : 1][10];
#pragma unroll
                for (int dd = 0; dd < DG; dd++) {
                    const int d = d0 + dd;
                    if (d < kIksT) {
                        const uint32_t dg = iks_digit(tj, d);
                        const uint32_t* row = Tl + ((size_t)(j * kIksT + d) * 3 + min(dg, 2u)) * kKsRowWords;
#pragma unroll
                        for (int w = 0; w < 10; w++) v[dd][w] = ldg_if(row + 32 * w, dg ^ 2u);
                    }
                }
#pragma unroll
                for (int dd = 0; dd < DG; dd++)
                    if (d0 + dd < kIksT) {
#pragma unroll
                        for (int w = 0; w < 10; w++) acc[w] = add16x2(acc[w], v[dd][w]);
                    }
            }
        }
    }
#pragma unroll
    for (int w = 0; w < 10; w++) s_red[warp][lane + 32 * w] = acc[w];
    __syncthreads();
    uint16_t* out = P.out + (size_t)it.slot * kLvl0Words;
    for (int c = threadIdx.x; c < kKsRowWords; c += blockDim.x) {
        uint32_t sum = 0u;
#pragma unroll
        for (int w = 0; w < WARPS; w++) sum = add16x2(sum, s_red[w][c]);
        const uint32_t lo = sum, hi = sum >> 16;
        const int c0 = 2 * c;
        if (c0 < kLwe) out[c0] = (uint16_t)lo;
        else if (c0 == kLwe) out[c0] = (uint16_t)(lo + iks_round_b(eb));
        if (c0 + 1 < kLwe) out[c0 + 1] = (uint16_t)hi;
        else if (c0 + 1 == kLwe) out[c0 + 1] = (uint16_t)(hi + iks_round_b(eb));
    }
}

// out[slot] = ca*A[s0] + cb*B[s1] + (0,..,0,off) elementwise, T = u16 (637 words) or
// u32 (1025 words).  Covers Not/Copy (src/bootstrap_gpu.cu:681-703) and the lvl1
// Mux/NMux combine (:718-741,756-780).
struct AffItem { uint32_t s0, s1; int8_t ca, cb; int8_t pad[2]; uint32_t off; uint32_t slot; };
template <typename T>
__global__ void k_affine(T* __restrict__ out, const T* __restrict__ A, const T* __restrict__ B, const AffItem* items,
                         int count, int words, int ca_u, int cb_u, uint32_t off_u)
{
    const int per = (words + 127) / 128;   // threads cover a ciphertext in chunks of 128
    (void)per;
    for (int item = blockIdx.x; item < count; item += gridDim.x) {
        AffItem it;
        if (items) it = items[item];
        else { it.s0 = it.s1 = it.slot = (uint32_t)item; it.ca = (int8_t)ca_u; it.cb = (int8_t)cb_u; it.off = off_u; }
        const T* x0 = A + (size_t)it.s0 * words;
        const T* x1 = (it.cb != 0 && B) ? B + (size_t)it.s1 * words : nullptr;
        T* o = out + (size_t)it.slot * words;
        for (int i = threadIdx.x; i < words; i += blockDim.x) {
            uint32_t v = (uint32_t)(int32_t)it.ca * (uint32_t)x0[i];
            if (x1) v += (uint32_t)(int32_t)it.cb * (uint32_t)x1[i];
            if (i == words - 1) v += it.off;
            o[i] = (T)v;
        }
    }
}

}  // namespace b2
