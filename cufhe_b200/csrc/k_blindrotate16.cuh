// k_blindrotate16.cuh -- batched blind rotation + SampleExtract(0), 128 threads per gate.
//
// Same mathematics, key-stationary batching and fused head/tail as k_blindrotate.cuh
// (replaces __BlindRotatePreAdd__/__BlindRotate__/Accumulate, reference
// include/gatebootstrapping_gpu.cuh:115-345, and __SampleExtractIndex__,
// src/bootstrap_gpu.cu:366-381), but a gate is FOUR warps: two 64-thread teams, one per
// NTT prime, each thread holding 16 coefficients (ntt16.cuh).  Per-thread state is
// 16 + 2 x 16 words instead of 32 + 2 x 32, so a CTA of G = 6 gates keeps 24 warps
// (6 per scheduler) resident at 80 registers instead of 12 warps at 168 -- the latency
// hiding the 32-coefficient kernel lacked (ncu: 0.51 issue/cycle, top stalls `wait`,
// short/long scoreboard on twiddle and key loads).
//
// Per CMUX step each team
//   1. rebuilds the four gadget-digit polynomials from the accumulator (shared memory),
//   2. forward-transforms each (3 in-register passes, 2 padded exchanges),
//   3. multiply-accumulates with the key row into 2 x 16 lazy residues,
//   4. inverse-transforms both outputs,
// then the teams swap one residue polynomial, each finishes the centred CRT of one output
// polynomial and updates the accumulator.  All synchronisation is the gate's own
// 128-thread named barrier (14 per step); gates of a CTA never wait for each other.
//
// Template parameter PAIR selects the schedule of those four phases (identical bits):
//   0  one TRGSW row at a time, as listed above (first version; b2fhe_set_br_variant(16));
//   3  the DEFAULT: the two digits of an accumulator polynomial are transformed together
//      (fwd16_digits2: one through exchange buffer A, one through B -- 16 independent butterflies
//      per stage per thread, twiddles fetched once), their two key rows share ONE Montgomery
//      reduction (umontmul2_p), and the two inverse transforms run together (inv16_2);
//   1, 2 intermediate steps, 4 = all four rows in one reduction (four transformed polynomials
//      live at once: spills at 128 registers) -- kept for A/B, see DESIGN.md 4.1.
#pragma once
#include "k_blindrotate.cuh"
#include "ntt16.cuh"

namespace b2 {

// shared-memory table accessors (Lane16Tables staged verbatim)
__device__ __forceinline__ const Tw* t16_fw2(const uint32_t* s, int q, int B) { return reinterpret_cast<const Tw*>(s) + q * 15 * 16 + B; }
__device__ __forceinline__ const Tw* t16_fw3(const uint32_t* s, int q, int T) { return reinterpret_cast<const Tw*>(s) + 2 * 15 * 16 + q * 12 * 64 + T; }
__device__ __forceinline__ const Tw* t16_iv2(const uint32_t* s, int q, int lo) { return reinterpret_cast<const Tw*>(s) + 2 * 15 * 16 + 2 * 12 * 64 + q * 15 * 16 + lo; }
__device__ __forceinline__ const Tw* t16_iv3(const uint32_t* s, int q, int T) { return reinterpret_cast<const Tw*>(s) + 4 * 15 * 16 + 2 * 12 * 64 + q * 12 * 64 + T; }
__device__ __forceinline__ const Tw* t16_scale(const uint32_t* s, int q, int T) { return reinterpret_cast<const Tw*>(s) + 4 * 15 * 16 + 4 * 12 * 64 + q * 16 * 64 + T; }

__device__ __forceinline__ const uint32_t* t16_lut1(const uint32_t* s, int q) { return s + kLut1Word + q * 256; }
__device__ __forceinline__ const U2* t16_lut24(const uint32_t* s, int q) { return reinterpret_cast<const U2*>(s + kLut1Word + 512) + q * 256; }
__device__ __forceinline__ const U2* t16_lut35(const uint32_t* s, int q) { return reinterpret_cast<const U2*>(s + kLut1Word + 512 + 1024) + q * 256; }

__device__ __forceinline__ void stage_tables16(uint32_t* s_tab, const Lane16Tables* g_tab)
{
    const uint4* src = reinterpret_cast<const uint4*>(g_tab);
    uint4* dst = reinterpret_cast<uint4*>(s_tab);
    for (int i = threadIdx.x; i < kLane16TableWords / 4; i += blockDim.x) dst[i] = __ldg(src + i);
}

// thread-constant exchange bases (word offsets into a padded buffer)
struct X16 {
    int wA, rB;      // A -> B : store T + 68 k            load 68 (T>>2) + (T&3) + 4 k
    int wB, rC;      // B -> C : store wB + 4 k + (k>>2)   load 17 T + k
    int rD;          // C -> D : store 17 T + k            load rD + 17 k
    int wD;          // D -> A : store wD + 16 k           load T + 64 k + 16 (k>>2)
};
__device__ __forceinline__ X16 make_x16(int T)
{
    X16 x;
    x.wA = T;
    x.rB = 68 * (T >> 2) + (T & 3);
    x.wB = 68 * (T >> 2) + (T & 3);           // padC(64 B + j + 4 k) = 68 B + j + 4 k + (k >> 2)
    x.rC = 17 * T;
    x.rD = 272 * (T >> 4) + (T & 15);         // padC(256 hi + lo + 16 k) = 272 hi + lo + 17 k
    x.wD = 272 * (T >> 4) + (T & 15);         // padDA(256 hi + lo + 16 k) = 272 hi + lo + 16 k
    return x;
}

// forward transform: layout A in, layout C out.  bufA / bufB: the team's two exchange
// buffers; `bar` = the gate's named barrier (all 128 threads of the gate call this).
// passes F2 and F3 (after the first pass left layout A in x)
__device__ __forceinline__ void fwd16_tail(uint32_t (&x)[16], uint32_t* bufA, uint32_t* bufB, const uint32_t* s_tab,
                                           const X16& a, int T, int q, uint32_t p, int bar, int barn)
{
#pragma unroll
    for (int k = 0; k < 16; k++) bufA[a.wA + 68 * k] = x[k];
    named_bar_sync(bar, barn);
#pragma unroll
    for (int k = 0; k < 16; k++) x[k] = bufA[a.rB + 4 * k];
    pass16_fwd<4, 8>(x, t16_fw2(s_tab, q, T >> 2), 16, p);
#pragma unroll
    for (int k = 0; k < 16; k++) bufB[a.wB + 4 * k + (k >> 2)] = x[k];
    named_bar_sync(bar, barn);
#pragma unroll
    for (int k = 0; k < 16; k++) x[k] = bufB[a.rC + k];
    pass16_fwd3(x, t16_fw3(s_tab, q, T), 64, p);
}
__device__ __forceinline__ void fwd16(uint32_t (&x)[16], uint32_t* bufA, uint32_t* bufB, const uint32_t* s_tab,
                                      const X16& a, int T, int q, uint32_t p, int bar, int barn = 128)
{
    pass16_fwd<4, 8>(x, c_fw1[q], 1, p);
    fwd16_tail(x, bufA, bufB, s_tab, a, T, q, p, bar, barn);
}
// the same transform of gadget digits: x[k] in [0, 256) stands for x[k] - 128 (pass16_fwd_digits)
__device__ __forceinline__ void fwd16_digits(uint32_t (&x)[16], uint32_t* bufA, uint32_t* bufB, const uint32_t* s_tab,
                                             const X16& a, int T, int q, uint32_t p, int bar, int barn = 128)
{
    pass16_fwd_digits(x, t16_lut1(s_tab, q), t16_lut24(s_tab, q), t16_lut35(s_tab, q), c_fw1[q], p);
    fwd16_tail(x, bufA, bufB, s_tab, a, T, q, p, bar, barn);
}

// two digit transforms at once (x0 through bufA, x1 through bufB): 16 independent butterflies per
// stage.  Both buffers are in use at every exchange, so each exchange's stores wait (barrier) for
// the previous exchange's loads -- four barriers, as for two single transforms.
__device__ __forceinline__ void fwd16_tail2(uint32_t (&x0)[16], uint32_t (&x1)[16], uint32_t* bufA, uint32_t* bufB,
                                            const uint32_t* s_tab, const X16& a, int T, int q, uint32_t p, int bar,
                                            bool lead, int barn)
{
    if (lead) named_bar_sync(bar, barn);        // not needed when a barrier already follows the buffers' last loads
#pragma unroll
    for (int k = 0; k < 16; k++) { bufA[a.wA + 68 * k] = x0[k]; bufB[a.wA + 68 * k] = x1[k]; }
    named_bar_sync(bar, barn);
#pragma unroll
    for (int k = 0; k < 16; k++) { x0[k] = bufA[a.rB + 4 * k]; x1[k] = bufB[a.rB + 4 * k]; }
    pass16_fwd<4, 8>(x0, t16_fw2(s_tab, q, T >> 2), 16, p);
    pass16_fwd<4, 8>(x1, t16_fw2(s_tab, q, T >> 2), 16, p);
    named_bar_sync(bar, barn);
#pragma unroll
    for (int k = 0; k < 16; k++) { bufA[a.wB + 4 * k + (k >> 2)] = x0[k]; bufB[a.wB + 4 * k + (k >> 2)] = x1[k]; }
    named_bar_sync(bar, barn);
#pragma unroll
    for (int k = 0; k < 16; k++) { x0[k] = bufA[a.rC + k]; x1[k] = bufB[a.rC + k]; }
    pass16_fwd3(x0, t16_fw3(s_tab, q, T), 64, p);
    pass16_fwd3(x1, t16_fw3(s_tab, q, T), 64, p);
}
__device__ __forceinline__ void fwd16_digits2(uint32_t (&x0)[16], uint32_t (&x1)[16], uint32_t* bufA, uint32_t* bufB,
                                              const uint32_t* s_tab, const X16& a, int T, int q, uint32_t p, int bar,
                                              bool lead = true, int barn = 128)
{
    pass16_fwd_digits(x0, t16_lut1(s_tab, q), t16_lut24(s_tab, q), t16_lut35(s_tab, q), c_fw1[q], p);
    pass16_fwd_digits(x1, t16_lut1(s_tab, q), t16_lut24(s_tab, q), t16_lut35(s_tab, q), c_fw1[q], p);
    fwd16_tail2(x0, x1, bufA, bufB, s_tab, a, T, q, p, bar, lead, barn);
}
// the same for two general inputs
__device__ __forceinline__ void fwd16_2(uint32_t (&x0)[16], uint32_t (&x1)[16], uint32_t* bufA, uint32_t* bufB,
                                        const uint32_t* s_tab, const X16& a, int T, int q, uint32_t p, int bar,
                                        bool lead = true, int barn = 128)
{
    pass16_fwd<4, 8>(x0, c_fw1[q], 1, p);
    pass16_fwd<4, 8>(x1, c_fw1[q], 1, p);
    fwd16_tail2(x0, x1, bufA, bufB, s_tab, a, T, q, p, bar, lead, barn);
}

// inverse transform incl. n^-1 psi^-e scaling: layout C in, layout A out, values in [0, 2p)
__device__ __forceinline__ void inv16(uint32_t (&x)[16], uint32_t* bufA, uint32_t* bufB, const uint32_t* s_tab,
                                      const X16& a, int T, int q, uint32_t p, int bar, int barn = 128)
{
    pass16_inv<true>(x, c_iv1[q], 1, p);
#pragma unroll
    for (int k = 0; k < 16; k++) bufA[a.rC + k] = x[k];
    named_bar_sync(bar, barn);
#pragma unroll
    for (int k = 0; k < 16; k++) x[k] = bufA[a.rD + 17 * k];
    pass16_inv<false>(x, t16_iv2(s_tab, q, T & 15), 16, p);
#pragma unroll
    for (int k = 0; k < 16; k++) bufB[a.wD + 16 * k] = x[k];
    named_bar_sync(bar, barn);
#pragma unroll
    for (int k = 0; k < 16; k++) x[k] = bufB[T + 64 * k + 16 * (k >> 2)];
    pass16_inv3(x, t16_iv3(s_tab, q, T), 64, p);
    const Tw* sc = t16_scale(s_tab, q, T);
#pragma unroll
    for (int k = 0; k < 16; k++) {
        const Tw f = sc[k * 64];
        x[k] = shoup_mul_p(x[k], f.w, f.ws, p);
    }
}

// two inverse transforms at once (x0 through bufA, x1 through bufB); the caller guarantees that
// nobody still reads either buffer when this starts
__device__ __forceinline__ void inv16_2(uint32_t (&x0)[16], uint32_t (&x1)[16], uint32_t* bufA, uint32_t* bufB,
                                        const uint32_t* s_tab, const X16& a, int T, int q, uint32_t p, int bar, int barn = 128)
{
    pass16_inv<true>(x0, c_iv1[q], 1, p);
    pass16_inv<true>(x1, c_iv1[q], 1, p);
#pragma unroll
    for (int k = 0; k < 16; k++) { bufA[a.rC + k] = x0[k]; bufB[a.rC + k] = x1[k]; }
    named_bar_sync(bar, barn);
#pragma unroll
    for (int k = 0; k < 16; k++) { x0[k] = bufA[a.rD + 17 * k]; x1[k] = bufB[a.rD + 17 * k]; }
    pass16_inv<false>(x0, t16_iv2(s_tab, q, T & 15), 16, p);
    pass16_inv<false>(x1, t16_iv2(s_tab, q, T & 15), 16, p);
    named_bar_sync(bar, barn);
#pragma unroll
    for (int k = 0; k < 16; k++) { bufA[a.wD + 16 * k] = x0[k]; bufB[a.wD + 16 * k] = x1[k]; }
    named_bar_sync(bar, barn);
#pragma unroll
    for (int k = 0; k < 16; k++) { x0[k] = bufA[T + 64 * k + 16 * (k >> 2)]; x1[k] = bufB[T + 64 * k + 16 * (k >> 2)]; }
    pass16_inv3(x0, t16_iv3(s_tab, q, T), 64, p);
    pass16_inv3(x1, t16_iv3(s_tab, q, T), 64, p);
    const Tw* sc = t16_scale(s_tab, q, T);
#pragma unroll
    for (int k = 0; k < 16; k++) {
        const Tw f = sc[k * 64];
        x0[k] = shoup_mul_p(x0[k], f.w, f.ws, p);
        x1[k] = shoup_mul_p(x1[k], f.w, f.ws, p);
    }
}

constexpr int kBr16GateSmemWords = 2 * kN + 4 * kXbufWords + 640 / 2;   // acc + 2 teams x 2 buffers + abar(u16)
__host__ __device__ constexpr int br16_smem_bytes(int gates) { return (kLane16TableWords + gates * kBr16GateSmemWords) * 4; }

// One code copy serves both primes (q is team-uniform).  sa = residues of output polynomial
// q ("mine", finished by this team), sb = residues of output polynomial 1-q (handed over).
// PAIR: 0 = one row at a time, 1 = row pairs share a Montgomery reduction, 2 = ... and are transformed together,
// 3 = ... and so are the two inverse transforms, 4 = all four rows share one reduction
template <int PAIR>
__device__ __forceinline__ void br16_gate_loop(const uint32_t* __restrict__ bk, uint32_t* acc, uint32_t* bufA,
                                               uint32_t* bufB, const uint32_t* bufA_other, const uint16_t* s_abar,
                                               const uint32_t* s_tab, const int T, const int q, const int bar)
{
    const uint32_t p = q ? Prime<1>::p : Prime<0>::p;
    const uint32_t pinv = q ? Prime<1>::pinv : Prime<0>::pinv;
    const X16 xa = make_x16(T);
    // key row layout [row][prime][out][256 uint4] in bk16_slot order; out = q first, then 1-q
    const uint4* bk4 = reinterpret_cast<const uint4*>(bk) + T + (size_t)q * 2 * (kN / 4);
    const int off_a = q * (kN / 4), off_b = (1 - q) * (kN / 4);
    uint32_t* acc_mine = acc + q * kN + T;
    for (int i = 0; i < kLwe; i++) {
        const uint32_t abar = s_abar[i];
        uint32_t sa[16], sb[16];
        if (PAIR != 4)
#pragma unroll
        for (int k = 0; k < 16; k++) sa[k] = sb[k] = PAIR ? 2u * p : 4u * p;   // + four products in (-p, p) -> (0, 8p); PAIR: + two in (-p, 2p) -> (0, 6p)

        const uint32_t am = abar & (kN - 1);
        const bool flip = (abar >> kNbit) != 0;
        const uint32_t rbase = ((uint32_t)T - abar) & (kN - 1);
        // coefficient e = T + 64k is negated iff (e < am) xor flip;  e < am <=> k < kth
        const int kth = ((int)am - T + 63) >> 6;
        if (PAIR == 4) {
            // All four digit polynomials are transformed (as two pairs) before the key multiply: the
            // four row products of an output coefficient are summed in 64 bits (< 2^34 p) and reduced
            // ONCE -- 11 multiply-pipe slots instead of 20 -- and no accumulator is live meanwhile.
            uint32_t y0[16], y1[16], x0[16], x1[16];
#pragma unroll 1
            for (int j = 0; j < 2; j++) {
                if (j == 1) {
#pragma unroll
                    for (int k = 0; k < 16; k++) { y0[k] = x0[k]; y1[k] = x1[k]; }
                }
                const uint32_t* aj = acc + j * kN;
#pragma unroll
                for (int k = 0; k < 16; k++) {
                    uint32_t rot = aj[(rbase + 64 * k) & (kN - 1)];
                    const bool neg = (k < kth) != flip;
                    rot = neg ? 0u - rot : rot;
                    const uint32_t t = rot - aj[T + 64 * k] + kDecompAdd;
                    x0[k] = t >> 24;
                    x1[k] = (t >> 16) & 0xFFu;
                }
                fwd16_digits2(x0, x1, bufA, bufB, s_tab, xa, T, q, p, bar, j != 0);
            }
            const uint4* row0 = bk4 + (size_t)(i * kRows) * (4 * (kN / 4));
            const uint4 *row1 = row0 + 4 * (kN / 4), *row2 = row0 + 8 * (kN / 4), *row3 = row0 + 12 * (kN / 4);
#pragma unroll
            for (int v = 0; v < 4; v++) {
                uint4 b0 = __ldg(row0 + off_a + v * 64), b1 = __ldg(row1 + off_a + v * 64);
                uint4 b2 = __ldg(row2 + off_a + v * 64), b3 = __ldg(row3 + off_a + v * 64);
                sa[4 * v + 0] = p + umontmul4_p(y0[4 * v + 0], b0.x, y1[4 * v + 0], b1.x, x0[4 * v + 0], b2.x, x1[4 * v + 0], b3.x, p, pinv);
                sa[4 * v + 1] = p + umontmul4_p(y0[4 * v + 1], b0.y, y1[4 * v + 1], b1.y, x0[4 * v + 1], b2.y, x1[4 * v + 1], b3.y, p, pinv);
                sa[4 * v + 2] = p + umontmul4_p(y0[4 * v + 2], b0.z, y1[4 * v + 2], b1.z, x0[4 * v + 2], b2.z, x1[4 * v + 2], b3.z, p, pinv);
                sa[4 * v + 3] = p + umontmul4_p(y0[4 * v + 3], b0.w, y1[4 * v + 3], b1.w, x0[4 * v + 3], b2.w, x1[4 * v + 3], b3.w, p, pinv);
                b0 = __ldg(row0 + off_b + v * 64), b1 = __ldg(row1 + off_b + v * 64);
                b2 = __ldg(row2 + off_b + v * 64), b3 = __ldg(row3 + off_b + v * 64);
                sb[4 * v + 0] = p + umontmul4_p(y0[4 * v + 0], b0.x, y1[4 * v + 0], b1.x, x0[4 * v + 0], b2.x, x1[4 * v + 0], b3.x, p, pinv);
                sb[4 * v + 1] = p + umontmul4_p(y0[4 * v + 1], b0.y, y1[4 * v + 1], b1.y, x0[4 * v + 1], b2.y, x1[4 * v + 1], b3.y, p, pinv);
                sb[4 * v + 2] = p + umontmul4_p(y0[4 * v + 2], b0.z, y1[4 * v + 2], b1.z, x0[4 * v + 2], b2.z, x1[4 * v + 2], b3.z, p, pinv);
                sb[4 * v + 3] = p + umontmul4_p(y0[4 * v + 3], b0.w, y1[4 * v + 3], b1.w, x0[4 * v + 3], b2.w, x1[4 * v + 3], b3.w, p, pinv);
            }
        } else if (PAIR) {
            // Both digits of acc_j are transformed before the key multiply, so a row PAIR shares one
            // Montgomery reduction: x0 b0 + x1 b1 < 2^33 p is accumulated in 64 bits (IMAD.WIDE with
            // accumulate) and reduced once -- 7 multiply-pipe slots per pair instead of 10.
#pragma unroll 1
            for (int j = 0; j < 2; j++) {
                uint32_t x0[16], x1[16];
                const uint32_t* aj = acc + j * kN;
#pragma unroll
                for (int k = 0; k < 16; k++) {
                    uint32_t rot = aj[(rbase + 64 * k) & (kN - 1)];
                    const bool neg = (k < kth) != flip;
                    rot = neg ? 0u - rot : rot;
                    const uint32_t t = rot - aj[T + 64 * k] + kDecompAdd;
                    x0[k] = t >> 24;
                    x1[k] = (t >> 16) & 0xFFu;
                }
                if (PAIR >= 2) {
                    fwd16_digits2(x0, x1, bufA, bufB, s_tab, xa, T, q, p, bar, j != 0);   // j = 0 follows the step's closing barrier
                } else {
                    fwd16_digits(x0, bufA, bufB, s_tab, xa, T, q, p, bar);
                    fwd16_digits(x1, bufA, bufB, s_tab, xa, T, q, p, bar);
                }
                const uint4* row0 = bk4 + (size_t)(i * kRows + 2 * j) * (4 * (kN / 4));
                const uint4* row1 = row0 + 4 * (kN / 4);
#pragma unroll
                for (int v = 0; v < 4; v++) {
                    const uint4 b0 = __ldg(row0 + off_a + v * 64), b1 = __ldg(row1 + off_a + v * 64);
                    sa[4 * v + 0] += umontmul2_p(x0[4 * v + 0], b0.x, x1[4 * v + 0], b1.x, p, pinv);
                    sa[4 * v + 1] += umontmul2_p(x0[4 * v + 1], b0.y, x1[4 * v + 1], b1.y, p, pinv);
                    sa[4 * v + 2] += umontmul2_p(x0[4 * v + 2], b0.z, x1[4 * v + 2], b1.z, p, pinv);
                    sa[4 * v + 3] += umontmul2_p(x0[4 * v + 3], b0.w, x1[4 * v + 3], b1.w, p, pinv);
                }
#pragma unroll
                for (int v = 0; v < 4; v++) {
                    const uint4 b0 = __ldg(row0 + off_b + v * 64), b1 = __ldg(row1 + off_b + v * 64);
                    sb[4 * v + 0] += umontmul2_p(x0[4 * v + 0], b0.x, x1[4 * v + 0], b1.x, p, pinv);
                    sb[4 * v + 1] += umontmul2_p(x0[4 * v + 1], b0.y, x1[4 * v + 1], b1.y, p, pinv);
                    sb[4 * v + 2] += umontmul2_p(x0[4 * v + 2], b0.z, x1[4 * v + 2], b1.z, p, pinv);
                    sb[4 * v + 3] += umontmul2_p(x0[4 * v + 3], b0.w, x1[4 * v + 3], b1.w, p, pinv);
                }
            }
        } else {
        uint32_t dlo[4];
#pragma unroll 1
        for (int r = 0; r < kRows; r++) {           // TRGSW row r = 2 j + d
            uint32_t x[16];
            if ((r & 1) == 0) {
                const uint32_t* aj = acc + (r >> 1) * kN;
#pragma unroll
                for (int k = 0; k < 16; k++) {
                    uint32_t rot = aj[(rbase + 64 * k) & (kN - 1)];
                    const bool neg = (k < kth) != flip;
                    rot = neg ? 0u - rot : rot;
                    const uint32_t t = rot - aj[T + 64 * k] + kDecompAdd;
                    x[k] = t >> 24;                                   // digit 0 (raw, stands for x - 128)
                    const uint32_t d1 = (t >> 16) & 0xFFu;
                    if ((k & 3) == 0) dlo[k >> 2] = d1; else dlo[k >> 2] |= d1 << (8 * (k & 3));
                }
            } else {
#pragma unroll
                for (int k = 0; k < 16; k++) x[k] = (dlo[k >> 2] >> (8 * (k & 3))) & 0xFFu;
            }
            fwd16_digits(x, bufA, bufB, s_tab, xa, T, q, p, bar);
            const uint4* row = bk4 + (size_t)(i * kRows + r) * (4 * (kN / 4));
#pragma unroll
            for (int v = 0; v < 4; v++) {
                const uint4 b = __ldg(row + off_a + v * 64);
                sa[4 * v + 0] += (uint32_t)umontmul_p(x[4 * v + 0], b.x, p, pinv);
                sa[4 * v + 1] += (uint32_t)umontmul_p(x[4 * v + 1], b.y, p, pinv);
                sa[4 * v + 2] += (uint32_t)umontmul_p(x[4 * v + 2], b.z, p, pinv);
                sa[4 * v + 3] += (uint32_t)umontmul_p(x[4 * v + 3], b.w, p, pinv);
            }
#pragma unroll
            for (int v = 0; v < 4; v++) {
                const uint4 b = __ldg(row + off_b + v * 64);
                sb[4 * v + 0] += (uint32_t)umontmul_p(x[4 * v + 0], b.x, p, pinv);
                sb[4 * v + 1] += (uint32_t)umontmul_p(x[4 * v + 1], b.y, p, pinv);
                sb[4 * v + 2] += (uint32_t)umontmul_p(x[4 * v + 2], b.z, p, pinv);
                sb[4 * v + 3] += (uint32_t)umontmul_p(x[4 * v + 3], b.w, p, pinv);
            }
        }
        }
        // two inverse transforms (two code copies: no register swap between them)
        if (PAIR >= 2) named_bar_sync(bar, 128);     // the paired transform's last loads were from both buffers
        if (PAIR >= 3) {
            inv16_2(sb, sa, bufA, bufB, s_tab, xa, T, q, p, bar);
            named_bar_sync(bar, 128);                // ... and so were this one's, and the hand-over goes through bufA
        } else {
            inv16(sb, bufA, bufB, s_tab, xa, T, q, p, bar);
            inv16(sa, bufA, bufB, s_tab, xa, T, q, p, bar);
        }
        // now sa = INTT(out q) (mine), sb = INTT(out 1-q): hand sb to the other team
#pragma unroll
        for (int k = 0; k < 16; k++) bufA[T + 64 * k] = sb[k];
        named_bar_sync(bar, 128);
#pragma unroll
        for (int k = 0; k < 16; k++) {
            const uint32_t t = bufA_other[T + 64 * k];         // other team's residue of out q
            const uint32_t s0 = q ? t : sa[k];
            const uint32_t s1 = q ? sa[k] : t;
            acc_mine[64 * k] += crt_lift32(s0, s1, c_crt_p0inv.w, c_crt_p0inv.ws);
        }
        named_bar_sync(bar, 128);
    }
}

template <int G, int PAIR = 0>
__global__ void __launch_bounds__(128 * G, 1) k_blind_rotate16(const BrParams P, const Lane16Tables* __restrict__ tab16)
{
    extern __shared__ __align__(16) uint32_t smem[];
    uint32_t* s_tab = smem;
    stage_tables16(s_tab, tab16);
    __syncthreads();

    const int gs = threadIdx.x >> 7, t128 = threadIdx.x & 127;
    const int q = t128 >> 6, T = t128 & 63;
    const int item = P.first + blockIdx.x * G + gs;
    if (item >= P.count) return;

    uint32_t* gbase = smem + kLane16TableWords + gs * kBr16GateSmemWords;
    uint32_t* acc = gbase;
    uint32_t* bufA = gbase + 2 * kN + q * 2 * kXbufWords;
    uint32_t* bufB = bufA + kXbufWords;
    const uint32_t* bufA_other = gbase + 2 * kN + (q ^ 1) * 2 * kXbufWords;
    uint16_t* s_abar = reinterpret_cast<uint16_t*>(gbase + 2 * kN + 4 * kXbufWords);
    const int bar = 1 + gs;

    // ---- pre-add + modulus switch (gatebootstrapping_gpu.cuh:316-345) ----
    BrItem it;
    if (P.items) it = P.items[item];
    else { it.g0 = it.g1 = it.slot = (uint32_t)item; it.ca = (int8_t)P.ca; it.cb = (int8_t)P.cb; it.src1 = 0; it.off = P.off; }
    const uint16_t* a0 = P.in0 + (size_t)it.g0 * kLvl0Words;
    const uint16_t* src1 = it.src1 ? P.in2 : P.in1;
    const uint16_t* a1 = (it.cb != 0 && src1) ? src1 + (size_t)it.g1 * kLvl0Words : nullptr;
    const int ca = it.ca, cb = a1 ? it.cb : 0;
    for (int i = t128; i < kLwe; i += 128) {
        const int c = ca * (int)a0[i] + (a1 ? cb * (int)a1[i] : 0);
        s_abar[i] = (uint16_t)modswitch_a((uint32_t)c & 0xFFFFu);
    }
    const int cbody = it.off + ca * (int)a0[kLwe] + (a1 ? cb * (int)a1[kLwe] : 0);
    const uint32_t bbar = modswitch_b((uint32_t)cbody & 0xFFFFu);
    for (int c = t128; c < kN; c += 128) {
        acc[c] = 0u;
        acc[kN + c] = testvec_b((uint32_t)c, bbar, P.mu);
    }
    named_bar_sync(bar, 128);

    br16_gate_loop<PAIR>(P.bk, acc, bufA, bufB, bufA_other, s_abar, s_tab, T, q, bar);

    // ---- SampleExtract(0) (src/bootstrap_gpu.cu:366-381) ----
    if (P.trlwe) {          // BootstrapTLWE2TRLWE (src/bootstrap_gpu.cu:316-322): no extraction
        uint32_t* out = P.out + (size_t)it.slot * 2 * kN;
        for (int m = t128; m < 2 * kN; m += 128) out[m] = acc[m];
        return;
    }
    uint32_t* out = P.out + (size_t)it.slot * kLvl1Words;
    for (int m = t128; m < kN; m += 128) out[m] = m == 0 ? acc[0] : 0u - acc[kN - m];
    if (t128 == 0) out[kN] = acc[kN];
}

// key polynomial storage order of the 32-coefficient kernel (bk_slot) -> this kernel's
// (bk16_slot); both index the same bit-reversed NTT positions.
__global__ void k_bk_to_layout16(const uint32_t* __restrict__ src, uint32_t* __restrict__ dst, size_t npoly)
{
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < npoly * kN; idx += (size_t)gridDim.x * blockDim.x) {
        const size_t poly = idx >> kNbit;
        const int pos = (int)(idx & (kN - 1));
        dst[poly * kN + bk16_slot(pos >> 4, pos & 15)] = src[poly * kN + bk_slot(pos >> 5, pos & 31)];
    }
}

}  // namespace b2
