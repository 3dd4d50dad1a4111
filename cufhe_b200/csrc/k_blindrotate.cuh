// k_blindrotate.cuh -- batched blind rotation + SampleExtract(0).
//
// Replaces __BlindRotatePreAdd__/__BlindRotate__/Accumulate
// (reference include/gatebootstrapping_gpu.cuh:115-345) and __SampleExtractIndex__
// (src/bootstrap_gpu.cu:366-381), which run ONE gate per 512-thread block and re-read
// the whole 41.7 MB key per gate.
//
// B200 design: a CTA owns G gates and walks the 636 CMUX steps once, so every key row
// is fetched once per CTA (L1/L2 resident) and shared by its gates.  A gate is two
// warps, one per NTT prime.  Per step each warp
//   1. rebuilds the four gadget-digit polynomials from the accumulator (shared memory),
//   2. forward-transforms each (registers + one warp transpose),
//   3. multiply-accumulates with the key row into 2 x 32 lazy residues (registers;
//      Montgomery form, the key carries the 2^32 factor),
//   4. inverse-transforms both outputs,
// then the two warps swap one residue polynomial through shared memory, each finishes
// the centred CRT for one output polynomial and updates the accumulator.  The only
// synchronisation is two 64-thread named barriers per step.
#pragma once
#include "dev_common.cuh"

namespace b2 {

// one blind-rotation work item: c = ca*in0[g0] + cb*in1[g1] + (0, off) over u16, result
// (a lvl1 TLWE, 1025 words) written to out[slot].
struct BrItem {
    uint32_t g0, g1;        // ciphertext indices into src0 / src1
    int8_t ca, cb;          // linear-combination signs
    int8_t src1;            // 0: second operand from in1, 1: from in2
    int8_t pad;
    int32_t off;            // offset added to the body term (u16 arithmetic)
    uint32_t slot;          // output slot
};

struct BrParams {
    const uint16_t* in0;    // [*][637] lvl0 ciphertexts
    const uint16_t* in1;    // may be null when cb == 0
    const uint16_t* in2;    // third operand array (Mux), may be null
    uint32_t* out;          // [*][1025] lvl1 TLWEs
    const BrItem* items;    // null: uniform batch (item i = gate i with ca/cb/off below)
    const uint32_t* bk;     // NTT-domain key [636][4 rows][2 primes][2 outs][1024] (bk_slot order)
    const LaneTables* tab;
    int count;
    int ca, cb, off;        // uniform coefficients
    uint32_t mu;            // test-vector constant (lvl1param::mu)
    int trlwe;              // 0: out[slot] = SampleExtract(0) (1025 words); 1: out[slot] = the TRLWE (2 x 1024 words)
    int first;              // k_blind_rotate16 / _s only: this launch covers items first .. count-1
    uint32_t dbg;           // always 0 in the product; tools/ablate_brs.cu uses it as an opaque zero
};

constexpr int kBrGateSmemWords = 2 * kN + 2 * kN + 640 / 2;   // acc + 2 exchange buffers + abar(u16)
__host__ __device__ constexpr int br_smem_bytes(int gates) { return (kLaneTableWords + gates * kBrGateSmemWords) * 4; }

// One code copy serves both primes (q is warp-uniform): the unrolled transforms are large
// and two specialised copies thrash the instruction cache (ncu: stall_no_instruction).
// Register arrays: sa = residues of output polynomial q ("mine", finished by this warp),
// sb = residues of output polynomial 1-q (handed to the partner warp).
__device__ __forceinline__ void br_gate_loop(const uint32_t* __restrict__ bk, uint32_t* acc, uint32_t* xbuf,
                                             const uint32_t* xbuf_other, const uint16_t* s_abar,
                                             const uint32_t* s_tab, const int lane, const int barid, const int q)
{
    const uint32_t p = q ? Prime<1>::p : Prime<0>::p;
    const uint32_t pinv = q ? Prime<1>::pinv : Prime<0>::pinv;
    const uint32_t dig_off = p - 128u;
    // key row layout [row][prime][out][256 uint4]; this warp reads out = q first, then 1-q
    const uint4* bk4 = reinterpret_cast<const uint4*>(bk) + lane + (size_t)q * 2 * (kN / 4);
    const int off_a = q * (kN / 4), off_b = (1 - q) * (kN / 4);
    uint32_t* acc_mine = acc + q * kN + lane;
    for (int i = 0; i < kLwe; i++) {
        const uint32_t abar = s_abar[i];
        // lazy residues: 4p + four Montgomery products in (-p, p)  ->  (0, 8p)
        uint32_t sa[32], sb[32];
#pragma unroll
        for (int k = 0; k < 32; k++) sa[k] = sb[k] = 4u * p;

        const uint32_t am = abar & (kN - 1);
        const bool flip = (abar >> kNbit) != 0;
        const uint32_t rbase = (lane - abar) & (kN - 1);
        // coefficient c = lane + 32k is negated iff (c < am) xor flip, c < am <=> k < kth
        const int kth = ((int)am - lane + 31) >> 5;
        uint32_t dlo[8];
#pragma unroll 1
        for (int r = 0; r < kRows; r++) {           // TRGSW row r = 2 j + d
            uint32_t x[32];
            if ((r & 1) == 0) {
                const uint32_t* aj = acc + (r >> 1) * kN;
#pragma unroll
                for (int k = 0; k < 32; k++) {
                    uint32_t rot = aj[(rbase + 32 * k) & (kN - 1)];
                    const bool neg = (k < kth) != flip;
                    rot = neg ? 0u - rot : rot;
                    const uint32_t t = rot - aj[lane + 32 * k] + kDecompAdd;
                    x[k] = (t >> 24) + dig_off;                       // digit 0 + p in [0, 2p)
                    const uint32_t d1 = (t >> 16) & 0xFFu;
                    if ((k & 3) == 0) dlo[k >> 2] = d1; else dlo[k >> 2] |= d1 << (8 * (k & 3));
                }
            } else {
#pragma unroll
                for (int k = 0; k < 32; k++) x[k] = ((dlo[k >> 2] >> (8 * (k & 3))) & 0xFFu) + dig_off;
            }
            fwd_dev_rt(x, xbuf, s_tab, lane, q, p);
            const uint4* row = bk4 + (size_t)(i * kRows + r) * (4 * (kN / 4));
#pragma unroll
            for (int v = 0; v < 8; v++) {
                const uint4 b = __ldg(row + off_a + v * 32);
                sa[4 * v + 0] += (uint32_t)umontmul_p(x[4 * v + 0], b.x, p, pinv);
                sa[4 * v + 1] += (uint32_t)umontmul_p(x[4 * v + 1], b.y, p, pinv);
                sa[4 * v + 2] += (uint32_t)umontmul_p(x[4 * v + 2], b.z, p, pinv);
                sa[4 * v + 3] += (uint32_t)umontmul_p(x[4 * v + 3], b.w, p, pinv);
            }
#pragma unroll
            for (int v = 0; v < 8; v++) {
                const uint4 b = __ldg(row + off_b + v * 32);
                sb[4 * v + 0] += (uint32_t)umontmul_p(x[4 * v + 0], b.x, p, pinv);
                sb[4 * v + 1] += (uint32_t)umontmul_p(x[4 * v + 1], b.y, p, pinv);
                sb[4 * v + 2] += (uint32_t)umontmul_p(x[4 * v + 2], b.z, p, pinv);
                sb[4 * v + 3] += (uint32_t)umontmul_p(x[4 * v + 3], b.w, p, pinv);
            }
        }
        // two inverse transforms through ONE code copy: transform sb, swap, transform (old) sa
#pragma unroll 1
        for (int t = 0; t < 2; t++) {
            inv_dev_rt(sb, xbuf, s_tab, lane, q, p);
#pragma unroll
            for (int k = 0; k < 32; k++) { const uint32_t tmp = sa[k]; sa[k] = sb[k]; sb[k] = tmp; }
        }
        // now sa = INTT(out q) (mine), sb = INTT(out 1-q): hand sb to the partner
        __syncwarp();
#pragma unroll
        for (int k = 0; k < 32; k++) xbuf[k * 32 + lane] = sb[k];
        named_bar_sync(barid, 64);
#pragma unroll
        for (int k = 0; k < 32; k++) {
            const uint32_t t = xbuf_other[k * 32 + lane];      // partner's residue of out q
            const uint32_t s0 = q ? t : sa[k];
            const uint32_t s1 = q ? sa[k] : t;
            acc_mine[32 * k] += crt_lift32(s0, s1, c_crt_p0inv.w, c_crt_p0inv.ws);
        }
        named_bar_sync(barid, 64);
    }
}

template <int G>
__global__ void __launch_bounds__(64 * G, 1) k_blind_rotate(const BrParams P)
{
    extern __shared__ __align__(16) uint32_t smem[];
    uint32_t* s_tab = smem;
    stage_lane_tables(s_tab, P.tab);
    __syncthreads();

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    constexpr int gates = G;
    const int gs = warp >> 1, q = warp & 1;
    const int item = blockIdx.x * gates + gs;
    if (item >= P.count) return;

    uint32_t* gbase = smem + kLaneTableWords + gs * kBrGateSmemWords;
    uint32_t* acc = gbase;
    uint32_t* xbuf = gbase + 2 * kN + q * kN;
    const uint32_t* xbuf_other = gbase + 2 * kN + (q ^ 1) * kN;
    uint16_t* s_abar = reinterpret_cast<uint16_t*>(gbase + 4 * kN);
    const int barid = 1 + gs;
    const int t64 = q * 32 + lane;

    // ---- pre-add + modulus switch (gatebootstrapping_gpu.cuh:316-345) ----
    BrItem it;
    if (P.items) it = P.items[item];
    else { it.g0 = it.g1 = it.slot = (uint32_t)item; it.ca = (int8_t)P.ca; it.cb = (int8_t)P.cb; it.src1 = 0; it.off = P.off; }
    const uint16_t* a0 = P.in0 + (size_t)it.g0 * kLvl0Words;
    const uint16_t* src1 = it.src1 ? P.in2 : P.in1;
    const uint16_t* a1 = (it.cb != 0 && src1) ? src1 + (size_t)it.g1 * kLvl0Words : nullptr;
    const int ca = it.ca, cb = a1 ? it.cb : 0;
    for (int i = t64; i < kLwe; i += 64) {
        const int c = ca * (int)a0[i] + (a1 ? cb * (int)a1[i] : 0);
        s_abar[i] = (uint16_t)modswitch_a((uint32_t)c & 0xFFFFu);
    }
    const int cbody = it.off + ca * (int)a0[kLwe] + (a1 ? cb * (int)a1[kLwe] : 0);
    const uint32_t bbar = modswitch_b((uint32_t)cbody & 0xFFFFu);
    for (int c = t64; c < kN; c += 64) {
        acc[c] = 0u;
        acc[kN + c] = testvec_b((uint32_t)c, bbar, P.mu);
    }
    named_bar_sync(barid, 64);

    br_gate_loop(P.bk, acc, xbuf, xbuf_other, s_abar, s_tab, lane, barid, q);

    // ---- SampleExtract(0) (src/bootstrap_gpu.cu:366-381) ----
    if (P.trlwe) {          // BootstrapTLWE2TRLWE (src/bootstrap_gpu.cu:316-322): no extraction
        uint32_t* out = P.out + (size_t)it.slot * 2 * kN;
        for (int m = t64; m < 2 * kN; m += 64) out[m] = acc[m];
        return;
    }
    uint32_t* out = P.out + (size_t)it.slot * kLvl1Words;
    for (int m = t64; m < kN; m += 64) out[m] = m == 0 ? acc[0] : 0u - acc[kN - m];
    if (t64 == 0) out[kN] = acc[kN];
}

}  // namespace b2
