// k_blindrotate_lat.cuh -- LOW-LATENCY blind rotation: one gate per CTA, 512 threads.
//
// For small batches (at most a couple of gates per SM) the throughput kernels leave most
// of the chip idle and a gate takes 636 x (4 forward + 2 inverse transforms) in sequence.
// Here the eight 64-thread teams of a CTA are (prime q, TRGSW row r): each team builds ONE
// digit polynomial, runs ONE forward transform and multiplies it with its key row; the
// four row products per (prime, output) are summed through shared memory by the team that
// then runs that output's inverse transform ((q, r) with r = output, r < 2), and teams
// (0, o) finish the centred CRT.  Critical path per CMUX step: 1 forward + 1 inverse
// transform instead of 4 + 2.  Same arithmetic (ntt16.cuh), same key layout and tables as
// k_blind_rotate16, bit-identical results.
//
// Replaces the reference's one-gate-per-block kernels for the single-gate latency case
// (__BlindRotatePreAdd__ + Accumulate, include/gatebootstrapping_gpu.cuh:115-345;
// SampleExtract src/bootstrap_gpu.cu:366-381).
#pragma once
#include "team16_dev.cuh"

namespace b2 {

constexpr int kLatThreads = 512;
// tables | acc | 8 teams x 2 exchange buffers | 8 teams x 2 outputs x 1024 partial products | abar
constexpr int kLatSmemWords = kLane16TableWords + 2 * kN + 16 * kXbufWords + 16 * kN + 640 / 2;
constexpr int kLatSmemBytes = kLatSmemWords * 4;

__global__ void __launch_bounds__(kLatThreads, 1) k_blind_rotate_lat(const BrParams P, const Lane16Tables* __restrict__ tab16)
{
    extern __shared__ __align__(16) uint32_t smem[];
    uint32_t* s_tab = smem;
    uint32_t* acc = smem + kLane16TableWords;
    uint32_t* bufs = acc + 2 * kN;
    uint32_t* part = bufs + 16 * kXbufWords;           // [team][out][k][T]
    uint16_t* s_abar = reinterpret_cast<uint16_t*>(part + 16 * kN);
    stage_tables16(s_tab, tab16);

    const int tid = threadIdx.x;
    const int team = tid >> 6, T = tid & 63;
    const int q = team >> 2, r = team & 3;              // prime, TRGSW row (r = 2 j + d)
    const int item = blockIdx.x;
    uint32_t* bufA = bufs + team * 2 * kXbufWords;
    uint32_t* bufB = bufA + kXbufWords;
    const int bar_team = 1 + team;                      // 64 threads
    const int bar_out = 9 + r;                          // teams (0, o) and (1, o): 128 threads (r < 2 only)

    // ---- pre-add + modulus switch (gatebootstrapping_gpu.cuh:316-345) ----
    BrItem it;
    if (P.items) it = P.items[item];
    else { it.g0 = it.g1 = it.slot = (uint32_t)item; it.ca = (int8_t)P.ca; it.cb = (int8_t)P.cb; it.src1 = 0; it.off = P.off; }
    const uint16_t* a0 = P.in0 + (size_t)it.g0 * kLvl0Words;
    const uint16_t* src1 = it.src1 ? P.in2 : P.in1;
    const uint16_t* a1 = (it.cb != 0 && src1) ? src1 + (size_t)it.g1 * kLvl0Words : nullptr;
    const int ca = it.ca, cb = a1 ? it.cb : 0;
    for (int i = tid; i < kLwe; i += kLatThreads) {
        const int c = ca * (int)a0[i] + (a1 ? cb * (int)a1[i] : 0);
        s_abar[i] = (uint16_t)modswitch_a((uint32_t)c & 0xFFFFu);
    }
    const int cbody = it.off + ca * (int)a0[kLwe] + (a1 ? cb * (int)a1[kLwe] : 0);
    const uint32_t bbar = modswitch_b((uint32_t)cbody & 0xFFFFu);
    for (int c = tid; c < kN; c += kLatThreads) {
        acc[c] = 0u;
        acc[kN + c] = testvec_b((uint32_t)c, bbar, P.mu);
    }
    __syncthreads();

    const uint32_t p = q ? Prime<1>::p : Prime<0>::p;
    const uint32_t pinv = q ? Prime<1>::pinv : Prime<0>::pinv;
    const uint32_t dig_off = p - 128u;
    const X16 xa = make_x16(T);
    // key row layout [i][row][prime][out][256 uint4] in bk16_slot order
    const uint4* bk4 = reinterpret_cast<const uint4*>(P.bk) + T + (size_t)(r * 2 + q) * 2 * (kN / 4);
    const uint32_t* aj = acc + (r >> 1) * kN;
    const int dshift = (r & 1) ? 16 : 24;
    uint32_t* my_part = part + (size_t)team * 2 * kN + T;                 // + o * kN + k * 64
    const uint32_t* sum_part = part + (size_t)(q * 4) * 2 * kN + (size_t)r * kN + T;   // out o = r: + team' * 2 kN + k * 64
    const uint32_t* crt_other = bufs + (size_t)(4 + r) * 2 * kXbufWords;  // bufA of team (1, r)

    for (int i = 0; i < kLwe; i++) {
        const uint32_t abar = s_abar[i];
        const uint32_t am = abar & (kN - 1);
        const bool flip = (abar >> kNbit) != 0;
        const uint32_t rbase = ((uint32_t)T - abar) & (kN - 1);
        const int kth = ((int)am - T + 63) >> 6;
        uint32_t x[16];
#pragma unroll
        for (int k = 0; k < 16; k++) {
            uint32_t rot = aj[(rbase + 64 * k) & (kN - 1)];
            const bool neg = (k < kth) != flip;
            rot = neg ? 0u - rot : rot;
            const uint32_t t = rot - aj[T + 64 * k] + kDecompAdd;
            x[k] = ((t >> dshift) & 0xFFu) + dig_off;                     // digit + p in [0, 2p)
        }
        // key row fetched BEFORE the transform (its barriers are compiler memory fences, so the
        // loads would otherwise be issued only after it and their L2 latency would be exposed)
        const uint4* row = bk4 + (size_t)i * kRows * (4 * (kN / 4));
        uint4 kb[8];
#pragma unroll
        for (int o = 0; o < 2; o++)
#pragma unroll
            for (int v = 0; v < 4; v++) kb[4 * o + v] = __ldg(row + o * (kN / 4) + v * 64);
        fwd16(x, bufA, bufB, s_tab, xa, T, q, p, bar_team, 64);
#pragma unroll
        for (int o = 0; o < 2; o++)
#pragma unroll
            for (int v = 0; v < 4; v++) {
                const uint4 b = kb[4 * o + v];
                uint32_t* dst = my_part + o * kN + (4 * v) * 64;
                dst[0] = (uint32_t)umontmul_p(x[4 * v + 0], b.x, p, pinv) + p;   // (0, 2p)
                dst[64] = (uint32_t)umontmul_p(x[4 * v + 1], b.y, p, pinv) + p;
                dst[128] = (uint32_t)umontmul_p(x[4 * v + 2], b.z, p, pinv) + p;
                dst[192] = (uint32_t)umontmul_p(x[4 * v + 3], b.w, p, pinv) + p;
            }
        __syncthreads();
        if (r < 2) {
            // team (q, o = r): sum the four row products of output o, inverse transform
            uint32_t y[16];
#pragma unroll
            for (int k = 0; k < 16; k++) {
                const uint32_t* s = sum_part + k * 64;
                y[k] = s[0] + s[2 * kN] + s[4 * kN] + s[6 * kN];                   // (0, 8p)
            }
            inv16(y, bufA, bufB, s_tab, xa, T, q, p, bar_team, 64);
            if (q == 1) {
#pragma unroll
                for (int k = 0; k < 16; k++) bufA[T + 64 * k] = y[k];
            }
            named_bar_sync(bar_out, 128);
            if (q == 0) {
                uint32_t* ao = acc + r * kN + T;
#pragma unroll
                for (int k = 0; k < 16; k++)
                    ao[64 * k] += crt_lift32(y[k], crt_other[T + 64 * k], c_crt_p0inv.w, c_crt_p0inv.ws);
            }
        }
        __syncthreads();
    }

    // ---- SampleExtract(0) (src/bootstrap_gpu.cu:366-381) ----
    if (P.trlwe) {          // BootstrapTLWE2TRLWE (src/bootstrap_gpu.cu:316-322): no extraction
        uint32_t* out = P.out + (size_t)it.slot * 2 * kN;
        for (int m = tid; m < 2 * kN; m += kLatThreads) out[m] = acc[m];
        return;
    }
    uint32_t* out = P.out + (size_t)it.slot * kLvl1Words;
    for (int m = tid; m < kN; m += kLatThreads) out[m] = m == 0 ? acc[0] : 0u - acc[kN - m];
    if (tid == 0) out[kN] = acc[kN];
}

}  // namespace b2
