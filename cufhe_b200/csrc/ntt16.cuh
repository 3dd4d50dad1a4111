// ntt16.cuh -- 1024-point negacyclic NTT for a TEAM of 64 threads x 16 coefficients.
//
// Same transform, primes, twiddles and bit-reversed output order as ntt_core.cuh (one
// warp x 32 coefficients); only the distribution differs.  Halving the per-thread state
// (16 + 2x16 accumulator words instead of 32 + 2x32) lets a CTA keep 24 warps resident
// instead of 12, which is what hides the IMAD.HI / LDS / LDG latencies that ncu showed
// the 32-coefficient kernel stalls on (profiles/r1_ncu_full_k_blind_rotate_v3_*).
//
// Replaces SmallForwardNTT_1024 / SmallInverseNTT_1024 (reference
// include/ntt_gpu/ntt_gpuntt.cuh:232-276,342-392).
//
// Thread T (0..63) of a team, register k (0..15); e = coefficient / position index:
//   layout A: e = T + 64 k                       (stride 64)
//   layout B: e = 64 (T>>2) + (T&3) + 4 k        (one 64-block per 4 threads, stride 4)
//   layout C: e = 16 T + k                       (contiguous)
//   layout D: e = 256 (T>>4) + (T&15) + 16 k     (stride 16)
//
// forward (merged-psi Cooley-Tukey, natural -> bit-reversed), butterfly strides
//   F1: 512..64 in layout A (15 warp-uniform twiddles, constant bank)
//   F2: 32..4   in layout B (15 twiddles per 64-block)
//   F3: 2, 1    in layout C (12 twiddles per thread)
// inverse (cyclic DIT with w^-1, bit-reversed -> natural, then * n^-1 psi^-e)
//   I1: 1..8    in layout C (15 warp-uniform twiddles; those equal to 1 skip the multiply)
//   I2: 16..128 in layout D (15 twiddles per (T&15))
//   I3: 256,512 in layout A (12 twiddles per thread) + 16 per-thread scale factors
// Between passes the team exchanges through PADDED shared-memory buffers; every
// store/load below is bank-conflict free with compile-time offsets:
//   A->B: slot e + 4 (e>>6)      B->C and C->D: slot e + (e>>4)      D->A: slot e + 16 (e>>8)
// Lazy-residue bounds are those of ntt_core.cuh (each stage adds 2p, p < 2^27).
#pragma once
#include "ntt_core.cuh"

namespace b2 {

constexpr int kXbufWords = 1088;                 // padded exchange buffer (max slot 1087)

struct alignas(8) U2 { uint32_t x, y; };
struct Lane16Tables {
    Tw fw2[2][15][16];      // [q][(1<<s')-1+g][block]
    Tw fw3[2][12][64];      // [q][slot][T]   slots 0..3: stride 2, 4..11: stride 1
    Tw iv2[2][15][16];      // [q][(1<<u')-1+g][T&15]
    Tw iv3[2][12][64];      // [q][slot][T]   slots 0..3: t=256, 4..11: t=512
    Tw scale[2][16][64];    // [q][k][T]      n^-1 psi^-(T + 64 k)
    // gadget-digit look-up tables for forward stages 0 and 1 (pass16_fwd_digits): entry d stands
    // for the signed digit d - 128; W1, W2a, W2b = twiddles of stage 0 and of the two stage-1 groups
    uint32_t lut1[2][256];  // W1 (d-128) mod p
    U2 lut24[2][256];       // { W2a (d-128), W2b (d-128) } mod p
    U2 lut35[2][256];       // { W2a W1 (d-128), W2b W1 (d-128) } mod p
};
constexpr int kLane16TableWords = sizeof(Lane16Tables) / 4;      // 14720 words = 58880 B
constexpr int kLut1Word = (sizeof(Tw) * (2 * 15 * 16 + 2 * 12 * 64 + 2 * 15 * 16 + 2 * 12 * 64 + 2 * 16 * 64)) / 4;
static_assert(kLut1Word == 12160 && kLane16TableWords == 12160 + 2 * 256 * 5, "table layout");

// F1 / I1 reuse the first 15 entries of NttTables::fw1 / iv1 (same formulas).
inline void make_tables16(Lane16Tables& L)
{
    for (int q = 0; q < 2; q++) {
        const uint32_t p = kP[q];
        const uint32_t psi = h_psi(p);
        const uint32_t ipsi = h_powmod(psi, p - 2, p);
        const uint32_t iw = h_mulmod(ipsi, ipsi, p);
        const uint32_t ninv = h_powmod(kN, p - 2, p);
        auto M = [&](uint32_t v) { return Tw{v, h_shoup(v, p)}; };
        auto psirev = [&](unsigned idx) { return h_powmod(psi, brv10(idx), p); };
        // forward stage s (m = 2^s groups): group i of element e is e >> (10 - s), twiddle psi^brv(m + i)
        for (int sp = 0; sp < 4; sp++)                      // stage 4 + sp, i = (B << sp) + g
            for (int g = 0; g < (1 << sp); g++)
                for (int B = 0; B < 16; B++)
                    L.fw2[q][(1 << sp) - 1 + g][B] = M(psirev((16u << sp) + ((unsigned)B << sp) + g));
        for (int T = 0; T < 64; T++) {
            for (int g = 0; g < 4; g++) L.fw3[q][g][T] = M(psirev(256u + 4u * T + g));          // stage 8
            for (int g = 0; g < 8; g++) L.fw3[q][4 + g][T] = M(psirev(512u + 8u * T + g));      // stage 9
        }
        // inverse DIT: half-size t uses w^-(j * 512 / t), j = e mod t
        for (int up = 0; up < 4; up++)                      // t = 16 << up, j = lo + 16 g
            for (int g = 0; g < (1 << up); g++)
                for (int lo = 0; lo < 16; lo++)
                    L.iv2[q][(1 << up) - 1 + g][lo] = M(h_powmod(iw, (uint64_t)(lo + 16 * g) * (32 >> up), p));
        for (int T = 0; T < 64; T++) {
            for (int g = 0; g < 4; g++) L.iv3[q][g][T] = M(h_powmod(iw, (uint64_t)(T + 64 * g) * 2, p));   // t = 256
            for (int g = 0; g < 8; g++) L.iv3[q][4 + g][T] = M(h_powmod(iw, (uint64_t)(T + 64 * g), p));   // t = 512
        }
        for (int k = 0; k < 16; k++)
            for (int T = 0; T < 64; T++)
                L.scale[q][k][T] = M(h_mulmod(ninv, h_powmod(ipsi, T + 64 * k, p), p));
        const uint32_t w1 = psirev(1), w2a = psirev(2), w2b = psirev(3);
        for (int d = 0; d < 256; d++) {
            const uint32_t v = d >= 128 ? (uint32_t)(d - 128) : p - (uint32_t)(128 - d);     // d - 128 mod p
            L.lut1[q][d] = h_mulmod(w1, v, p);
            L.lut24[q][d] = U2{h_mulmod(w2a, v, p), h_mulmod(w2b, v, p)};
            L.lut35[q][d] = U2{h_mulmod(h_mulmod(w2a, w1, p), v, p), h_mulmod(h_mulmod(w2b, w1, p), v, p)};
        }
    }
}

// ---- padded exchange slots ---------------------------------------------------------
B2_HD int padAB(int e) { return e + 4 * (e >> 6); }
B2_HD int padC(int e) { return e + (e >> 4); }
B2_HD int padDA(int e) { return e + 16 * (e >> 8); }
B2_HD int eA(int T, int k) { return T + 64 * k; }
B2_HD int eB(int T, int k) { return 64 * (T >> 2) + (T & 3) + 4 * k; }
B2_HD int eC(int T, int k) { return 16 * T + k; }
B2_HD int eD(int T, int k) { return 256 * (T >> 4) + (T & 15) + 16 * k; }

// storage slot of NTT position (T, k) = 16 T + k inside one 1024-word key polynomial:
// [v = k/4][T][e = k%4], so a warp's uint4 load is 512 contiguous bytes.
B2_HD int bk16_slot(int T, int k) { return ((k >> 2) * 64 + T) * 4 + (k & 3); }

// ---- in-register passes on 16 values ---------------------------------------------------
// `nst` stages with sub-strides first, first/2, ...; twiddle index (2^s - 1) + k / (2 tp)
template <int NST, int FIRST>
B2_HD void pass16_fwd(uint32_t (&x)[16], const Tw* tw, const int stride, const uint32_t p)
{
    const uint32_t p2 = 2u * p;
#pragma unroll
    for (int s = 0; s < NST; s++) {
        const int tp = FIRST >> s;
#pragma unroll
        for (int k = 0; k < 16; k++)
            if ((k & tp) == 0) butterfly_p(x[k], x[k + tp], tw[((1 << s) - 1 + k / (2 * tp)) * stride], p, p2);
    }
}
// Pass F1 on raw gadget digits x[k] in [0, 256) (meaning x[k] - 128): stages 0 and 1 are linear
// in the four digits of a radix-4 group, so each term is ONE table entry -- five look-ups and
// eight additions replace four Shoup butterflies; stages 2 and 3 follow as usual.  Outputs are
// congruent to pass16_fwd<4, 8> on (digit - 128 + p) and obey its bounds (< 4p + 128 after the
// table stages, where the butterflies allow 6p).
B2_HD void pass16_fwd_digits(uint32_t (&x)[16], const uint32_t* l1, const U2* l24, const U2* l35, const Tw* tw,
                             const uint32_t p)
{
    const uint32_t p2 = 2u * p;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const uint32_t d0 = x[k], t1 = l1[x[k + 8]];
        const U2 u = l24[x[k + 4]], v = l35[x[k + 12]];
        const uint32_t A = d0 + (p - 128u) + t1;          // x0 + W1 x2
        const uint32_t C = d0 + (p2 - 128u) - t1;         // x0 - W1 x2
        const uint32_t B = u.x + v.x;                     // W2a (x1 + W1 x3)
        const uint32_t D = u.y - v.y + p;                 // W2b (x1 - W1 x3)
        x[k] = A + B;
        x[k + 4] = A - B + p2;
        x[k + 8] = C + D;
        x[k + 12] = C - D + p2;
    }
#pragma unroll
    for (int s = 2; s < 4; s++) {
        const int tp = 8 >> s;
#pragma unroll
        for (int k = 0; k < 16; k++)
            if ((k & tp) == 0) butterfly_p(x[k], x[k + tp], tw[(1 << s) - 1 + k / (2 * tp)], p, p2);
    }
}
// F3: sub-strides 2 then 1; slots 0..3 then 4..11
B2_HD void pass16_fwd3(uint32_t (&x)[16], const Tw* tw, const int stride, const uint32_t p)
{
    const uint32_t p2 = 2u * p;
#pragma unroll
    for (int k = 0; k < 16; k++)
        if ((k & 2) == 0) butterfly_p(x[k], x[k + 2], tw[(k >> 2) * stride], p, p2);
#pragma unroll
    for (int k = 0; k < 16; k += 2) butterfly_p(x[k], x[k + 1], tw[(4 + (k >> 1)) * stride], p, p2);
}

// reduction of any 32-bit lazy residue into [0, 2p): v - (v >> 27) p  (needs 31 (2^27 - p) < 2p - 2^27; both primes are 2^27 - 2^k + 1, k <= 21)
B2_HD uint32_t lazy_reduce(uint32_t v, uint32_t p) { return v - (v >> 27) * p; }

// inverse sub-strides 1, 2, 4, 8; twiddle index (2^u - 1) + (k mod tp).  UNIT: twiddle
// 0 of every stage is 1 (pass I1): replace the Shoup product by a plain reduction.
template <bool UNIT>
B2_HD void pass16_inv(uint32_t (&x)[16], const Tw* tw, const int stride, const uint32_t p)
{
    const uint32_t p2 = 2u * p;
#pragma unroll
    for (int u = 0; u < 4; u++) {
        const int tp = 1 << u;
#pragma unroll
        for (int k = 0; k < 16; k++)
            if ((k & tp) == 0) {
                if (UNIT && (k & (tp - 1)) == 0) {
                    const uint32_t r = lazy_reduce(x[k + tp], p);
                    const uint32_t a = x[k];
                    x[k] = a + r;
                    x[k + tp] = a - r + p2;
                } else {
                    butterfly_p(x[k], x[k + tp], tw[((1 << u) - 1 + (k & (tp - 1))) * stride], p, p2);
                }
            }
    }
}
// I3: sub-strides 4 (t = 256) then 8 (t = 512); slots 0..3 then 4..11
B2_HD void pass16_inv3(uint32_t (&x)[16], const Tw* tw, const int stride, const uint32_t p)
{
    const uint32_t p2 = 2u * p;
#pragma unroll
    for (int k = 0; k < 16; k++)
        if ((k & 4) == 0) butterfly_p(x[k], x[k + 4], tw[(k & 3) * stride], p, p2);
#pragma unroll
    for (int k = 0; k < 8; k++) butterfly_p(x[k], x[k + 8], tw[(4 + k) * stride], p, p2);
}

}  // namespace b2
