// ntt16s.cuh -- 1024-point negacyclic NTT for a TEAM of 64 threads x 16
// coefficients: SIGNED lazy residues, 128-bit exchanges, just-in-time twiddle vectors.
//
// The transform: merged-psi Cooley-Tukey forward, natural -> bit-reversed (stage s, 2^s groups: group i
// of element e is e >> (10 - s), twiddle psi^brv(2^s + i)); cyclic DIT inverse with w^-1 (half-size t
// uses w^-(j 512 / t), j = e mod t) followed by n^-1 psi^-e.  NTT-domain positions are stored in
// bk16_slot order (ntt_core.cuh), Montgomery form in [0, p): the layout of the bootstrapping key.
// Replaces SmallForwardNTT_1024 / SmallInverseNTT_1024 (reference
// include/ntt_gpu/ntt_gpuntt.cuh:232-276,342-392).
//
// Cost model this file is written against (measured on B200, profiles/r2_ubench_mix_issue_port.log):
// IMAD.HI / IMAD.WIDE take 4 cycles of a sub-partition during which NO ALU instruction overlaps
// (IMAD.HI + k IADD3 = 4 + 2k cycles), IMAD (2 cycles) overlaps ALU work perfectly.  So a step costs
// about 4 N(HI, WIDE) + N(everything else) issue cycles per warp, and the job is fewer instructions:
//
//  * residues are int32 in (-2^31, 2^31).  Shoup multiplication with a CENTRED twiddle w and the
//    ROUNDED companion ws = round(w 2^32 / p):  q = mulhi.s32(v, ws), r = v w - q p lies in
//    [-p/4, 5p/4) for ANY int32 v.  A butterfly is then
//         q = IMAD.HI(v, ws);  t = IMAD(v, w, u);  u' = IMAD(q, -p, t);  v' = IADD3(u, u, -u')
//    -- 4 instructions instead of 5, and no "+2p" offsets.  Bounds: every stage widens a value by at
//    most 5p/4; forward: table stages leave |x| < 3p/2 + 128, + 8 stages -> < 11.6 p; inverse: key
//    products |s| < 2.45 p, first stage reduced, + 9 stages -> < 14.5 p; 2^31 > 16 p
//    (tests/emul/emul_ntt16s.cpp propagates the intervals over the real butterfly structure).
//  * both primes are Solinas numbers p = 2^27 - 2^k + 1 (k = 11, 21).  Computing q p as
//    (q << 27) - (q << k) + q on the ALU pipe (3 multiply slots instead of 4) was built and MEASURED
//    SLOWER in every mix (profiles/r2_probe_s_v1_first_signed_kernel.log, r2_ablate_v1.log) -- it
//    trades one IMAD for three ALU instructions, the wrong direction under the cost model above.
//  * forward passes: [table stages 0,1 in layout A] -> [strides 128..16 in layout D] -> [8..1 in
//    layout C]; inverse: C -> D -> A.  Every exchange moves 128-bit vectors on at least one side,
//    and the D<->C exchanges stay inside a half-warp (no barrier, __syncwarp only).
//  * twiddles come as 128-bit vectors {w, ws, w', ws'} fetched stage by stage (at most 18 twiddle
//    registers live instead of 32).
//
// Thread T (0..63), register k (0..15); e = coefficient / position index:
//   layout A: e = T + 64 k                        layout C: e = 16 T + k
//   layout D: e = 256 (T>>4) + (T&15) + 16 k
// An exchange array is 64 chunks of 20 words (16 used), each 16-thread group skewed by 16 more
// words (half the banks): chunk(T) = 20 T + 16 (T>>4).  A chunk is only ever READ by its owner T
// with four LDS.128; writers scatter vectors or words into other threads' chunks.  All accesses are
// bank-conflict free (checked in tests/emul/emul_ntt16s.cpp).
#pragma once
#include <cstddef>
#include "ntt_core.cuh"

namespace b2 {

struct alignas(8) Tws { int32_t w, ws; };            // centred twiddle, rounded Shoup companion
struct alignas(16) V4 { int32_t x, y, z, w; };       // two Tws / four residues
struct alignas(8) I2 { int32_t x, y; };

constexpr int kSolK[2] = {11, 21};                   // p_q = 2^27 - 2^k + 1
static_assert(Prime<0>::p == (1u << 27) - (1u << 11) + 1u && Prime<1>::p == (1u << 27) - (1u << 21) + 1u, "Solinas primes");
static_assert(16ull * Prime<2>::p > (1ull << 31) - (1ull << 24), "third prime: 2^31 / p > 15.9");

constexpr int kChunkWords = 20;
constexpr int kXsWords = 64 * kChunkWords + 48;      // 1328 words per exchange array
B2_HD int chunk_of(int T) { return kChunkWords * T + 16 * (T >> 4); }

B2_HD int32_t smulhi(int32_t a, int32_t b)
{
#ifdef __CUDA_ARCH__
    return __mulhi(a, b);
#else
    return (int32_t)(((int64_t)a * (int64_t)b) >> 32);
#endif
}

// v * w mod p in [-p/4, 5p/4) for any int32 v
B2_HD int32_t sshoup(int32_t v, int32_t w, int32_t ws, int32_t p)
{
    const uint32_t q = (uint32_t)smulhi(v, ws);
    return (int32_t)((uint32_t)v * (uint32_t)w - q * (uint32_t)p);
}
// (u, v) -> (u + w v, u - w v)
B2_HD void sbfly(int32_t& u, int32_t& v, const Tws t, const int32_t p)
{
    const uint32_t q = (uint32_t)smulhi(v, t.ws);
    const uint32_t s = (uint32_t)v * (uint32_t)t.w + (uint32_t)u;
    const uint32_t un = s - q * (uint32_t)p;
    v = (int32_t)((uint32_t)u + (uint32_t)u - un);
    u = (int32_t)un;
}
// the Solinas form of the same product (kept for the emulator's equality check and the record)
B2_HD int32_t sshoup_solinas(int32_t v, int32_t w, int32_t ws, int k)
{
    const uint32_t q = (uint32_t)smulhi(v, ws);
    return (int32_t)((uint32_t)v * (uint32_t)w + (q << k) - ((q << 27) + q));
}
// v - round(v / 2^27) p for |v| < 2^31 - 2^26: |result| <= 2^26 + 16 (2^k - 1)  (< 0.51 p for k = 11, < 0.76 p for k = 21)
B2_HD int32_t sreduce(int32_t v, int32_t p)
{
    const int32_t h = (int32_t)((uint32_t)v + (1u << 26)) >> 27;
    return (int32_t)((uint32_t)v - (uint32_t)h * (uint32_t)p);
}
// twiddle 1: (u, v) -> (u + v, u - v) with v reduced first
B2_HD void sbfly_unit(int32_t& u, int32_t& v, const int32_t p)
{
    const int32_t r = sreduce(v, p);
    const int32_t a = u;
    u = (int32_t)((uint32_t)a + (uint32_t)r);
    v = (int32_t)((uint32_t)a - (uint32_t)r);
}

// a0 b0 + a1 b1 in one Montgomery reduction: |a| < 2^31 (lazy), 0 <= b < p (key word, Montgomery
// form): |sum| < 2^32 p, result = sum / 2^32 mod p with |result| < |sum| / 2^32 + p/2.
B2_HD int32_t smontmul2(int32_t a0, uint32_t b0, int32_t a1, uint32_t b1, int32_t p, uint32_t pinv)
{
    const int64_t prod = (int64_t)a0 * (int64_t)(int32_t)b0 + (int64_t)a1 * (int64_t)(int32_t)b1;
    const int32_t m = (int32_t)((uint32_t)prod * pinv);
    return (int32_t)(prod >> 32) - smulhi(m, p);
}
// all four TRGSW rows in one reduction: |sum| < 2^33 p, |result| < |sum| / 2^32 + p/2
B2_HD int32_t smontmul4(int32_t a0, uint32_t b0, int32_t a1, uint32_t b1, int32_t a2, uint32_t b2, int32_t a3, uint32_t b3,
                        int32_t p, uint32_t pinv)
{
    const int64_t prod = (int64_t)a0 * (int64_t)(int32_t)b0 + (int64_t)a1 * (int64_t)(int32_t)b1 +
                         (int64_t)a2 * (int64_t)(int32_t)b2 + (int64_t)a3 * (int64_t)(int32_t)b3;
    const int32_t m = (int32_t)((uint32_t)prod * pinv);
    return (int32_t)(prod >> 32) - smulhi(m, p);
}
B2_HD int32_t smontmul1(int32_t a0, uint32_t b0, int32_t p, uint32_t pinv)
{
    const int64_t prod = (int64_t)a0 * (int64_t)(int32_t)b0;
    const int32_t m = (int32_t)((uint32_t)prod * pinv);
    return (int32_t)(prod >> 32) - smulhi(m, p);
}

// centred CRT of signed lazy residues (|s| < 1.5 * 2^27): the integer S, |S| < P/2 - 2^28, truncated
// to 32 bits == the reference's lift rule (include/gatebootstrapping_gpu.cuh:258,270-281)
B2_HD uint32_t scrt_lift32(int32_t s0, int32_t s1, int32_t c, int32_t cs)
{
    constexpr int32_t p1 = (int32_t)Prime<1>::p;
    int32_t u = sshoup(s1 - s0, c, cs, p1);                  // (s1 - s0) / p0 mod p1 in [-p1/4, 5 p1/4)
    u = u > (p1 - 1) / 2 ? u - p1 : u;
    return (uint32_t)s0 + (uint32_t)u * Prime<0>::p;
}

// three primes (Garner, centred at every level): the integer S, |S| < P/2 ~ 2^80, truncated to 32 bits
B2_HD uint32_t scrt3_lift32(int32_t s0, int32_t s1, int32_t s2, const Tws c01, const Tws c02, const Tws c12, uint32_t p01)
{
    constexpr int32_t p1 = (int32_t)Prime<1>::p, p2 = (int32_t)Prime<2>::p;
    int32_t u1 = sshoup(s1 - s0, c01.w, c01.ws, p1);
    u1 = u1 > (p1 - 1) / 2 ? u1 - p1 : u1;                     // S = s0 + u1 p0 (mod p0 p1), centred
    const int32_t t2 = sshoup(s2 - s0, c02.w, c02.ws, p2);    // (s2 - s0) / p0 mod p2
    int32_t u2 = sshoup(t2 - u1, c12.w, c12.ws, p2);          // ((s2 - s0) / p0 - u1) / p1 mod p2
    u2 = u2 > (p2 - 1) / 2 ? u2 - p2 : u2;
    return (uint32_t)s0 + (uint32_t)u1 * Prime<0>::p + (uint32_t)u2 * p01;
}

// ---- tables ---------------------------------------------------------------------------------
// Sixteen Tws per pass and thread = eight V4 vectors {w, ws, w', ws'}; index (2^s - 1) + g.
// [vec][T] order: a warp's LDS.128 is 512 contiguous bytes.  The three per-thread
// sets share one base pointer.
constexpr int kVecF3 = 0, kVecI3 = 8, kVecSc = 14, kVecPerT = 22;
struct PrimeSTables {           // one prime; the shared-memory image a team works from
    V4 perT[kVecPerT][64];      // [vec][T]: forward strides 8..1 (layout C) | inverse t = 256 (slots 0..3), 512 (4..11) | n^-1 psi^-(T + 64 k)
    V4 i2[8][16];               // [vec][lo]      inverse t = 16..128     (layout D)
    V4 f2[4][8];                // [hi][vec]      forward strides 128..16 (layout D); half-warp broadcast
    // gadget-digit tables indexed by the RAW digit byte d (standing for d - 128), centred residues:
    //   lut0 (4-byte entries: a warp's random look-ups spread over all 32 banks, measured 0.9 % of a blind
    //         rotation against 8-byte entries with an unused half, profiles/r2_ablate_v5_rejected_experiments.log):
    //         W1 (d-128)
    //   lut12[0]: { W2a (d-128), W2b (d-128) };  lut12[1]: { W2a W1 (d-128), W2b W1 (d-128) }
    int32_t lut0[256];
    I2 lut12[2][256];
};
constexpr int kLut1Off = 1024, kLut2Off = 1024 + 2048;    // byte offsets of lut12[0], lut12[1] from lut0
static_assert(sizeof(PrimeSTables) == 22528 + 2048 + 512 + 1024 + 4096, "table layout");
static_assert(offsetof(PrimeSTables, lut12) - offsetof(PrimeSTables, lut0) == kLut1Off, "table layout");
struct TeamSTables { PrimeSTables pr[2]; };          // the bootstrapping path: two primes
struct TeamSTables3 { PrimeSTables pr[3]; };         // the exact polynomial product: three
constexpr int kTeamSTableWords = sizeof(TeamSTables) / 4;
struct TeamSConst {
    Tws iv1[3][16];         // inverse t = 1..8, warp-uniform: index (2^u - 1) + j (those with j = 0 are 1)
    Tws fw01[3][4];         // forward stages 0, 1 for general inputs: psi^brv(1), psi^brv(2), psi^brv(3)
    Tws crt;                // p0^-1 mod p1
    Tws crt02, crt12;       // p0^-1 mod p2, p1^-1 mod p2 (three-prime Garner)
    uint32_t p01;           // p0 p1 mod 2^32
    int32_t r2[3];          // 2^64 mod p (to Montgomery form), centred
};

inline int32_t h_centre(uint32_t v, uint32_t p) { return v > p / 2 ? (int32_t)v - (int32_t)p : (int32_t)v; }
inline Tws h_tws(uint32_t v, uint32_t p)
{
    const int64_t w = h_centre(v % p, p);
    // round(w 2^32 / p), floor division of (2 w 2^32 + p) by 2p
    const __int128 num = ((__int128)w << 33) + p;
    const __int128 den = 2 * (__int128)p;
    __int128 qv = num / den;
    if (num % den != 0 && num < 0) qv -= 1;
    return Tws{(int32_t)w, (int32_t)qv};
}

inline void make_prime_tables16s(PrimeSTables& L, TeamSConst& C, int q)
{
    const uint32_t p = kP[q];
    const uint32_t psi = h_psi(p);
    const uint32_t ipsi = h_powmod(psi, p - 2, p);
    const uint32_t iw = h_mulmod(ipsi, ipsi, p);
    const uint32_t ninv = h_powmod(kN, p - 2, p);
    auto psirev = [&](unsigned idx) { return h_powmod(psi, brv10(idx), p); };
    auto put = [&](V4* vec, int vstride, int idx, Tws t) {      // Tws number idx of a set
        V4& v = vec[(idx >> 1) * vstride];
        if (idx & 1) { v.z = t.w; v.w = t.ws; } else { v.x = t.w; v.y = t.ws; }
    };
    // forward stage st (2^st groups): element e uses psi^brv(2^st + (e >> (10 - st)))
    for (int hi = 0; hi < 4; hi++) {
        for (int s = 0; s < 4; s++)
            for (int g = 0; g < (1 << s); g++)
                put(&L.f2[hi][0], 1, (1 << s) - 1 + g, h_tws(psirev((4u << s) + ((unsigned)hi << s) + g), p));
        put(&L.f2[hi][0], 1, 15, Tws{0, 0});
    }
    for (int T = 0; T < 64; T++) {
        for (int s = 0; s < 4; s++)
            for (int g = 0; g < (1 << s); g++)
                put(&L.perT[kVecF3][T], 64, (1 << s) - 1 + g, h_tws(psirev((64u << s) + ((unsigned)T << s) + g), p));
        put(&L.perT[kVecF3][T], 64, 15, Tws{0, 0});
    }
    // inverse DIT: half-size t uses w^-(j 512 / t), j = e mod t
    for (int u = 0; u < 4; u++)
        for (int j = 0; j < (1 << u); j++) C.iv1[q][(1 << u) - 1 + j] = h_tws(h_powmod(iw, (uint64_t)j * (512 >> u), p), p);
    C.iv1[q][15] = Tws{0, 0};
    for (int lo = 0; lo < 16; lo++) {
        for (int u = 0; u < 4; u++)                   // t = 16 << u, j = lo + 16 g
            for (int g = 0; g < (1 << u); g++)
                put(&L.i2[0][lo], 16, (1 << u) - 1 + g, h_tws(h_powmod(iw, (uint64_t)(lo + 16 * g) * (32 >> u), p), p));
        put(&L.i2[0][lo], 16, 15, Tws{0, 0});
    }
    for (int T = 0; T < 64; T++) {
        for (int g = 0; g < 4; g++) put(&L.perT[kVecI3][T], 64, g, h_tws(h_powmod(iw, (uint64_t)(T + 64 * g) * 2, p), p));   // t = 256
        for (int g = 0; g < 8; g++) put(&L.perT[kVecI3][T], 64, 4 + g, h_tws(h_powmod(iw, (uint64_t)(T + 64 * g), p), p));   // t = 512
        for (int k = 0; k < 16; k++) put(&L.perT[kVecSc][T], 64, k, h_tws(h_mulmod(ninv, h_powmod(ipsi, T + 64 * k, p), p), p));
    }
    const uint32_t w1 = psirev(1), w2a = psirev(2), w2b = psirev(3);
    C.fw01[q][0] = h_tws(w1, p); C.fw01[q][1] = h_tws(w2a, p); C.fw01[q][2] = h_tws(w2b, p); C.fw01[q][3] = Tws{0, 0};
    C.r2[q] = h_centre((uint32_t)((((unsigned __int128)1) << 64) % p), p);
    for (int d = 0; d < 256; d++) {
        const uint32_t v = d >= 128 ? (uint32_t)(d - 128) : p - (uint32_t)(128 - d);     // d - 128 mod p
        L.lut0[d] = h_centre(h_mulmod(w1, v, p), p);
        L.lut12[0][d] = I2{h_centre(h_mulmod(w2a, v, p), p), h_centre(h_mulmod(w2b, v, p), p)};
        L.lut12[1][d] = I2{h_centre(h_mulmod(h_mulmod(w2a, w1, p), v, p), p), h_centre(h_mulmod(h_mulmod(w2b, w1, p), v, p), p)};
    }
}
inline void make_tables16s(TeamSTables3& L, TeamSConst& C)
{
    for (int q = 0; q < 3; q++) make_prime_tables16s(L.pr[q], C, q);
    C.crt = h_tws(h_powmod(kP[0] % kP[1], kP[1] - 2, kP[1]), kP[1]);
    C.crt02 = h_tws(h_powmod(kP[0] % kP[2], kP[2] - 2, kP[2]), kP[2]);
    C.crt12 = h_tws(h_powmod(kP[1] % kP[2], kP[2] - 2, kP[2]), kP[2]);
    C.p01 = kP[0] * kP[1];
}

B2_HD Tws tws_lo(const V4& v) { return Tws{v.x, v.y}; }
B2_HD Tws tws_hi(const V4& v) { return Tws{v.z, v.w}; }

// ---- in-register passes ---------------------------------------------------------------------
// Loader functors: ld(v) returns twiddle vector v of the pass (Tws 2v and 2v + 1); vectors are
// fetched stage by stage so that at most 18 twiddle registers are live.
//
// Table stages 0 and 1 on gadget digits, layout A.  The four digits of a radix-4 group
// (k, k+4, k+8, k+12): x[k] holds the SIGNED digit, the other three hold table HANDLES (whatever
// the look-up functors take: the raw digit byte on the host, a shared-memory address on the
// device).  Three look-ups + six 3-input additions.  |output| < 3p/2 + 128.
template <class L0, class L1, class L2>
B2_HD void spass_digits(int32_t (&x)[16], L0 l0, L1 l1, L2 l2)
{
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int32_t d0 = x[k], t1 = l0(x[k + 8]);
        const I2 u = l1(x[k + 4]), v = l2(x[k + 12]);
        const int32_t B = u.x + v.x, D = u.y - v.y;       // W2a (x1 + W1 x3),  W2b (x1 - W1 x3)
        x[k] = d0 + t1 + B;                               // x0 + W1 x2 + B
        x[k + 4] = d0 + t1 - B;
        x[k + 8] = d0 - t1 + D;                           // x0 - W1 x2 + D
        x[k + 12] = d0 - t1 - D;
    }
}
// stages 0 and 1 by butterflies for general inputs (key conversion, standalone transforms):
// w1, w2a, w2b = psi^brv(1), psi^brv(2), psi^brv(3) (TeamSConst::fw01)
B2_HD void spass_fwd01(int32_t (&x)[16], const Tws w1, const Tws w2a, const Tws w2b, const int32_t p)
{
#pragma unroll
    for (int k = 0; k < 8; k++) sbfly(x[k], x[k + 8], w1, p);
#pragma unroll
    for (int k = 0; k < 4; k++) { sbfly(x[k], x[k + 4], w2a, p); sbfly(x[k + 8], x[k + 12], w2b, p); }
}
// one stage of NP (1 or 2) polynomials: butterflies (k, k + tp), twiddle tw(k)
template <int NP, class TW>
B2_HD void sstage(int32_t (&x0)[16], int32_t (&x1)[16], const int tp, TW tw, const int32_t p)
{
#pragma unroll
    for (int k = 0; k < 16; k++)
        if ((k & tp) == 0) {
            const Tws t = tw(k);
            sbfly(x0[k], x0[k + tp], t, p);
            if (NP == 2) sbfly(x1[k], x1[k + tp], t, p);
        }
}
// four forward stages on register strides 8, 4, 2, 1; twiddle (2^s - 1) + k / (2 tp)
template <int NP, class LD>
B2_HD void spass_fwd(int32_t (&x0)[16], int32_t (&x1)[16], LD ld, const int32_t p)
{
    const V4 c0 = ld(0), c1 = ld(1);
    sstage<NP>(x0, x1, 8, [&](int) { return tws_lo(c0); }, p);
    const V4 c2 = ld(2), c3 = ld(3);
    sstage<NP>(x0, x1, 4, [&](int k) { return (k >> 3) ? tws_lo(c1) : tws_hi(c0); }, p);
    const V4 c4 = ld(4), c5 = ld(5), c6 = ld(6), c7 = ld(7);
    sstage<NP>(x0, x1, 2, [&](int k) { const int g = k >> 2; return g == 0 ? tws_hi(c1) : g == 1 ? tws_lo(c2) : g == 2 ? tws_hi(c2) : tws_lo(c3); }, p);
    sstage<NP>(x0, x1, 1, [&](int k) {
        const int g = k >> 1;           // Tws 7 + g
        return g == 0 ? tws_hi(c3) : g == 1 ? tws_lo(c4) : g == 2 ? tws_hi(c4) : g == 3 ? tws_lo(c5) : g == 4 ? tws_hi(c5)
             : g == 5 ? tws_lo(c6) : g == 6 ? tws_hi(c6) : tws_lo(c7); }, p);
}
// inverse pass I1: register strides 1, 2, 4, 8 with WARP-UNIFORM twiddles tw(i), i = (2^u - 1) + (k mod tp);
// twiddle 0 of every stage is 1: plain reduction instead of the Shoup product
template <int NP, class TW>
B2_HD void spass_inv1(int32_t (&x0)[16], int32_t (&x1)[16], TW tw, const int32_t p)
{
#pragma unroll
    for (int u = 0; u < 4; u++) {
        const int tp = 1 << u;
#pragma unroll
        for (int k = 0; k < 16; k++)
            if ((k & tp) == 0) {
                if ((k & (tp - 1)) == 0) {
                    sbfly_unit(x0[k], x0[k + tp], p);
                    if (NP == 2) sbfly_unit(x1[k], x1[k + tp], p);
                } else {
                    const Tws t = tw((1 << u) - 1 + (k & (tp - 1)));
                    sbfly(x0[k], x0[k + tp], t, p);
                    if (NP == 2) sbfly(x1[k], x1[k + tp], t, p);
                }
            }
    }
}
// inverse pass I2: register strides 1, 2, 4, 8; twiddle (2^u - 1) + (k mod tp) from eight vectors
template <int NP, class LD>
B2_HD void spass_inv2(int32_t (&x0)[16], int32_t (&x1)[16], LD ld, const int32_t p)
{
    const V4 c0 = ld(0), c1 = ld(1);
    sstage<NP>(x0, x1, 1, [&](int) { return tws_lo(c0); }, p);
    const V4 c2 = ld(2), c3 = ld(3);
    sstage<NP>(x0, x1, 2, [&](int k) { return (k & 1) ? tws_lo(c1) : tws_hi(c0); }, p);
    const V4 c4 = ld(4), c5 = ld(5), c6 = ld(6), c7 = ld(7);
    sstage<NP>(x0, x1, 4, [&](int k) { const int g = k & 3; return g == 0 ? tws_hi(c1) : g == 1 ? tws_lo(c2) : g == 2 ? tws_hi(c2) : tws_lo(c3); }, p);
    sstage<NP>(x0, x1, 8, [&](int k) {
        const int g = k & 7;            // Tws 7 + g
        return g == 0 ? tws_hi(c3) : g == 1 ? tws_lo(c4) : g == 2 ? tws_hi(c4) : g == 3 ? tws_lo(c5) : g == 4 ? tws_hi(c5)
             : g == 5 ? tws_lo(c6) : g == 6 ? tws_hi(c6) : tws_lo(c7); }, p);
}
// I3: register strides 4 (t = 256, Tws 0..3) then 8 (t = 512, Tws 4..11) from six vectors, then the
// sixteen scale factors from eight vectors
template <int NP, class LD, class LS>
B2_HD void spass_inv3(int32_t (&x0)[16], int32_t (&x1)[16], LD ld, LS ls, const int32_t p)
{
    const V4 c0 = ld(0), c1 = ld(1);
    sstage<NP>(x0, x1, 4, [&](int k) { const int g = k & 3; return g == 0 ? tws_lo(c0) : g == 1 ? tws_hi(c0) : g == 2 ? tws_lo(c1) : tws_hi(c1); }, p);
    const V4 c2 = ld(2), c3 = ld(3), c4 = ld(4), c5 = ld(5);
    sstage<NP>(x0, x1, 8, [&](int k) {
        const int g = k & 7;            // Tws 4 + g
        return g == 0 ? tws_lo(c2) : g == 1 ? tws_hi(c2) : g == 2 ? tws_lo(c3) : g == 3 ? tws_hi(c3) : g == 4 ? tws_lo(c4)
             : g == 5 ? tws_hi(c4) : g == 6 ? tws_lo(c5) : tws_hi(c5); }, p);
#pragma unroll
    for (int v = 0; v < 8; v++) {
        const V4 f = ls(v);
        x0[2 * v] = sshoup(x0[2 * v], f.x, f.y, p);
        x0[2 * v + 1] = sshoup(x0[2 * v + 1], f.z, f.w, p);
        if (NP == 2) {
            x1[2 * v] = sshoup(x1[2 * v], f.x, f.y, p);
            x1[2 * v + 1] = sshoup(x1[2 * v + 1], f.z, f.w, p);
        }
    }
}

// ---- exchange maps (word offsets inside one array) --------------------------------------------
// A -> D : writer T sends vector kh = {x[4 kh .. 4 kh + 3]} to chunk(16 kh + (T&15)) slot 4 (T>>4);
//          reader's vector th lands in x[th], x[th+4], x[th+8], x[th+12]
B2_HD int xs_AD_w(int T, int kh) { return chunk_of(16 * kh + (T & 15)) + 4 * (T >> 4); }
// D -> C and C -> D : writer (g = T>>4, t = T&15) sends word k to chunk(16 g + k) slot t; reader's
//          vector v lands in x[4 v .. 4 v + 3]
B2_HD int xs_DC_w(int T, int k) { return chunk_of(16 * (T >> 4) + k) + (T & 15); }
// D -> A : writer (hi = T>>4, lo = T&15) sends vector kl = {x[kl], x[kl+4], x[kl+8], x[kl+12]} to
//          chunk(lo + 16 kl) slot 4 hi; reader's vector h lands in x[4 h .. 4 h + 3]
B2_HD int xs_DA_w(int T, int kl) { return chunk_of((T & 15) + 16 * kl) + 4 * (T >> 4); }

}  // namespace b2
