// ntt16s.cuh -- second-generation 1024-point negacyclic NTT for a TEAM of 64 threads x 16
// coefficients: SIGNED lazy residues, Solinas primes, vectorised exchanges.
//
// Same transform as ntt16.cuh / ntt_core.cuh (merged-psi Cooley-Tukey forward, natural ->
// bit-reversed; cyclic DIT inverse with w^-1 followed by n^-1 psi^-e), same NTT-domain
// positions, so the bootstrapping key (bk16_slot order, Montgomery form in [0, p)) is shared.
// Replaces SmallForwardNTT_1024 / SmallInverseNTT_1024 (reference
// include/ntt_gpu/ntt_gpuntt.cuh:232-276,342-392).  What changed and why (VERDICT r1, weak 5:
// the multiply pipe was 71 % busy at 1.74 issued instructions per multiply slot):
//
//  * residues are int32 in (-2^31, 2^31).  Shoup multiplication with a CENTRED twiddle w and the
//    ROUNDED companion ws = round(w 2^32 / p):  q = mulhi.s32(v, ws), r = v w - q p lies in
//    [-p/4, 5p/4) for ANY int32 v.  A butterfly is then
//         q = IMAD.HI(v, ws);  t = IMAD(v, w, u);  u' = IMAD(q, -p, t);  v' = IADD3(u, u, -u')
//    -- 4 instructions instead of 5, and no "+2p" offsets.  Bounds: every stage widens a value by at
//    most 5p/4; forward: table stages leave |x| < 3p/2 + 128, + 8 stages -> < 11.6 p; inverse: key
//    products |s| < 2.45 p, first stage reduced, + 9 stages -> < 14.5 p; 2^31 > 16 p.
//  * both primes are Solinas numbers p = 2^27 - 2^k + 1 (k = 11, 21), so q p = (q << 27) - (q << k) + q:
//    the SOL form of the butterfly trades the third multiply for SHF + LEA + IADD3 on the ALU pipe
//    (3 multiply-pipe slots instead of 4); which passes use it is a template mask, chosen by
//    measurement so that the two pipes balance.
//  * forward passes: [table stages 0,1 in layout A] -> [strides 128..16 in layout D] -> [8..1 in
//    layout C]; inverse: C -> D -> A as before.  Every exchange moves 128-bit vectors on at least
//    one side, and the D<->C exchanges stay inside a half-warp (no barrier, __syncwarp only).
//
// Thread T (0..63), register k (0..15); e = coefficient / position index:
//   layout A: e = T + 64 k                        layout C: e = 16 T + k
//   layout D: e = 256 (T>>4) + (T&15) + 16 k
// An exchange array is 64 chunks of 20 words (16 used), each 16-thread group skewed by 16 more
// words (half the banks): chunk(T) = 20 T + 16 (T>>4).  A chunk is only ever READ by its owner T with four
// LDS.128; writers scatter vectors or words into other threads' chunks.  All accesses are
// bank-conflict free (checked in tests/emul/emul_ntt16s.cpp).
#pragma once
#include "ntt16.cuh"

namespace b2 {

struct alignas(8) Tws { int32_t w, ws; };            // centred twiddle, rounded Shoup companion
struct alignas(16) V4 { int32_t x, y, z, w; };       // two Tws / four residues
struct alignas(8) I2 { int32_t x, y; };

constexpr int kSolK[2] = {11, 21};                   // p_q = 2^27 - 2^k + 1
static_assert(Prime<0>::p == (1u << 27) - (1u << 11) + 1u && Prime<1>::p == (1u << 27) - (1u << 21) + 1u, "Solinas primes");

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
template <bool SOL>
B2_HD int32_t sshoup(int32_t v, int32_t w, int32_t ws, int32_t p, int k)
{
    const uint32_t q = (uint32_t)smulhi(v, ws);
    const uint32_t t = (uint32_t)v * (uint32_t)w;
    return (int32_t)(SOL ? t + (q << k) - ((q << 27) + q) : t - q * (uint32_t)p);
}
// (u, v) -> (u + w v, u - w v)
template <bool SOL>
B2_HD void sbfly(int32_t& u, int32_t& v, const Tws t, const int32_t p, const int k)
{
    const uint32_t q = (uint32_t)smulhi(v, t.ws);
    const uint32_t s = (uint32_t)v * (uint32_t)t.w + (uint32_t)u;
    const uint32_t un = SOL ? s + (q << k) - ((q << 27) + q) : s - q * (uint32_t)p;
    v = (int32_t)((uint32_t)u + (uint32_t)u - un);
    u = (int32_t)un;
}
// v - round(v / 2^27) p: |result| <= 2^26 + 16 (2^k - 1)  (< 0.51 p for k = 11, < 0.76 p for k = 21)
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
    int32_t u = sshoup<false>(s1 - s0, c, cs, p1, 0);        // (s1 - s0) / p0 mod p1 in [-p1/4, 5 p1/4)
    u = u > (p1 - 1) / 2 ? u - p1 : u;
    return (uint32_t)s0 + (uint32_t)u * Prime<0>::p;
}

// ---- tables ---------------------------------------------------------------------------------
// Sixteen Tws per pass and thread = eight V4 vectors {w, ws, w', ws'}; index (2^s - 1) + g as in
// ntt16.cuh.  [vec][T] order: a warp's LDS.128 is 512 contiguous bytes.
struct TeamSTables {
    V4 f2[2][4][8];         // [q][hi][vec]      forward strides 128..16 (layout D); half-warp broadcast
    V4 f3[2][8][64];        // [q][vec][T]       forward strides 8..1    (layout C)
    V4 i2[2][8][16];        // [q][vec][lo]      inverse t = 16..128     (layout D)
    V4 i3[2][6][64];        // [q][vec][T]       inverse t = 256 (slots 0..3), 512 (slots 4..11)
    V4 sc[2][8][64];        // [q][vec][T]       n^-1 psi^-(T + 64 k), k = 2 vec, 2 vec + 1
    int32_t lut1[2][256];   // gadget-digit tables (entry d stands for d - 128), centred residues:
    I2 lut24[2][256];       //   W1 (d-128);  { W2a (d-128), W2b (d-128) };  { W2a W1 (d-128), W2b W1 (d-128) }
    I2 lut35[2][256];
};
constexpr int kTeamSTableWords = sizeof(TeamSTables) / 4;
static_assert(sizeof(TeamSTables) == 1024 + 16384 + 4096 + 12288 + 16384 + 10240, "table layout");
struct TeamSConst {
    Tws iv1[2][16];         // inverse t = 1..8, warp-uniform: index (2^u - 1) + j (those with j = 0 are 1)
    Tws fw01[2][4];         // forward stages 0, 1 for general inputs: psi^brv(1), psi^brv(2), psi^brv(3)
    Tws crt;                // p0^-1 mod p1
    int32_t r2[2];          // 2^64 mod p (to Montgomery form), centred
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

inline void make_tables16s(TeamSTables& L, TeamSConst& C)
{
    for (int q = 0; q < 2; q++) {
        const uint32_t p = kP[q];
        const uint32_t psi = h_psi(p);
        const uint32_t ipsi = h_powmod(psi, p - 2, p);
        const uint32_t iw = h_mulmod(ipsi, ipsi, p);
        const uint32_t ninv = h_powmod(kN, p - 2, p);
        auto psirev = [&](unsigned idx) { return h_powmod(psi, brv10(idx), p); };
        auto put = [&](V4* vec, int vstride, int idx, Tws t) {      // Tws number idx of a 16-entry set
            V4& v = vec[(idx >> 1) * vstride];
            if (idx & 1) { v.z = t.w; v.w = t.ws; } else { v.x = t.w; v.y = t.ws; }
        };
        // forward stage st (2^st groups): element e uses psi^brv(2^st + (e >> (10 - st)))
        for (int hi = 0; hi < 4; hi++) {
            for (int s = 0; s < 4; s++)
                for (int g = 0; g < (1 << s); g++)
                    put(&L.f2[q][hi][0], 1, (1 << s) - 1 + g, h_tws(psirev((4u << s) + ((unsigned)hi << s) + g), p));
            put(&L.f2[q][hi][0], 1, 15, Tws{0, 0});
        }
        for (int T = 0; T < 64; T++) {
            for (int s = 0; s < 4; s++)
                for (int g = 0; g < (1 << s); g++)
                    put(&L.f3[q][0][T], 64, (1 << s) - 1 + g, h_tws(psirev((64u << s) + ((unsigned)T << s) + g), p));
            put(&L.f3[q][0][T], 64, 15, Tws{0, 0});
        }
        // inverse DIT: half-size t uses w^-(j 512 / t), j = e mod t
        for (int u = 0; u < 4; u++)
            for (int j = 0; j < (1 << u); j++) C.iv1[q][(1 << u) - 1 + j] = h_tws(h_powmod(iw, (uint64_t)j * (512 >> u), p), p);
        C.iv1[q][15] = Tws{0, 0};
        for (int lo = 0; lo < 16; lo++) {
            for (int u = 0; u < 4; u++)                   // t = 16 << u, j = lo + 16 g
                for (int g = 0; g < (1 << u); g++)
                    put(&L.i2[q][0][lo], 16, (1 << u) - 1 + g, h_tws(h_powmod(iw, (uint64_t)(lo + 16 * g) * (32 >> u), p), p));
            put(&L.i2[q][0][lo], 16, 15, Tws{0, 0});
        }
        for (int T = 0; T < 64; T++) {
            for (int g = 0; g < 4; g++) put(&L.i3[q][0][T], 64, g, h_tws(h_powmod(iw, (uint64_t)(T + 64 * g) * 2, p), p));   // t = 256
            for (int g = 0; g < 8; g++) put(&L.i3[q][0][T], 64, 4 + g, h_tws(h_powmod(iw, (uint64_t)(T + 64 * g), p), p));   // t = 512
            for (int k = 0; k < 16; k++) put(&L.sc[q][0][T], 64, k, h_tws(h_mulmod(ninv, h_powmod(ipsi, T + 64 * k, p), p), p));
        }
        const uint32_t w1 = psirev(1), w2a = psirev(2), w2b = psirev(3);
        C.fw01[q][0] = h_tws(w1, p); C.fw01[q][1] = h_tws(w2a, p); C.fw01[q][2] = h_tws(w2b, p); C.fw01[q][3] = Tws{0, 0};
        C.r2[q] = h_centre((uint32_t)((((unsigned __int128)1) << 64) % p), p);
        for (int d = 0; d < 256; d++) {
            const uint32_t v = d >= 128 ? (uint32_t)(d - 128) : p - (uint32_t)(128 - d);     // d - 128 mod p
            L.lut1[q][d] = h_centre(h_mulmod(w1, v, p), p);
            L.lut24[q][d] = I2{h_centre(h_mulmod(w2a, v, p), p), h_centre(h_mulmod(w2b, v, p), p)};
            L.lut35[q][d] = I2{h_centre(h_mulmod(h_mulmod(w2a, w1, p), v, p), p), h_centre(h_mulmod(h_mulmod(w2b, w1, p), v, p), p)};
        }
    }
    C.crt = h_tws(h_powmod(kP[0] % kP[1], kP[1] - 2, kP[1]), kP[1]);
}

// Tws number idx of a 16-entry set held as eight V4
B2_HD Tws tws_of(const V4 (&tv)[8], int idx)
{
    const V4& v = tv[idx >> 1];
    return (idx & 1) ? Tws{v.z, v.w} : Tws{v.x, v.y};
}

// ---- in-register passes ---------------------------------------------------------------------
// Table stages 0 and 1 on raw gadget digits x[k] in [0, 256) (meaning x[k] - 128), layout A:
// the four digits of a radix-4 group (k, k+4, k+8, k+12) -> five table entries + eight additions.
// |output| < 3p/2 + 128.
B2_HD void spass_digits(int32_t (&x)[16], const int32_t* l1, const I2* l24, const I2* l35)
{
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int32_t d0 = x[k] - 128, t1 = l1[x[k + 8]];
        const I2 u = l24[x[k + 4]], v = l35[x[k + 12]];
        const int32_t A = d0 + t1, C = d0 - t1;           // x0 +- W1 x2
        const int32_t B = u.x + v.x, D = u.y - v.y;       // W2a (x1 + W1 x3),  W2b (x1 - W1 x3)
        x[k] = A + B;
        x[k + 4] = A - B;
        x[k + 8] = C + D;
        x[k + 12] = C - D;
    }
}
// stages 0 and 1 by butterflies for general inputs (key conversion, standalone transforms):
// w1, w2a, w2b = psi^brv(1), psi^brv(2), psi^brv(3) (TeamSConst::fw01)
B2_HD void spass_fwd01(int32_t (&x)[16], const Tws w1, const Tws w2a, const Tws w2b, const int32_t p)
{
#pragma unroll
    for (int k = 0; k < 8; k++) sbfly<false>(x[k], x[k + 8], w1, p, 0);
#pragma unroll
    for (int k = 0; k < 4; k++) { sbfly<false>(x[k], x[k + 4], w2a, p, 0); sbfly<false>(x[k + 8], x[k + 12], w2b, p, 0); }
}
// four forward stages on register strides 8, 4, 2, 1; twiddle (2^s - 1) + k / (2 tp)
template <bool SOL>
B2_HD void spass_fwd(int32_t (&x)[16], const V4 (&tv)[8], const int32_t p, const int sk)
{
#pragma unroll
    for (int s = 0; s < 4; s++) {
        const int tp = 8 >> s;
#pragma unroll
        for (int k = 0; k < 16; k++)
            if ((k & tp) == 0) sbfly<SOL>(x[k], x[k + tp], tws_of(tv, (1 << s) - 1 + k / (2 * tp)), p, sk);
    }
}
template <bool SOL>
B2_HD void spass_fwd2(int32_t (&x0)[16], int32_t (&x1)[16], const V4 (&tv)[8], const int32_t p, const int sk)
{
#pragma unroll
    for (int s = 0; s < 4; s++) {
        const int tp = 8 >> s;
#pragma unroll
        for (int k = 0; k < 16; k++)
            if ((k & tp) == 0) {
                const Tws t = tws_of(tv, (1 << s) - 1 + k / (2 * tp));
                sbfly<SOL>(x0[k], x0[k + tp], t, p, sk);
                sbfly<SOL>(x1[k], x1[k + tp], t, p, sk);
            }
    }
}
// inverse register strides 1, 2, 4, 8; twiddle (2^u - 1) + (k mod tp).  UNIT: twiddle 0 of every
// stage is 1 (pass I1, warp-uniform table): plain reduction instead of the Shoup product.
template <bool SOL, bool UNIT, class TW>
B2_HD void spass_inv(int32_t (&x)[16], const TW& tw, const int32_t p, const int sk)
{
#pragma unroll
    for (int u = 0; u < 4; u++) {
        const int tp = 1 << u;
#pragma unroll
        for (int k = 0; k < 16; k++)
            if ((k & tp) == 0) {
                if (UNIT && (k & (tp - 1)) == 0) sbfly_unit(x[k], x[k + tp], p);
                else sbfly<SOL>(x[k], x[k + tp], tw((1 << u) - 1 + (k & (tp - 1))), p, sk);
            }
    }
}
// I3: register strides 4 (t = 256, slots 0..3) then 8 (t = 512, slots 4..11), then the scale factors
template <bool SOL, bool SOLSC>
B2_HD void spass_inv3(int32_t (&x)[16], const V4 (&tv)[6], const V4 (&sv)[8], const int32_t p, const int sk)
{
    auto tw = [&](int idx) { const V4& v = tv[idx >> 1]; return (idx & 1) ? Tws{v.z, v.w} : Tws{v.x, v.y}; };
#pragma unroll
    for (int k = 0; k < 16; k++)
        if ((k & 4) == 0) sbfly<SOL>(x[k], x[k + 4], tw(k & 3), p, sk);
#pragma unroll
    for (int k = 0; k < 8; k++) sbfly<SOL>(x[k], x[k + 8], tw(4 + k), p, sk);
#pragma unroll
    for (int k = 0; k < 16; k++) {
        const Tws f = tws_of(sv, k);
        x[k] = sshoup<SOLSC>(x[k], f.w, f.ws, p, sk);
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
