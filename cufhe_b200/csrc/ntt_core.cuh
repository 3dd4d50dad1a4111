// ntt_core.cuh -- pieces shared by both generations of the 1024-point negacyclic NTT: the twiddle
// type, the warp-uniform twiddle tables of the first passes, the unsigned Shoup butterfly.
//
// Replaces the twiddle handling of SmallForwardNTT_1024 / SmallInverseNTT_1024 (reference
// include/ntt_gpu/ntt_gpuntt.cuh:232-276,342-392: twiddles fetched from global memory).
//
// forward  (merged-psi Cooley-Tukey, natural -> bit-reversed): stage s (2^s groups), group i of
//          element e is e >> (10 - s), twiddle psi^brv(2^s + i)
// inverse  (cyclic DIT with w^-1, bit-reversed -> natural, then * n^-1 psi^-c): the stage with
//          half-size t uses w^-(j 512 / t), j = e mod t
//
// First-generation butterflies (ntt16.cuh) use Shoup multiplication (twiddle w + companion ws) on
// lazy UNSIGNED residues: t = w*v mod p lands in [0, 2p) for any 32-bit v, and (u + t, u - t + 2p)
// grows the bound by 2p per stage, so with p < 2^27 no correction is ever needed:
//   forward: input < 2p (digit + p, or a canonical key residue), + 20p  -> < 22p < 2^32
//   inverse: input < 8p (sum of 4 Montgomery products + 4p),     + 20p  -> < 28p < 2^32.
#pragma once
#include "modarith.cuh"

namespace b2 {

struct alignas(8) Tw { uint32_t w, ws; };     // twiddle and its Shoup companion

struct NttTables {
    Tw fw1[2][32];          // forward stages 0..4, index (2^s - 1) + g           (31 used)
    Tw iv1[2][32];          // inverse t = 1..16, index (2^u - 1) + j             (31 used)
    uint32_t r2[2];         // 2^64 mod p  (to Montgomery form)
    Tw crt_p0inv;           // p0^-1 mod p1
};

inline unsigned brv10(unsigned x)
{
    unsigned r = 0;
    for (int i = 0; i < kNbit; i++) r |= ((x >> i) & 1u) << (kNbit - 1 - i);
    return r;
}

inline void make_tables(NttTables& T)
{
    for (int q = 0; q < 2; q++) {
        const uint32_t p = kP[q];
        const uint32_t psi = h_psi(p);
        const uint32_t ipsi = h_powmod(psi, p - 2, p);
        const uint32_t iw = h_mulmod(ipsi, ipsi, p);            // w^-1, w = psi^2
        auto M = [&](uint32_t v) { return Tw{v, h_shoup(v, p)}; };
        auto psirev = [&](unsigned idx) { return h_powmod(psi, brv10(idx), p); };
        for (int s = 0; s < 5; s++)
            for (int g = 0; g < (1 << s); g++) T.fw1[q][(1 << s) - 1 + g] = M(psirev((1u << s) + g));
        T.fw1[q][31] = Tw{0, 0};
        for (int u = 0; u < 5; u++)
            for (int j = 0; j < (1 << u); j++)
                T.iv1[q][(1 << u) - 1 + j] = M(h_powmod(iw, (uint64_t)j * (512 >> u), p));
        T.iv1[q][31] = Tw{0, 0};
        T.r2[q] = (uint32_t)((((unsigned __int128)1) << 64) % p);
    }
    {
        const uint32_t c = h_powmod(kP[0] % kP[1], kP[1] - 2, kP[1]);
        T.crt_p0inv = Tw{c, h_shoup(c, kP[1])};
    }
}

// The prime is a (warp-uniform) run-time value in the *_p forms so that both primes share
// ONE copy of the unrolled code; the <Q> forms bind it at compile time.
B2_HD void butterfly_p(uint32_t& u, uint32_t& v, const Tw t, const uint32_t p, const uint32_t p2)
{
    const uint32_t r = shoup_mul_p(v, t.w, t.ws, p);
    const uint32_t a = u;
    u = a + r;
    v = a - r + p2;
}
template <int Q>
B2_HD void butterfly(uint32_t& u, uint32_t& v, const Tw t) { butterfly_p(u, v, t, Prime<Q>::p, 2u * Prime<Q>::p); }

}  // namespace b2
