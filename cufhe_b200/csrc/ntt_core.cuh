// ntt_core.cuh -- 1024-point negacyclic NTT as two in-register radix-32 passes.
//
// Replaces SmallForwardNTT_1024 / SmallInverseNTT_1024 (reference
// include/ntt_gpu/ntt_gpuntt.cuh:232-276,342-392: 512 threads x 2 coefficients,
// 10 radix-2 stages through shared memory, twiddles fetched from global memory).
//
// B200 design: ONE WARP transforms one polynomial for one prime.  Each lane holds
// 32 coefficients in registers; five radix-2 stages run entirely in registers, the
// warp transposes the 32x32 tile once through (swizzled, conflict-free) shared
// memory, and the remaining five stages run in registers again.  No block barrier,
// one __syncwarp pair per transform.
//
//   layout A: lane L holds x[k] = a[L + 32 k]       (coefficient order, stride 32)
//   layout B: lane L holds y[k] = a[32 L + k]       (contiguous 32)
//
// forward  (merged-psi Cooley-Tukey, natural -> bit-reversed):
//     pass 1 in layout A (strides 512..32; twiddles warp-uniform -> constant bank)
//     pass 2 in layout B (strides 16..1;   31 per-lane twiddles)
// inverse  (cyclic DIT with w^-1, bit-reversed -> natural, then * n^-1 psi^-c):
//     pass 1 in layout B (strides 1..16;   twiddles warp-uniform)
//     pass 2 in layout A (strides 32..512; 31 per-lane twiddles) + 32 per-lane scales
//
// Butterflies use Shoup multiplication (twiddle w + companion ws) on lazy UNSIGNED
// residues: t = w*v mod p lands in [0, 2p) for any 32-bit v, and (u + t, u - t + 2p)
// grows the bound by 2p per stage, so with p < 2^27 no correction is ever needed:
//   forward: input < 2p (digit + p, or a canonical key residue), + 20p  -> < 22p < 2^32
//   inverse: input < 8p (sum of 4 Montgomery products + 4p),     + 20p  -> < 28p < 2^32.
#pragma once
#include "modarith.cuh"

namespace b2 {

struct alignas(8) Tw { uint32_t w, ws; };     // twiddle and its Shoup companion

// twiddle table layout shared by host generator, device constant/global copies and
// the CPU emulation.
struct NttTables {
    Tw fw1[2][32];          // forward pass 1, index (2^s - 1) + g           (31 used)
    Tw iv1[2][32];          // inverse pass 1, index (2^u - 1) + j           (31 used)
    Tw fw2[2][31][32];      // forward pass 2, [q][lane]
    Tw iv2[2][31][32];      // inverse pass 2, [q][lane]
    Tw scale[2][32][32];    // n^-1 psi^-(lane + 32 k), [k][lane]
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
        const uint32_t ninv = h_powmod(kN, p - 2, p);
        auto M = [&](uint32_t v) { return Tw{v, h_shoup(v, p)}; };
        auto psirev = [&](unsigned idx) { return h_powmod(psi, brv10(idx), p); };
        // forward: stage s (m = 2^s groups), group i uses psi^brv10(m + i)
        for (int s = 0; s < 5; s++)
            for (int g = 0; g < (1 << s); g++) T.fw1[q][(1 << s) - 1 + g] = M(psirev((1u << s) + g));
        T.fw1[q][31] = Tw{0, 0};
        for (int s = 0; s < 5; s++)                              // global stage 5 + s
            for (int g = 0; g < (1 << s); g++)
                for (int lane = 0; lane < 32; lane++)
                    T.fw2[q][(1 << s) - 1 + g][lane] = M(psirev((32u << s) + ((unsigned)lane << s) + g));
        // inverse DIT: stage with half-size t uses w^-(j * 512 / t), j = index mod t
        for (int u = 0; u < 5; u++)
            for (int j = 0; j < (1 << u); j++)
                T.iv1[q][(1 << u) - 1 + j] = M(h_powmod(iw, (uint64_t)j * (512 >> u), p));
        T.iv1[q][31] = Tw{0, 0};
        for (int u = 0; u < 5; u++)                              // t = 32 << u
            for (int g = 0; g < (1 << u); g++)
                for (int lane = 0; lane < 32; lane++)
                    T.iv2[q][(1 << u) - 1 + g][lane] = M(h_powmod(iw, (uint64_t)(lane + 32 * g) * (16 >> u), p));
        for (int k = 0; k < 32; k++)
            for (int lane = 0; lane < 32; lane++)
                T.scale[q][k][lane] = M(h_mulmod(ninv, h_powmod(ipsi, lane + 32 * k, p), p));
        T.r2[q] = (uint32_t)((((unsigned __int128)1) << 64) % p);
    }
    {
        const uint32_t c = h_powmod(kP[0] % kP[1], kP[1] - 2, kP[1]);
        T.crt_p0inv = Tw{c, h_shoup(c, kP[1])};
    }
}

// ---- in-register radix-32 passes --------------------------------------------
// tw[q * stride] is twiddle q of this lane (stride 1: warp-uniform table; stride 32:
// per-lane table with the lane offset already applied to the pointer).

// The prime is a (warp-uniform) run-time value in the *_p forms so that both primes share
// ONE copy of the unrolled code (instruction-cache footprint halves); the <Q> forms bind
// it at compile time.
B2_HD void butterfly_p(uint32_t& u, uint32_t& v, const Tw t, const uint32_t p, const uint32_t p2)
{
    const uint32_t r = shoup_mul_p(v, t.w, t.ws, p);
    const uint32_t a = u;
    u = a + r;
    v = a - r + p2;
}
template <int Q>
B2_HD void butterfly(uint32_t& u, uint32_t& v, const Tw t) { butterfly_p(u, v, t, Prime<Q>::p, 2u * Prime<Q>::p); }

// forward ordering: sub-strides 16, 8, 4, 2, 1; one twiddle per group k / (2 tp)
B2_HD void pass_fwd_p(uint32_t (&x)[32], const Tw* tw, const int stride, const uint32_t p)
{
    const uint32_t p2 = 2u * p;
#pragma unroll
    for (int s = 0; s < 5; s++) {
        const int tp = 16 >> s;
#pragma unroll
        for (int k = 0; k < 32; k++)
            if ((k & tp) == 0) butterfly_p(x[k], x[k + tp], tw[((1 << s) - 1 + k / (2 * tp)) * stride], p, p2);
    }
}
// inverse (DIT) ordering: sub-strides 1, 2, 4, 8, 16; one twiddle per offset k mod tp
B2_HD void pass_inv_p(uint32_t (&x)[32], const Tw* tw, const int stride, const uint32_t p)
{
    const uint32_t p2 = 2u * p;
#pragma unroll
    for (int u = 0; u < 5; u++) {
        const int tp = 1 << u;
#pragma unroll
        for (int k = 0; k < 32; k++)
            if ((k & tp) == 0) butterfly_p(x[k], x[k + tp], tw[((1 << u) - 1 + (k & (tp - 1))) * stride], p, p2);
    }
}
template <int Q>
B2_HD void pass_fwd(uint32_t (&x)[32], const Tw* tw, const int stride) { pass_fwd_p(x, tw, stride, Prime<Q>::p); }
template <int Q>
B2_HD void pass_inv(uint32_t (&x)[32], const Tw* tw, const int stride) { pass_inv_p(x, tw, stride, Prime<Q>::p); }

// ---- the 32x32 warp transpose through shared memory ---------------------------
// writer lane L stores its 8 uint4 vectors in row L at swizzled vector slots;
// reader lane L reads word L of every row j.  Both sides are bank-conflict free.
B2_HD int xchg_wr_word(int lane, int v) { return 32 * lane + 4 * (v ^ (lane & 7)); }
B2_HD int xchg_rd_word(int lane, int j) { return 32 * j + 4 * ((lane >> 2) ^ (j & 7)) + (lane & 3); }

// storage slot (in words) of NTT position (lane, k) inside one 1024-word polynomial of
// the NTT-domain bootstrapping key: [v = k/4][lane][e = k%4] so a warp's uint4 load is
// 512 contiguous bytes.
B2_HD int bk_slot(int lane, int k) { return ((k >> 2) * 32 + lane) * 4 + (k & 3); }

}  // namespace b2
