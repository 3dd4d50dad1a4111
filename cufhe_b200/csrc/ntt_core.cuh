// ntt_core.cuh -- the unsigned Shoup butterfly kept for the integer-pipe peak probe (k_int_peak in
// k_ntt_s.cuh: the roofline denominator of DESIGN.md 4 was measured with this instruction mix and
// stays frozen), the bit-reversal helper of the table generators and the index maps of the team schedule.  The transforms themselves are
// in ntt16s.cuh (signed lazy residues).
#pragma once
#include "modarith.cuh"

namespace b2 {

struct alignas(8) Tw { uint32_t w, ws; };     // twiddle and its Shoup companion (unsigned)

inline unsigned brv10(unsigned x)
{
    unsigned r = 0;
    for (int i = 0; i < kNbit; i++) r |= ((x >> i) & 1u) << (kNbit - 1 - i);
    return r;
}

// index maps of the team schedule (64 threads x 16 values): thread T, register k <-> polynomial position
B2_HD int eA(int T, int k) { return T + 64 * k; }                              // layout A
B2_HD int eC(int T, int k) { return 16 * T + k; }                              // layout C (NTT positions)
B2_HD int eD(int T, int k) { return 256 * (T >> 4) + (T & 15) + 16 * k; }      // layout D
// storage slot of NTT position (T, k) = 16 T + k inside one 1024-word key polynomial:
// [v = k/4][T][e = k%4], so a warp's uint4 load is 512 contiguous bytes.
B2_HD int bk16_slot(int T, int k) { return ((k >> 2) * 64 + T) * 4 + (k & 3); }

// (u, v) -> (u + w v, u - w v + 2p) on lazy unsigned residues: IMAD.HI + 2 IMAD + IADD3 + IADD
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
