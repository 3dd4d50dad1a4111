// modarith.cuh -- 32-bit signed Montgomery arithmetic over two 27-bit NTT primes.
//
// Replaces the reference's 64-bit Barrett field (include/ntt_gpu/ntt_gpuntt.cuh:51-224,
// FFP / barrett_mult over p ~ 2^60).  SURVEY.md F2: the external product is an exact
// integer negacyclic convolution with |S| < 2^51, so any modulus P > 2^52 gives the
// same ciphertext bits.  We use P = p0 * p1 (~2^54) with two primes < 2^27, each
// = 1 (mod 2048), and recombine with a centred CRT.  Every modular multiply is then
// 3 integer-pipe instructions (IMAD.WIDE, IMAD, IMAD.HI) on sm_100a instead of a
// 64x64->128 Barrett product, and lazy signed residues need no conditional corrections.
//
// All functions are host+device so the same code is checked on the CPU
// (tests/emul) before it runs on a GPU.
#pragma once
#include <stdint.h>

#ifdef __CUDACC__
#define B2_HD __host__ __device__ __forceinline__
#else
#define B2_HD inline
#endif

namespace b2 {

constexpr int kN = 1024;        // lvl1 ring degree (TFHEpp params/128bit.hpp:42-72)
constexpr int kNbit = 10;
constexpr int kLwe = 636;       // lvl0 dimension (128bit.hpp:7-23)
constexpr int kL = 2;           // gadget levels
constexpr int kRows = 4;        // (k+1)*l TRGSW rows
constexpr int kIksT = 7;        // key-switch digits
constexpr int kLvl0Words = kLwe + 1;
constexpr int kLvl1Words = kN + 1;
constexpr uint32_t kMu0 = 1u << 13;
constexpr uint32_t kMu1 = 1u << 29;

constexpr uint32_t inv32(uint32_t p)
{   // p^-1 mod 2^32 by Newton iteration (p odd)
    uint32_t x = p;
    for (int i = 0; i < 5; i++) x *= 2u - p * x;
    return x;
}

template <int Q> struct Prime;
template <> struct Prime<0> {
    static constexpr uint32_t p = 134215681u;           // 65535 * 2048 + 1 = 2^27 - 2^11 + 1
    static constexpr uint32_t pinv = inv32(134215681u);
};
template <> struct Prime<1> {
    static constexpr uint32_t p = 132120577u;           // 63 * 2^21 + 1 = 2^27 - 2^21 + 1
    static constexpr uint32_t pinv = inv32(132120577u);
};
// third prime: ONLY for the exact polynomial product of the microbenchmark entry point (any u32 x int32
// inputs, |sum| < 2^73 < P/2 ~ 2^80); the bootstrapping path needs two (|sum| < 2^51 < P/2 ~ 2^53)
template <> struct Prime<2> {
    static constexpr uint32_t p = 134250497u;           // 2^27 + 2^15 + 1
    static constexpr uint32_t pinv = inv32(134250497u);
};
constexpr uint32_t kP[3] = {Prime<0>::p, Prime<1>::p, Prime<2>::p};
constexpr uint32_t kPinv[3] = {Prime<0>::pinv, Prime<1>::pinv, Prime<2>::pinv};

// Shoup multiplication by a constant w < p with companion ws = floor(w * 2^32 / p):
// v * w mod p in [0, 2p) for ANY 32-bit v.  One IMAD.HI + two IMAD on sm_100a, which
// measures 1.7x faster than the Montgomery form (profiles/r1_ubench_butterfly_mix.log)
// because IMAD.WIDE and IMAD.HI are both half-rate.
B2_HD uint32_t shoup_mul_p(uint32_t v, uint32_t w, uint32_t ws, uint32_t p)
{
#ifdef __CUDA_ARCH__
    const uint32_t q = __umulhi(v, ws);
#else
    const uint32_t q = (uint32_t)(((uint64_t)v * (uint64_t)ws) >> 32);
#endif
    return v * w - q * p;
}

// ---- host-only helpers for table generation --------------------------------
inline uint32_t h_mulmod(uint32_t a, uint32_t b, uint32_t p) { return (uint32_t)((uint64_t)a * b % p); }
inline uint32_t h_powmod(uint32_t b, uint64_t e, uint32_t p)
{
    uint32_t r = 1;
    while (e) { if (e & 1) r = h_mulmod(r, b, p); b = h_mulmod(b, b, p); e >>= 1; }
    return r;
}
inline uint32_t h_shoup(uint32_t w, uint32_t p) { return (uint32_t)((((uint64_t)w) << 32) / p); }
// primitive 2N-th root of unity (smallest generator candidate g with (g^((p-1)/2N))^N == -1)
inline uint32_t h_psi(uint32_t p)
{
    for (uint32_t g = 2;; g++) {
        uint32_t psi = h_powmod(g, (p - 1) / (2 * kN), p);
        if (h_powmod(psi, kN, p) == p - 1) return psi;
    }
}

}  // namespace b2
