// k_blindrotate_s.cuh -- batched blind rotation + SampleExtract(0), second generation.
//
// Same mathematics and key-stationary batching as k_blindrotate16.cuh (replaces
// __BlindRotatePreAdd__/__BlindRotate__/Accumulate, reference
// include/gatebootstrapping_gpu.cuh:115-345, and __SampleExtractIndex__,
// src/bootstrap_gpu.cu:366-381): a CTA owns G gates and walks the 636 CMUX steps once; a gate
// is four warps = two 64-thread teams, one per NTT prime, 16 coefficients per thread; the two
// gadget digits of an accumulator polynomial are transformed together, their two key rows share
// one Montgomery reduction, the two inverse transforms run together.
//
// What is new (ntt16s.cuh has the arithmetic and the measured cost model: a step costs about
// 4 N(IMAD.HI, IMAD.WIDE) + N(other instructions) issue cycles per warp, so the work is cutting
// instructions):
//   * signed lazy residues: 4-instruction butterflies, no offsets;
//   * forward passes A -> D -> C so that every exchange is 128-bit on at least one side; per thread
//     and step the exchanges cost 96 LDS/STS instead of 384;
//   * synchronisation that matches who actually shares data: D <-> C exchanges stay inside a
//     half-warp (__syncwarp), A -> D and D -> A cross the team's two warps (64-thread named
//     barrier, 5 per step), only the residue hand-over and the accumulator update cross teams
//     (128-thread named barrier, 2 per step) -- the first generation used fourteen 128-thread
//     barriers per step;
//   * the accumulator is kept as [-acc | +acc] (2 x 1024 words per polynomial): coefficient e of
//     X^a acc is word 1024 + e - a, sign included, for every e -- no wrap test, no select, no
//     negation per coefficient; a >= 1024 (all signs flipped) takes a warp-uniform branch;
//   * table look-ups take shared-memory ADDRESSES built straight from the digit byte (SHF/PRMT + LEA);
//   * all shared-memory traffic goes through explicit 32-bit addresses with immediate offsets, so
//     ptxas keeps eight address registers instead of re-deriving pointers inside the loop.
#pragma once
#include "k_blindrotate.cuh"
#include "ntt16s.cuh"

namespace b2 {

__constant__ TeamSConst c_s;

constexpr int kBrsAccWords = 2 * 2 * kN;                               // two polynomials x [-acc | +acc]
constexpr int kBrsGateSmemWords = kBrsAccWords + 4 * kXsWords + 640 / 2;      // acc + 2 teams x 2 arrays + abar(u16)
__host__ __device__ constexpr int brs_smem_bytes(int gates) { return (kTeamSTableWords + gates * kBrsGateSmemWords) * 4; }

// timing ablations (tools/ablate_brs.cu only; results are WRONG with any of them set): what does each
// component of a CMUX step cost on the real machine?
enum : int { ABL_LUT = 1,      // table look-ups at conflict-free (lane) indices
             ABL_DIG = 2,      // no rotation / decomposition (one accumulator load per coefficient remains)
             ABL_KEY = 4,      // no key loads (register constants)
             ABL_BAR = 8,      // no named barriers / __syncwarp
             ABL_XCH = 16,     // no shared-memory exchanges
             ABL_TWD = 64,     // twiddles from registers (no table loads)
             // schedule fuzzing (RIGHT results, b2fhe_set_br_variant(41)): every warp sleeps a random 0..1 us before and
             // after every barrier / __syncwarp, so a missing or misplaced synchronisation shows up as wrong bits
             // (tests/test_gpu_parity.py::test_schedule_jitter_is_bit_exact; compute-sanitizer is closed on this pool)
             ABL_JIT = 32 };

// ---- shared memory through explicit 32-bit window addresses ------------------------------------
#ifdef B2_HOST_TRACE
template <int OFF> inline int32_t lds32(uint32_t a) { return (int32_t)trace_ld(a + OFF, 0); }
template <int OFF> inline I2 lds64(uint32_t a) { I2 r; r.x = (int32_t)trace_ld(a + OFF, 0); r.y = (int32_t)trace_ld(a + OFF, 1); return r; }
template <int OFF> inline V4 lds128(uint32_t a)
{
    V4 r;
    r.x = (int32_t)trace_ld(a + OFF, 0); r.y = (int32_t)trace_ld(a + OFF, 1); r.z = (int32_t)trace_ld(a + OFF, 2); r.w = (int32_t)trace_ld(a + OFF, 3);
    return r;
}
template <int OFF> inline void sts32(uint32_t a, int32_t v) { trace_st(a + OFF, 0, (uint32_t)v); }
template <int OFF> inline void sts128(uint32_t a, int32_t x, int32_t y, int32_t z, int32_t w)
{
    trace_st(a + OFF, 0, (uint32_t)x); trace_st(a + OFF, 1, (uint32_t)y); trace_st(a + OFF, 2, (uint32_t)z); trace_st(a + OFF, 3, (uint32_t)w);
}
inline uint32_t __byte_perm(uint32_t x, uint32_t y, uint32_t s)
{
    const uint64_t v = ((uint64_t)y << 32) | x;
    uint32_t r = 0;
    for (int i = 0; i < 4; i++) r |= (uint32_t)((v >> (8 * ((s >> (4 * i)) & 7))) & 0xFF) << (8 * i);
    return r;
}
#else
template <int OFF> __device__ __forceinline__ int32_t lds32(uint32_t a)
{
    int32_t r;
    asm volatile("ld.shared.s32 %0, [%1 + %2];" : "=r"(r) : "r"(a), "n"(OFF));
    return r;
}
template <int OFF> __device__ __forceinline__ I2 lds64(uint32_t a)
{
    I2 r;
    asm volatile("ld.shared.v2.s32 {%0, %1}, [%2 + %3];" : "=r"(r.x), "=r"(r.y) : "r"(a), "n"(OFF));
    return r;
}
template <int OFF> __device__ __forceinline__ V4 lds128(uint32_t a)
{
    V4 r;
    asm volatile("ld.shared.v4.s32 {%0, %1, %2, %3}, [%4 + %5];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"(a), "n"(OFF));
    return r;
}
template <int OFF> __device__ __forceinline__ void sts32(uint32_t a, int32_t v)
{
    asm volatile("st.shared.s32 [%0 + %1], %2;" ::"r"(a), "n"(OFF), "r"(v) : "memory");
}
template <int OFF> __device__ __forceinline__ void sts128(uint32_t a, int32_t x, int32_t y, int32_t z, int32_t w)
{
    asm volatile("st.shared.v4.s32 [%0 + %1], {%2, %3, %4, %5};" ::"r"(a), "n"(OFF), "r"(x), "r"(y), "r"(z), "r"(w) : "memory");
}

#endif

// key vector load flavours (A/B in tools/ablate_brs.cu): 0 = ld.global.nc (through L1), 1 = ld.global.cg (L2 only),
// 2 = ld.global.cs (streaming), 3 = ld.global.nc.L1::no_allocate
template <int KEYLD> __device__ __forceinline__ uint4 ld_key(const uint4* p)
{
#ifdef B2_HOST_TRACE
    return *p;
#else
    if (KEYLD == 1) return __ldcg(p);
    if (KEYLD == 2) return __ldcs(p);
    if (KEYLD == 3) {
        uint4 r;
        asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
        return r;
    }
    return __ldg(p);
#endif
}

struct SAddr {              // per-thread shared-window byte addresses
    uint32_t own;           // chunk_of(T) in array 0 of my team; array 1 = + 4 kXsWords
    uint32_t vw;            // vector scatter base 20 (T&15) + 4 (T>>4) words  (+ 336 words per kh)   A->D, D->A
    uint32_t sw;            // word scatter base 336 (T>>4) + (T&15) words     (+ 20 words per k)     D<->C
    uint32_t oth;           // the other team's chunk_of(T) in its array 0 (residue hand-over)
    uint32_t tT;            // &tab.pr[q].perT[0][T]   (+ 1024 bytes per vector)
    uint32_t tLo;           // &tab.pr[q].i2[0][T&15]   (+ 256 bytes per vector)
    uint32_t tHi;           // &tab.pr[q].f2[T>>4][0]   (+ 16 bytes per vector)
    uint32_t lut;           // &tab.pr[q].lut0[0]   (4-byte entries; lut12[0], lut12[1] at + kLut1Off, + kLut2Off, 8-byte entries)
};
constexpr int kXsBytes = kXsWords * 4;

__device__ __forceinline__ void stage_tables_s(uint32_t* s_tab, const TeamSTables* g_tab)
{
    const uint4* src = reinterpret_cast<const uint4*>(g_tab);
    uint4* dst = reinterpret_cast<uint4*>(s_tab);
    for (int i = threadIdx.x; i < kTeamSTableWords / 4; i += blockDim.x) dst[i] = __ldg(src + i);
}

#ifndef B2_HOST_TRACE
// warp-uniform random sleep of 0..1023 ns (clock-seeded: different in every launch).  It moves WARPS against each
// other: tools/fuzz_mutants.cu shows that dropping any of the six cross-warp barriers of a step is caught (all gates
// wrong under fuzzing; the 'lead' barrier alone corrupts only ~5 of 1184 gates without it).  The four __syncwarp()
// of the half-warp exchanges order LANES of one warp; a converged warp issues its shared-memory instructions in
// program order, so no timing perturbation can make their absence visible (per-lane sleeps were tried: a single
// NANOSLEEP does not split a warp) -- they are kept because the PTX memory model requires them.
__device__ __forceinline__ void jitter_sleep(uint32_t salt)
{
    uint32_t h = ((uint32_t)clock() ^ salt) * 0x9E3779B9u;
    h ^= h >> 15; h *= 0x2C1B3C6Du; h ^= h >> 13;
    h = __shfl_sync(0xFFFFFFFFu, h, 0);
    __nanosleep(h & 1023u);
}
#else
inline void jitter_sleep(uint32_t) {}
inline void __syncwarp() { trace_warpsync(); }
#endif
// SITE numbers the ten synchronisation points of a step; a MUTANT build (tools/fuzz_mutants.cu) drops one of them
// to show that the fuzzing notices: ABL = ABL_JIT | site << kMutShift
constexpr int kMutShift = 8;
#ifdef B2_HOST_TRACE
template <int ABL, int SITE> inline bool sync_kept() { return trace_sync_kept(SITE); }
#else
template <int ABL, int SITE> __host__ __device__ constexpr bool sync_kept() { return !(ABL & ABL_BAR) && ((ABL >> kMutShift) & 15) != SITE; }
#endif
template <int ABL, int SITE = 0> __device__ __forceinline__ void tbar_sync(int id, int n)
{
    if (ABL & ABL_JIT) jitter_sleep(threadIdx.x >> 5);
    if (sync_kept<ABL, SITE>()) named_bar_sync(id, n);
    if (ABL & ABL_JIT) jitter_sleep(0x55u + (threadIdx.x >> 5));
}
template <int ABL, int SITE = 0> __device__ __forceinline__ void warp_sync()
{
    if (ABL & ABL_JIT) jitter_sleep(0xAAu + (threadIdx.x >> 5));
    if (sync_kept<ABL, SITE>()) __syncwarp();
    if (ABL & ABL_JIT) jitter_sleep(0x33u + (threadIdx.x >> 5));
}

// owner's four LDS.128 of its chunk in array ARR: vector v -> x[4 v .. 4 v + 3]
template <int ARR> __device__ __forceinline__ void ld_own(int32_t (&x)[16], const uint32_t own)
{
    const V4 r0 = lds128<ARR * kXsBytes + 0>(own), r1 = lds128<ARR * kXsBytes + 16>(own);
    const V4 r2 = lds128<ARR * kXsBytes + 32>(own), r3 = lds128<ARR * kXsBytes + 48>(own);
    x[0] = r0.x; x[1] = r0.y; x[2] = r0.z; x[3] = r0.w;
    x[4] = r1.x; x[5] = r1.y; x[6] = r1.z; x[7] = r1.w;
    x[8] = r2.x; x[9] = r2.y; x[10] = r2.z; x[11] = r2.w;
    x[12] = r3.x; x[13] = r3.y; x[14] = r3.z; x[15] = r3.w;
}
// ... the same after A -> D: vector th -> x[th], x[th + 4], x[th + 8], x[th + 12]
template <int ARR> __device__ __forceinline__ void ld_own_strided(int32_t (&x)[16], const uint32_t own)
{
    const V4 r0 = lds128<ARR * kXsBytes + 0>(own), r1 = lds128<ARR * kXsBytes + 16>(own);
    const V4 r2 = lds128<ARR * kXsBytes + 32>(own), r3 = lds128<ARR * kXsBytes + 48>(own);
    x[0] = r0.x; x[4] = r0.y; x[8] = r0.z; x[12] = r0.w;
    x[1] = r1.x; x[5] = r1.y; x[9] = r1.z; x[13] = r1.w;
    x[2] = r2.x; x[6] = r2.y; x[10] = r2.z; x[14] = r2.w;
    x[3] = r3.x; x[7] = r3.y; x[11] = r3.z; x[15] = r3.w;
}
// word scatter (D <-> C): word k to sw + 80 k bytes
template <int ARR, int K = 0> __device__ __forceinline__ void st_words(const int32_t (&x)[16], const uint32_t sw)
{
    if constexpr (K < 16) {
        sts32<ARR * kXsBytes + 80 * K>(sw, x[K]);
        st_words<ARR, K + 1>(x, sw);
    }
}
// vector scatter, contiguous registers (A -> D): vector kh = x[4 kh .. 4 kh + 3] to vw + 1344 kh bytes
template <int ARR> __device__ __forceinline__ void st_vec_contig(const int32_t (&x)[16], const uint32_t vw)
{
    sts128<ARR * kXsBytes + 0>(vw, x[0], x[1], x[2], x[3]);
    sts128<ARR * kXsBytes + 1344>(vw, x[4], x[5], x[6], x[7]);
    sts128<ARR * kXsBytes + 2688>(vw, x[8], x[9], x[10], x[11]);
    sts128<ARR * kXsBytes + 4032>(vw, x[12], x[13], x[14], x[15]);
}
// vector scatter, strided registers (D -> A): vector kl = x[kl], x[kl + 4], x[kl + 8], x[kl + 12]
template <int ARR> __device__ __forceinline__ void st_vec_strided(const int32_t (&x)[16], const uint32_t vw)
{
    sts128<ARR * kXsBytes + 0>(vw, x[0], x[4], x[8], x[12]);
    sts128<ARR * kXsBytes + 1344>(vw, x[1], x[5], x[9], x[13]);
    sts128<ARR * kXsBytes + 2688>(vw, x[2], x[6], x[10], x[14]);
    sts128<ARR * kXsBytes + 4032>(vw, x[3], x[7], x[11], x[15]);
}
// twiddle vector loader (vector stride STRIDE bytes from BASE), or register stand-ins for the TWD ablation
template <int STRIDE, int BASE, bool FAKE>
struct TwLd {
    uint32_t a; int32_t p; int T;
    __device__ __forceinline__ V4 operator()(int v) const
    {
        if (FAKE) return V4{p ^ v, T + v, p - v, T ^ v};
        // v is a compile-time constant after unrolling: one LDS.128 with an immediate offset
        switch (v) {
            case 0: return lds128<BASE + 0 * STRIDE>(a);
            case 1: return lds128<BASE + 1 * STRIDE>(a);
            case 2: return lds128<BASE + 2 * STRIDE>(a);
            case 3: return lds128<BASE + 3 * STRIDE>(a);
            case 4: return lds128<BASE + 4 * STRIDE>(a);
            case 5: return lds128<BASE + 5 * STRIDE>(a);
            case 6: return lds128<BASE + 6 * STRIDE>(a);
            default: return lds128<BASE + 7 * STRIDE>(a);
        }
    }
};

// two digit polynomials (slots k < 4: signed digits, k >= 4: table addresses; layout A) -> NTT domain (layout C)
template <int ABL>
__device__ __forceinline__ void sfwd2_dev(int32_t (&x0)[16], int32_t (&x1)[16], const SAddr& a, const int32_t p, const int tbar,
                                          const bool lead, const int T)
{
    {
        auto l0 = [](int32_t h) { return lds32<0>((uint32_t)h); };
        auto l1 = [](int32_t h) { return lds64<kLut1Off>((uint32_t)h); };
        auto l2 = [](int32_t h) { return lds64<kLut2Off>((uint32_t)h); };
        spass_digits(x0, l0, l1, l2);
        spass_digits(x1, l0, l1, l2);
    }
    if (lead) tbar_sync<ABL, 1>(tbar, 64);          // owners have read their chunks (previous D->C)
    if (!(ABL & ABL_XCH)) { st_vec_contig<0>(x0, a.vw); st_vec_contig<1>(x1, a.vw); }
    tbar_sync<ABL, 2>(tbar, 64);
    if (!(ABL & ABL_XCH)) { ld_own_strided<0>(x0, a.own); ld_own_strided<1>(x1, a.own); }
    spass_fwd<2>(x0, x1, TwLd<16, 0, (ABL & ABL_TWD) != 0>{a.tHi, p, T}, p);
    warp_sync<ABL, 3>();                            // group mates have read their chunks
    if (!(ABL & ABL_XCH)) { st_words<0>(x0, a.sw); st_words<1>(x1, a.sw); }
    warp_sync<ABL, 4>();
    if (!(ABL & ABL_XCH)) { ld_own<0>(x0, a.own); ld_own<1>(x1, a.own); }
    spass_fwd<2>(x0, x1, TwLd<1024, 1024 * kVecF3, (ABL & ABL_TWD) != 0>{a.tT, p, T}, p);
}

// two inverse transforms incl. n^-1 psi^-e: layout C in, layout A out, values in [-p/4, 5p/4)
template <int ABL>
__device__ __forceinline__ void sinv2_dev(int32_t (&x0)[16], int32_t (&x1)[16], const SAddr& a, const int q, const int32_t p,
                                          const int tbar, const int T)
{
    {
        const Tws* c1 = c_s.iv1[q];
        spass_inv1<2>(x0, x1, [&](int i) { return c1[i]; }, p);
    }
    warp_sync<ABL, 5>();                            // group mates' last loads of their chunks (forward D->C)
    if (!(ABL & ABL_XCH)) { st_words<0>(x0, a.sw); st_words<1>(x1, a.sw); }
    warp_sync<ABL, 6>();
    if (!(ABL & ABL_XCH)) { ld_own<0>(x0, a.own); ld_own<1>(x1, a.own); }
    spass_inv2<2>(x0, x1, TwLd<256, 0, (ABL & ABL_TWD) != 0>{a.tLo, p, T}, p);
    tbar_sync<ABL, 7>(tbar, 64);                    // every owner has read its chunk
    if (!(ABL & ABL_XCH)) { st_vec_strided<0>(x0, a.vw); st_vec_strided<1>(x1, a.vw); }
    tbar_sync<ABL, 8>(tbar, 64);
    if (!(ABL & ABL_XCH)) { ld_own<0>(x0, a.own); ld_own<1>(x1, a.own); }
    spass_inv3<2>(x0, x1, TwLd<1024, 1024 * kVecI3, (ABL & ABL_TWD) != 0>{a.tT, p, T},
                  TwLd<1024, 1024 * kVecSc, (ABL & ABL_TWD) != 0>{a.tT, p, T}, p);
}

// digit extraction of accumulator polynomial j: slots k < 4 get the signed digits, slots k >= 4 the
// address of their table entry.  own = &D_j[1024 + T] (my coefficients), rot = &D_j[1024 + T - am]
// (coefficient e - am of the rotated polynomial, sign included), both + 256 bytes per k.
template <bool FLIP, int K = 0>
__device__ __forceinline__ void digits_dev(int32_t (&x0)[16], int32_t (&x1)[16], const uint32_t own, const uint32_t rot, const uint32_t lut)
{
    if constexpr (K < 16) {
        const uint32_t r = (uint32_t)lds32<256 * K>(rot), o = (uint32_t)lds32<256 * K>(own);
        const uint32_t t = FLIP ? kDecompAdd - o - r : kDecompAdd - o + r;
        if (K < 4) {
            x0[K] = (int32_t)(t >> 24) - 128;
            x1[K] = (int32_t)((t >> 16) & 0xFFu) - 128;
        } else if (K >= 8 && K < 12) {          // table 0: 4-byte entries
            x0[K] = (int32_t)(lut + ((t >> 24) << 2));
            x1[K] = (int32_t)(lut + (__byte_perm(t, 0, 0x4442) << 2));
        } else {
            x0[K] = (int32_t)(lut + ((t >> 24) << 3));
            x1[K] = (int32_t)(lut + (__byte_perm(t, 0, 0x4442) << 3));
        }
        digits_dev<FLIP, K + 1>(x0, x1, own, rot, lut);
    }
}
// accumulator update of my sixteen coefficients: acc += S, both copies
template <int K = 0>
__device__ __forceinline__ void acc_update(const uint32_t mine, const int32_t (&s0)[16], const int32_t (&s1)[16], const int32_t cw, const int32_t cws)
{
    if constexpr (K < 16) {
        const uint32_t nv = (uint32_t)lds32<256 * K>(mine) + scrt_lift32(s0[K], s1[K], cw, cws);
        sts32<256 * K>(mine, (int32_t)nv);
        sts32<256 * K - 4096>(mine, (int32_t)(0u - nv));
        acc_update<K + 1>(mine, s0, s1, cw, cws);
    }
}

// One code copy serves both primes (q is team-uniform).  sa = residues of output polynomial q
// ("mine", finished by this team), sb = residues of output polynomial 1-q (handed over).
template <bool ROW0, int ABL, int KEYLD>
__device__ __forceinline__ void brs_gate_loop(const uint32_t* __restrict__ bk, const uint32_t accb, const SAddr& a,
                                              const uint16_t* s_abar, const int T, const int q, const int tbar, const int gbar,
                                              const uint32_t dbg, const int nsteps = kLwe)
{
    const int32_t p = (int32_t)(q ? Prime<1>::p : Prime<0>::p);
    const uint32_t pinv = q ? Prime<1>::pinv : Prime<0>::pinv;
    // key row layout [row][prime][out][256 uint4] in bk16_slot order; out = q first, then 1-q
    const uint4* bk4 = reinterpret_cast<const uint4*>(bk) + T + (size_t)q * 2 * (kN / 4);
    const int off_a = q * (kN / 4), off_b = (1 - q) * (kN / 4);
    const uint32_t own0 = accb + 4 * (kN + T);                     // &D_0[1024 + T]; D_1 = + 8192 bytes
    const uint32_t mine = own0 + q * 8192;                         // my output polynomial's +acc copy
    const uint4* row0 = bk4;                                       // rows 2j, 2j + 1 of step i: advances by two rows per j
    for (int i = 0; i < nsteps; i++) {
        const uint32_t abar = s_abar[i];
        const uint32_t rot0 = own0 - 4 * (abar & (kN - 1));
        const bool flip = (abar >> kNbit) != 0;
        // the four digit polynomials of the step: (y0, y1) = digits of acc_0, (x0, x1) = digits of acc_1
        int32_t y0[16], y1[16], x0[16], x1[16];
        if (ABL & ABL_DIG) {
#pragma unroll
            for (int k = 0; k < 16; k++) {
                const uint32_t t = ((uint32_t)lds32<0>(own0 + 256 * k) & dbg) + (uint32_t)(T + 5 * k);
                y0[k] = k < 4 ? (int32_t)(t & 0xFFu) - 128 : (int32_t)(a.lut + ((t & 0xFFu) << ((k >> 2) == 2 ? 2 : 3)));
                y1[k] = k < 4 ? (int32_t)((t + 7u) & 0xFFu) - 128 : (int32_t)(a.lut + (((t + 7u) & 0xFFu) << ((k >> 2) == 2 ? 2 : 3)));
            }
        } else if (flip) digits_dev<true>(y0, y1, own0, rot0, a.lut);
        else digits_dev<false>(y0, y1, own0, rot0, a.lut);
        if (ABL & ABL_LUT) {
#pragma unroll
            for (int k = 4; k < 16; k++) {
                y0[k] = (int32_t)((((uint32_t)y0[k] - a.lut) & dbg) + a.lut + 8 * ((T & 31) | ((k & 7) << 5)));
                y1[k] = (int32_t)((((uint32_t)y1[k] - a.lut) & dbg) + a.lut + 8 * ((T & 31) | (((k + 3) & 7) << 5)));
            }
        }
        sfwd2_dev<ABL>(y0, y1, a, p, tbar, false, T);        // follows the step's closing barrier
        if (ABL & ABL_DIG) {
#pragma unroll
            for (int k = 0; k < 16; k++) {
                const uint32_t t = ((uint32_t)lds32<8192>(own0 + 256 * k) & dbg) + (uint32_t)(T + 3 * k);
                x0[k] = k < 4 ? (int32_t)(t & 0xFFu) - 128 : (int32_t)(a.lut + ((t & 0xFFu) << ((k >> 2) == 2 ? 2 : 3)));
                x1[k] = k < 4 ? (int32_t)((t + 7u) & 0xFFu) - 128 : (int32_t)(a.lut + (((t + 7u) & 0xFFu) << ((k >> 2) == 2 ? 2 : 3)));
            }
        } else if (flip) digits_dev<true>(x0, x1, own0 + 8192, rot0 + 8192, a.lut);
        else digits_dev<false>(x0, x1, own0 + 8192, rot0 + 8192, a.lut);
        if (ABL & ABL_LUT) {
#pragma unroll
            for (int k = 4; k < 16; k++) {
                x0[k] = (int32_t)((((uint32_t)x0[k] - a.lut) & dbg) + a.lut + 8 * ((T & 31) | ((k & 7) << 5)));
                x1[k] = (int32_t)((((uint32_t)x1[k] - a.lut) & dbg) + a.lut + 8 * ((T & 31) | (((k + 3) & 7) << 5)));
            }
        }
        sfwd2_dev<ABL>(x0, x1, a, p, tbar, true, T);
        // key multiply: the four row products of an output coefficient are summed in 64 bits (< 2^33 p)
        // and reduced ONCE; out 1-q first (its inverse transform input is sb), then out q
        int32_t sa[16], sb[16];
        const uint4* r1 = row0 + 4 * (kN / 4);
        const uint4* r2 = row0 + 8 * (kN / 4);
        const uint4* r3 = row0 + 12 * (kN / 4);
#pragma unroll
        for (int v = 0; v < 4; v++) {
            uint4 b0, b1, b2, b3;
            if (ABL & ABL_KEY) { b0 = make_uint4(i + v, 5, T, p - v); b1 = make_uint4(T + v, i, 9, p - 2 * v); b2 = make_uint4(i, T, v, p - 5); b3 = make_uint4(v, v + i, T, p - 6); }
            else { b0 = ld_key<KEYLD>(row0 + off_b + v * 64); b1 = ld_key<KEYLD>(r1 + off_b + v * 64); b2 = ld_key<KEYLD>(r2 + off_b + v * 64); b3 = ld_key<KEYLD>(r3 + off_b + v * 64); }
            sb[4 * v + 0] = smontmul4(y0[4 * v + 0], b0.x, y1[4 * v + 0], b1.x, x0[4 * v + 0], b2.x, x1[4 * v + 0], b3.x, p, pinv);
            sb[4 * v + 1] = smontmul4(y0[4 * v + 1], b0.y, y1[4 * v + 1], b1.y, x0[4 * v + 1], b2.y, x1[4 * v + 1], b3.y, p, pinv);
            sb[4 * v + 2] = smontmul4(y0[4 * v + 2], b0.z, y1[4 * v + 2], b1.z, x0[4 * v + 2], b2.z, x1[4 * v + 2], b3.z, p, pinv);
            sb[4 * v + 3] = smontmul4(y0[4 * v + 3], b0.w, y1[4 * v + 3], b1.w, x0[4 * v + 3], b2.w, x1[4 * v + 3], b3.w, p, pinv);
        }
#pragma unroll
        for (int v = 0; v < 4; v++) {
            uint4 b0, b1, b2, b3;
            if (ABL & ABL_KEY) { b0 = make_uint4(i + v, 5, T, p - v); b1 = make_uint4(T + v, i, 9, p - 2 * v); b2 = make_uint4(i, T, v, p - 5); b3 = make_uint4(v, v + i, T, p - 6); }
            else { b0 = ld_key<KEYLD>(row0 + off_a + v * 64); b1 = ld_key<KEYLD>(r1 + off_a + v * 64); b2 = ld_key<KEYLD>(r2 + off_a + v * 64); b3 = ld_key<KEYLD>(r3 + off_a + v * 64); }
            sa[4 * v + 0] = smontmul4(y0[4 * v + 0], b0.x, y1[4 * v + 0], b1.x, x0[4 * v + 0], b2.x, x1[4 * v + 0], b3.x, p, pinv);
            sa[4 * v + 1] = smontmul4(y0[4 * v + 1], b0.y, y1[4 * v + 1], b1.y, x0[4 * v + 1], b2.y, x1[4 * v + 1], b3.y, p, pinv);
            sa[4 * v + 2] = smontmul4(y0[4 * v + 2], b0.z, y1[4 * v + 2], b1.z, x0[4 * v + 2], b2.z, x1[4 * v + 2], b3.z, p, pinv);
            sa[4 * v + 3] = smontmul4(y0[4 * v + 3], b0.w, y1[4 * v + 3], b1.w, x0[4 * v + 3], b2.w, x1[4 * v + 3], b3.w, p, pinv);
        }
        if (!ROW0) row0 += 16 * (kN / 4);                // timing experiment ROW0: key row 0 at every step
        sinv2_dev<ABL>(sb, sa, a, q, p, tbar, T);
        // sa = INTT(out q) (mine), sb = INTT(out 1-q): hand sb to the other team through my own chunk
        sts128<0>(a.own, sb[0], sb[1], sb[2], sb[3]);
        sts128<16>(a.own, sb[4], sb[5], sb[6], sb[7]);
        sts128<32>(a.own, sb[8], sb[9], sb[10], sb[11]);
        sts128<48>(a.own, sb[12], sb[13], sb[14], sb[15]);
        tbar_sync<ABL, 9>(gbar, 128);
        int32_t t[16];
        ld_own<0>(t, a.oth);                             // other team's residue of out q
        const int32_t cw = c_s.crt.w, cws = c_s.crt.ws;
        if (q) acc_update(mine, t, sa, cw, cws);
        else acc_update(mine, sa, t, cw, cws);
        tbar_sync<ABL, 10>(gbar, 128);
    }
}

// the shared-window addresses of thread T of team q (xs / xs_other: the two exchange arrays of my / the other team)
__device__ __forceinline__ SAddr brs_addr(const TeamSTables* tab, const uint32_t* xs, const uint32_t* xs_other, const int q, const int T)
{
    SAddr a;
    a.own = smem_u32(xs + chunk_of(T));
    a.vw = smem_u32(xs + 20 * (T & 15) + 4 * (T >> 4));
    a.sw = smem_u32(xs + 336 * (T >> 4) + (T & 15));
    a.oth = smem_u32(xs_other + chunk_of(T));
    a.tT = smem_u32(&tab->pr[q].perT[0][T]);
    a.tLo = smem_u32(&tab->pr[q].i2[0][T & 15]);
    a.tHi = smem_u32(&tab->pr[q].f2[T >> 4][0]);
    a.lut = smem_u32(&tab->pr[q].lut0[0]);
    return a;
}

#ifndef B2_BRS_STEPS
#define B2_BRS_STEPS kLwe          // 636; the host trace build (tests/emul/emul_brs.cpp) walks a few steps only
#endif
template <int G, bool ROW0 = false, int ABL = 0, int KEYLD = 0>
__global__ void __launch_bounds__(128 * G, 1) k_blind_rotate_s(const BrParams P, const TeamSTables* __restrict__ tabs)
{
    B2_DYN_SMEM(smem);
    stage_tables_s(smem, tabs);
    __syncthreads();

    const int gs = threadIdx.x >> 7, t128 = threadIdx.x & 127;
    const int q = t128 >> 6, T = t128 & 63;
    const int item = P.first + blockIdx.x * G + gs;
    if (item >= P.count) return;

    const TeamSTables* tab = reinterpret_cast<const TeamSTables*>(smem);
    uint32_t* gbase = smem + kTeamSTableWords + gs * kBrsGateSmemWords;
    uint32_t* acc = gbase;                                   // D_0 = [-acc_0 | +acc_0], D_1 = [-acc_1 | +acc_1]
    uint32_t* xs = gbase + kBrsAccWords + q * 2 * kXsWords;  // my team's two arrays
    uint32_t* xs_other = gbase + kBrsAccWords + (q ^ 1) * 2 * kXsWords;
    uint16_t* s_abar = reinterpret_cast<uint16_t*>(gbase + kBrsAccWords + 4 * kXsWords);
    const int gbar = 1 + gs, tbar = 1 + G + 2 * gs + q;      // named barriers (0 = __syncthreads)

    const SAddr a = brs_addr(tab, xs, xs_other, q, T);

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
        const uint32_t tv = testvec_b((uint32_t)c, bbar, P.mu);
        B2_SMEM_WR(acc + c, 0u);
        B2_SMEM_WR(acc + kN + c, 0u);
        B2_SMEM_WR(acc + 2 * kN + c, 0u - tv);
        B2_SMEM_WR(acc + 3 * kN + c, tv);
    }
    named_bar_sync(gbar, 128);

    brs_gate_loop<ROW0, ABL, KEYLD>(P.bk, smem_u32(acc), a, s_abar, T, q, tbar, gbar, P.dbg, B2_BRS_STEPS);

    // ---- SampleExtract(0) (src/bootstrap_gpu.cu:366-381) ----
    const uint32_t* acc0 = acc + kN;                         // +acc_0
    const uint32_t* acc1 = acc + 3 * kN;                     // +acc_1
    if (P.trlwe) {          // BootstrapTLWE2TRLWE (src/bootstrap_gpu.cu:316-322): no extraction
        uint32_t* out = P.out + (size_t)it.slot * 2 * kN;
        for (int m = t128; m < kN; m += 128) { out[m] = B2_SMEM_RD(acc0 + m); out[kN + m] = B2_SMEM_RD(acc1 + m); }
        return;
    }
    uint32_t* out = P.out + (size_t)it.slot * kLvl1Words;
    for (int m = t128; m < kN; m += 128) out[m] = m == 0 ? B2_SMEM_RD(acc0) : 0u - B2_SMEM_RD(acc0 + kN - m);
    if (t128 == 0) out[kN] = B2_SMEM_RD(acc1);
}


}  // namespace b2
