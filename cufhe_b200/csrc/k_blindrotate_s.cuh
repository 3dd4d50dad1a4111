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
// What is new (ntt16s.cuh has the arithmetic): signed lazy residues (4-instruction butterflies),
// Solinas primes (SOLM selects the passes whose q*p product runs on the ALU pipe), forward
// passes A -> D -> C so that every exchange is 128-bit on at least one side, and
// synchronisation that matches who actually shares data:
//   * D <-> C exchanges stay inside a half-warp            -> __syncwarp
//   * A -> D and D -> A cross the team's two warps         -> 64-thread named barrier (5 per step)
//   * residue hand-over + accumulator update cross teams   -> 128-thread named barrier (2 per step)
// (the first generation used fourteen 128-thread barriers per step).  Per thread and step the
// exchanges cost 96 LDS/STS instead of 384.
#pragma once
#include "k_blindrotate.cuh"
#include "ntt16s.cuh"

namespace b2 {

__constant__ TeamSConst c_s;

constexpr int kBrsGateSmemWords = 2 * kN + 4 * kXsWords + 640 / 2;      // acc + 2 teams x 2 arrays + abar(u16)
__host__ __device__ constexpr int brs_smem_bytes(int gates) { return (kTeamSTableWords + gates * kBrsGateSmemWords) * 4; }

enum : int { SOL_F2 = 1, SOL_F3 = 2, SOL_I1 = 4, SOL_I2 = 8, SOL_I3 = 16, SOL_SC = 32 };
// timing ablations (tools/ablate_brs.cu only; results are WRONG with any of them set): what does each
// component of a CMUX step cost on the real machine?
enum : int { ABL_LUT = 1,      // table look-ups at conflict-free (lane) indices
             ABL_DIG = 2,      // no rotation / decomposition (one accumulator load per coefficient remains)
             ABL_KEY = 4,      // no key loads (register constants)
             ABL_BAR = 8,      // no named barriers / __syncwarp
             ABL_XCH = 16,     // no shared-memory exchanges
             ABL_FMA = 32,     // butterflies and key products without multiplies
             ABL_TWD = 64 };   // twiddles from registers (no table loads)

// smem image of TeamSTables: per-T vectors f3 | i3 | sc are addressed from ONE pointer
// (vector v of pass X at tbT[64 (off_X + v)]) -- they are NOT adjacent in the struct, so the kernel
// uses three immediates relative to &f3[q][0][T]
struct SPtr {
    int32_t* own;           // chunk_of(T) in array 0 of my team; array 1 = + kXsWords
    int32_t* vw;            // vector scatter base 20 (T&15) + 4 (T>>4)   (+ 336 kh)    A->D, D->A
    int32_t* sw;            // word scatter base 336 (T>>4) + (T&15)      (+ 20 k)      D<->C
    const V4* f2;           // &tab.f2[q][T>>4][0]
    const V4* f3;           // &tab.f3[q][0][T]   (vector stride 64)
    const V4* i2;           // &tab.i2[q][0][T&15] (vector stride 16)
    const V4* i3;           // &tab.i3[q][0][T]
    const V4* sc;           // &tab.sc[q][0][T]
    const int32_t* lut;     // &tab.lut1[q][0]; lut24 / lut35 of prime q follow via tab offsets
    const I2* lut24;
    const I2* lut35;
};

__device__ __forceinline__ void stage_tables_s(uint32_t* s_tab, const TeamSTables* g_tab)
{
    const uint4* src = reinterpret_cast<const uint4*>(g_tab);
    uint4* dst = reinterpret_cast<uint4*>(s_tab);
    for (int i = threadIdx.x; i < kTeamSTableWords / 4; i += blockDim.x) dst[i] = __ldg(src + i);
}

__device__ __forceinline__ void ld_own(int32_t (&x)[16], const int32_t* own)
{
#pragma unroll
    for (int v = 0; v < 4; v++) {
        const V4 r = *reinterpret_cast<const V4*>(own + 4 * v);
        x[4 * v] = r.x; x[4 * v + 1] = r.y; x[4 * v + 2] = r.z; x[4 * v + 3] = r.w;
    }
}

template <int ABL> __device__ __forceinline__ void tbar_sync(int id, int n) { if (!(ABL & ABL_BAR)) named_bar_sync(id, n); }
template <int ABL> __device__ __forceinline__ void warp_sync() { if (!(ABL & ABL_BAR)) __syncwarp(); }
// ablation stand-ins: same data flow, no multiplies
__device__ __forceinline__ void abfly(int32_t& u, int32_t& v, const Tws t)
{
    const int32_t r = v ^ t.w, a = u;
    u = a + r;
    v = a - (r ^ t.ws);
}
template <int N>
__device__ __forceinline__ void load_tw(V4 (&tv)[N], const V4* src, const int stride, const bool fake, const int32_t p, const int T)
{
#pragma unroll
    for (int v = 0; v < N; v++) tv[v] = fake ? V4{p ^ v, T + v, p - v, T ^ v} : src[stride * v];
}

// two digit polynomials (raw digits, layout A) -> NTT domain (layout C)
template <int SOLM, int ABL>
__device__ __forceinline__ void sfwd2_dev(int32_t (&x0)[16], int32_t (&x1)[16], const SPtr& a, const int32_t p, const int sk,
                                          const int tbar, const bool lead, const int T, const uint32_t dbg)
{
    if (ABL & ABL_LUT) {
#pragma unroll
        for (int k = 0; k < 16; k++) {
            x0[k] = (x0[k] & (int32_t)dbg) | (T & 31) | ((k & 7) << 5);
            x1[k] = (x1[k] & (int32_t)dbg) | (T & 31) | (((k + 3) & 7) << 5);
        }
    }
    spass_digits(x0, a.lut, a.lut24, a.lut35);
    spass_digits(x1, a.lut, a.lut24, a.lut35);
    if (lead) tbar_sync<ABL>(tbar, 64);          // owners have read their chunks (previous D->C)
    if (!(ABL & ABL_XCH)) {
#pragma unroll
        for (int kh = 0; kh < 4; kh++) {
            *reinterpret_cast<V4*>(a.vw + 336 * kh) = V4{x0[4 * kh], x0[4 * kh + 1], x0[4 * kh + 2], x0[4 * kh + 3]};
            *reinterpret_cast<V4*>(a.vw + kXsWords + 336 * kh) = V4{x1[4 * kh], x1[4 * kh + 1], x1[4 * kh + 2], x1[4 * kh + 3]};
        }
    }
    tbar_sync<ABL>(tbar, 64);
    if (!(ABL & ABL_XCH)) {
#pragma unroll
        for (int th = 0; th < 4; th++) {
            const V4 r0 = *reinterpret_cast<const V4*>(a.own + 4 * th);
            const V4 r1 = *reinterpret_cast<const V4*>(a.own + kXsWords + 4 * th);
            x0[th] = r0.x; x0[th + 4] = r0.y; x0[th + 8] = r0.z; x0[th + 12] = r0.w;
            x1[th] = r1.x; x1[th + 4] = r1.y; x1[th + 8] = r1.z; x1[th + 12] = r1.w;
        }
    }
    {
        V4 tv[8];
        load_tw(tv, a.f2, 1, (ABL & ABL_TWD) != 0, p, T);
        if (ABL & ABL_FMA) {
#pragma unroll
            for (int s = 0; s < 4; s++)
#pragma unroll
                for (int k = 0; k < 16; k++)
                    if ((k & (8 >> s)) == 0) { const Tws t = tws_of(tv, (1 << s) - 1 + k / (16 >> s)); abfly(x0[k], x0[k + (8 >> s)], t); abfly(x1[k], x1[k + (8 >> s)], t); }
        } else spass_fwd2<(SOLM & SOL_F2) != 0>(x0, x1, tv, p, sk);
    }
    warp_sync<ABL>();                            // group mates have read their chunks
    if (!(ABL & ABL_XCH)) {
#pragma unroll
        for (int k = 0; k < 16; k++) { a.sw[20 * k] = x0[k]; a.sw[kXsWords + 20 * k] = x1[k]; }
    }
    warp_sync<ABL>();
    if (!(ABL & ABL_XCH)) {
        ld_own(x0, a.own);
        ld_own(x1, a.own + kXsWords);
    }
    {
        V4 tv[8];
        load_tw(tv, a.f3, 64, (ABL & ABL_TWD) != 0, p, T);
        if (ABL & ABL_FMA) {
#pragma unroll
            for (int s = 0; s < 4; s++)
#pragma unroll
                for (int k = 0; k < 16; k++)
                    if ((k & (8 >> s)) == 0) { const Tws t = tws_of(tv, (1 << s) - 1 + k / (16 >> s)); abfly(x0[k], x0[k + (8 >> s)], t); abfly(x1[k], x1[k + (8 >> s)], t); }
        } else spass_fwd2<(SOLM & SOL_F3) != 0>(x0, x1, tv, p, sk);
    }
}

// two inverse transforms incl. n^-1 psi^-e: layout C in, layout A out, values in [-p/4, 5p/4)
template <int SOLM, int ABL>
__device__ __forceinline__ void sinv2_dev(int32_t (&x0)[16], int32_t (&x1)[16], const SPtr& a, const int q, const int32_t p,
                                          const int sk, const int tbar, const int T)
{
    if (ABL & ABL_FMA) {
#pragma unroll
        for (int u = 0; u < 4; u++)
#pragma unroll
            for (int k = 0; k < 16; k++)
                if ((k & (1 << u)) == 0) { const Tws t = Tws{p + u, k}; abfly(x0[k], x0[k + (1 << u)], t); abfly(x1[k], x1[k + (1 << u)], t); }
    } else {
        const Tws* c1 = c_s.iv1[q];
        auto tw = [&](int i) { return c1[i]; };
        spass_inv<(SOLM & SOL_I1) != 0, true>(x0, tw, p, sk);
        spass_inv<(SOLM & SOL_I1) != 0, true>(x1, tw, p, sk);
    }
    warp_sync<ABL>();                            // group mates' last loads of their chunks (forward D->C)
    if (!(ABL & ABL_XCH)) {
#pragma unroll
        for (int k = 0; k < 16; k++) { a.sw[20 * k] = x0[k]; a.sw[kXsWords + 20 * k] = x1[k]; }
    }
    warp_sync<ABL>();
    if (!(ABL & ABL_XCH)) {
        ld_own(x0, a.own);
        ld_own(x1, a.own + kXsWords);
    }
    {
        V4 tv[8];
        load_tw(tv, a.i2, 16, (ABL & ABL_TWD) != 0, p, T);
        if (ABL & ABL_FMA) {
#pragma unroll
            for (int u = 0; u < 4; u++)
#pragma unroll
                for (int k = 0; k < 16; k++)
                    if ((k & (1 << u)) == 0) { const Tws t = tws_of(tv, (1 << u) - 1 + (k & ((1 << u) - 1))); abfly(x0[k], x0[k + (1 << u)], t); abfly(x1[k], x1[k + (1 << u)], t); }
        } else {
            auto tw = [&](int i) { return tws_of(tv, i); };
            spass_inv<(SOLM & SOL_I2) != 0, false>(x0, tw, p, sk);
            spass_inv<(SOLM & SOL_I2) != 0, false>(x1, tw, p, sk);
        }
    }
    tbar_sync<ABL>(tbar, 64);                    // every owner has read its chunk
    if (!(ABL & ABL_XCH)) {
#pragma unroll
        for (int kl = 0; kl < 4; kl++) {
            *reinterpret_cast<V4*>(a.vw + 336 * kl) = V4{x0[kl], x0[kl + 4], x0[kl + 8], x0[kl + 12]};
            *reinterpret_cast<V4*>(a.vw + kXsWords + 336 * kl) = V4{x1[kl], x1[kl + 4], x1[kl + 8], x1[kl + 12]};
        }
    }
    tbar_sync<ABL>(tbar, 64);
    if (!(ABL & ABL_XCH)) {
        ld_own(x0, a.own);
        ld_own(x1, a.own + kXsWords);
    }
    {
        V4 tv[6], sv[8];
        load_tw(tv, a.i3, 64, (ABL & ABL_TWD) != 0, p, T);
        load_tw(sv, a.sc, 64, (ABL & ABL_TWD) != 0, p, T);
        if (ABL & ABL_FMA) {
#pragma unroll
            for (int k = 0; k < 16; k++)
                if ((k & 4) == 0) { abfly(x0[k], x0[k + 4], Tws{tv[k & 3].x, tv[k & 3].y}); abfly(x1[k], x1[k + 4], Tws{tv[k & 3].x, tv[k & 3].y}); }
#pragma unroll
            for (int k = 0; k < 8; k++) { abfly(x0[k], x0[k + 8], tws_of(sv, k)); abfly(x1[k], x1[k + 8], tws_of(sv, k)); }
#pragma unroll
            for (int k = 0; k < 16; k++) { x0[k] ^= tws_of(sv, k).w; x1[k] ^= tws_of(sv, k).ws; }
        } else {
            spass_inv3<(SOLM & SOL_I3) != 0, (SOLM & SOL_SC) != 0>(x0, tv, sv, p, sk);
            spass_inv3<(SOLM & SOL_I3) != 0, (SOLM & SOL_SC) != 0>(x1, tv, sv, p, sk);
        }
    }
}

// One code copy serves both primes (q is team-uniform).  sa = residues of output polynomial q
// ("mine", finished by this team), sb = residues of output polynomial 1-q (handed over).
template <int SOLM, bool ROW0, int ABL>
__device__ __forceinline__ void brs_gate_loop(const uint32_t* __restrict__ bk, uint32_t* acc, const SPtr& a,
                                              const int32_t* other_own, const uint16_t* s_abar, const int T, const int q,
                                              const int tbar, const int gbar, const uint32_t dbg)
{
    const int32_t p = (int32_t)(q ? Prime<1>::p : Prime<0>::p);
    const uint32_t pinv = q ? Prime<1>::pinv : Prime<0>::pinv;
    const int sk = q ? 21 : 11;                  // kSolK
    // key row layout [row][prime][out][256 uint4] in bk16_slot order; out = q first, then 1-q
    const uint4* bk4 = reinterpret_cast<const uint4*>(bk) + T + (size_t)q * 2 * (kN / 4);
    const int off_a = q * (kN / 4), off_b = (1 - q) * (kN / 4);
    uint32_t* acc_mine = acc + q * kN + T;
    for (int i = 0; i < kLwe; i++) {
        const uint32_t abar = s_abar[i];
        const int am = (int)(abar & (kN - 1));
        const bool flip = (abar >> kNbit) != 0;
        // coefficient e = T + 64k of X^abar * acc is -/+ acc[e - am (+ 1024)]: wrapped (and negated) iff e < am <=> k < kth
        const int kth = (am - T + 63) >> 6;
        int32_t sa[16], sb[16];
#pragma unroll
        for (int k = 0; k < 16; k++) sa[k] = sb[k] = 0;
#pragma unroll 1
        for (int j = 0; j < 2; j++) {
            int32_t x0[16], x1[16];
            const uint32_t* aj = acc + j * kN + T;
            const uint32_t* lin = aj - am;             // element e - am
            if (ABL & ABL_DIG) {
#pragma unroll
                for (int k = 0; k < 16; k++) {
                    const uint32_t t = (aj[64 * k] & dbg) + (uint32_t)(T + 5 * k);
                    x0[k] = (int32_t)(t & 0xFFu);
                    x1[k] = (int32_t)((t + 7u) & 0xFFu);
                }
            } else {
#pragma unroll
            for (int k = 0; k < 16; k++) {
                const bool wrap = k < kth;
                const uint32_t* src = wrap ? lin + kN : lin;
                uint32_t rot = src[64 * k];
                rot = (wrap != flip) ? 0u - rot : rot;
                const uint32_t t = rot - aj[64 * k] + kDecompAdd;
                x0[k] = (int32_t)(t >> 24);
                x1[k] = (int32_t)((t >> 16) & 0xFFu);
            }
            }
            sfwd2_dev<SOLM, ABL>(x0, x1, a, p, sk, tbar, j != 0, T, dbg);       // j = 0 follows the step's closing barrier
            const uint4* row0 = bk4 + (size_t)((ROW0 ? 0 : i) * kRows + 2 * j) * (4 * (kN / 4));
            const uint4* row1 = row0 + 4 * (kN / 4);
#pragma unroll
            for (int v = 0; v < 4; v++) {
                uint4 b0, b1;
                if (ABL & ABL_KEY) { b0 = make_uint4(i + v, j + 5, T, p - v); b1 = make_uint4(T + v, i, 9 + j, p - 2 * v); }
                else { b0 = __ldg(row0 + off_a + v * 64); b1 = __ldg(row1 + off_a + v * 64); }
                if (ABL & ABL_FMA) {
                    sa[4 * v + 0] += (x0[4 * v + 0] ^ (int32_t)b0.x) + (x1[4 * v + 0] ^ (int32_t)b1.x);
                    sa[4 * v + 1] += (x0[4 * v + 1] ^ (int32_t)b0.y) + (x1[4 * v + 1] ^ (int32_t)b1.y);
                    sa[4 * v + 2] += (x0[4 * v + 2] ^ (int32_t)b0.z) + (x1[4 * v + 2] ^ (int32_t)b1.z);
                    sa[4 * v + 3] += (x0[4 * v + 3] ^ (int32_t)b0.w) + (x1[4 * v + 3] ^ (int32_t)b1.w);
                } else {
                sa[4 * v + 0] += smontmul2(x0[4 * v + 0], b0.x, x1[4 * v + 0], b1.x, p, pinv);
                sa[4 * v + 1] += smontmul2(x0[4 * v + 1], b0.y, x1[4 * v + 1], b1.y, p, pinv);
                sa[4 * v + 2] += smontmul2(x0[4 * v + 2], b0.z, x1[4 * v + 2], b1.z, p, pinv);
                sa[4 * v + 3] += smontmul2(x0[4 * v + 3], b0.w, x1[4 * v + 3], b1.w, p, pinv);
                }
            }
#pragma unroll
            for (int v = 0; v < 4; v++) {
                uint4 b0, b1;
                if (ABL & ABL_KEY) { b0 = make_uint4(i + v, j + 5, T, p - v); b1 = make_uint4(T + v, i, 9 + j, p - 2 * v); }
                else { b0 = __ldg(row0 + off_b + v * 64); b1 = __ldg(row1 + off_b + v * 64); }
                if (ABL & ABL_FMA) {
                    sb[4 * v + 0] += (x0[4 * v + 0] ^ (int32_t)b0.x) + (x1[4 * v + 0] ^ (int32_t)b1.x);
                    sb[4 * v + 1] += (x0[4 * v + 1] ^ (int32_t)b0.y) + (x1[4 * v + 1] ^ (int32_t)b1.y);
                    sb[4 * v + 2] += (x0[4 * v + 2] ^ (int32_t)b0.z) + (x1[4 * v + 2] ^ (int32_t)b1.z);
                    sb[4 * v + 3] += (x0[4 * v + 3] ^ (int32_t)b0.w) + (x1[4 * v + 3] ^ (int32_t)b1.w);
                } else {
                sb[4 * v + 0] += smontmul2(x0[4 * v + 0], b0.x, x1[4 * v + 0], b1.x, p, pinv);
                sb[4 * v + 1] += smontmul2(x0[4 * v + 1], b0.y, x1[4 * v + 1], b1.y, p, pinv);
                sb[4 * v + 2] += smontmul2(x0[4 * v + 2], b0.z, x1[4 * v + 2], b1.z, p, pinv);
                sb[4 * v + 3] += smontmul2(x0[4 * v + 3], b0.w, x1[4 * v + 3], b1.w, p, pinv);
                }
            }
        }
        sinv2_dev<SOLM, ABL>(sb, sa, a, q, p, sk, tbar, T);
        // sa = INTT(out q) (mine), sb = INTT(out 1-q): hand sb to the other team through my own chunk
#pragma unroll
        for (int v = 0; v < 4; v++) *reinterpret_cast<V4*>(a.own + 4 * v) = V4{sb[4 * v], sb[4 * v + 1], sb[4 * v + 2], sb[4 * v + 3]};
        tbar_sync<ABL>(gbar, 128);
        int32_t t[16];
        ld_own(t, other_own);                            // other team's residue of out q
        const int32_t cw = c_s.crt.w, cws = c_s.crt.ws;
        if (q) {
#pragma unroll
            for (int k = 0; k < 16; k++) acc_mine[64 * k] += scrt_lift32(t[k], sa[k], cw, cws);
        } else {
#pragma unroll
            for (int k = 0; k < 16; k++) acc_mine[64 * k] += scrt_lift32(sa[k], t[k], cw, cws);
        }
        tbar_sync<ABL>(gbar, 128);
    }
}

template <int G, int SOLM, bool ROW0 = false, int ABL = 0>
__global__ void __launch_bounds__(128 * G, 1) k_blind_rotate_s(const BrParams P, const TeamSTables* __restrict__ tabs)
{
    extern __shared__ __align__(16) uint32_t smem[];
    stage_tables_s(smem, tabs);
    __syncthreads();

    const int gs = threadIdx.x >> 7, t128 = threadIdx.x & 127;
    const int q = t128 >> 6, T = t128 & 63;
    const int item = P.first + blockIdx.x * G + gs;
    if (item >= P.count) return;

    const TeamSTables* tab = reinterpret_cast<const TeamSTables*>(smem);
    uint32_t* gbase = smem + kTeamSTableWords + gs * kBrsGateSmemWords;
    uint32_t* acc = gbase;
    int32_t* xs = reinterpret_cast<int32_t*>(gbase + 2 * kN) + q * 2 * kXsWords;           // my team's two arrays
    const int32_t* xs_other = reinterpret_cast<const int32_t*>(gbase + 2 * kN) + (q ^ 1) * 2 * kXsWords;
    uint16_t* s_abar = reinterpret_cast<uint16_t*>(gbase + 2 * kN + 4 * kXsWords);
    const int gbar = 1 + gs, tbar = 1 + G + 2 * gs + q;                                    // named barriers (0 = __syncthreads)

    SPtr a;
    a.own = xs + chunk_of(T);
    a.vw = xs + 20 * (T & 15) + 4 * (T >> 4);
    a.sw = xs + 336 * (T >> 4) + (T & 15);
    a.f2 = &tab->f2[q][T >> 4][0];
    a.f3 = &tab->f3[q][0][T];
    a.i2 = &tab->i2[q][0][T & 15];
    a.i3 = &tab->i3[q][0][T];
    a.sc = &tab->sc[q][0][T];
    a.lut = &tab->lut1[q][0];
    a.lut24 = &tab->lut24[q][0];
    a.lut35 = &tab->lut35[q][0];

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
        acc[c] = 0u;
        acc[kN + c] = testvec_b((uint32_t)c, bbar, P.mu);
    }
    named_bar_sync(gbar, 128);

    brs_gate_loop<SOLM, ROW0, ABL>(P.bk, acc, a, xs_other + chunk_of(T), s_abar, T, q, tbar, gbar, P.dbg);

    // ---- SampleExtract(0) (src/bootstrap_gpu.cu:366-381) ----
    if (P.trlwe) {          // BootstrapTLWE2TRLWE (src/bootstrap_gpu.cu:316-322): no extraction
        uint32_t* out = P.out + (size_t)it.slot * 2 * kN;
        for (int m = t128; m < 2 * kN; m += 128) out[m] = acc[m];
        return;
    }
    uint32_t* out = P.out + (size_t)it.slot * kLvl1Words;
    for (int m = t128; m < kN; m += 128) out[m] = m == 0 ? acc[0] : 0u - acc[kN - m];
    if (t128 == 0) out[kN] = acc[kN];
}

}  // namespace b2
