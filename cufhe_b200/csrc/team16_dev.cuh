// team16_dev.cuh -- device side of the first-generation 64-thread x 16-coefficient team transforms
// (ntt16.cuh: unsigned lazy residues, padded scalar exchanges, 128-thread named barriers).  Still
// used by the latency kernels (k_blindrotate_cl4.cuh, k_blindrotate_lat.cuh) and the stand-alone
// CMUX (k_cmux.cuh); the throughput path moved to ntt16s.cuh / k_blindrotate_s.cuh.
#pragma once
#include "k_blindrotate.cuh"
#include "ntt16.cuh"

namespace b2 {

// shared-memory table accessors (Lane16Tables staged verbatim)
__device__ __forceinline__ const Tw* t16_fw2(const uint32_t* s, int q, int B) { return reinterpret_cast<const Tw*>(s) + q * 15 * 16 + B; }
__device__ __forceinline__ const Tw* t16_fw3(const uint32_t* s, int q, int T) { return reinterpret_cast<const Tw*>(s) + 2 * 15 * 16 + q * 12 * 64 + T; }
__device__ __forceinline__ const Tw* t16_iv2(const uint32_t* s, int q, int lo) { return reinterpret_cast<const Tw*>(s) + 2 * 15 * 16 + 2 * 12 * 64 + q * 15 * 16 + lo; }
__device__ __forceinline__ const Tw* t16_iv3(const uint32_t* s, int q, int T) { return reinterpret_cast<const Tw*>(s) + 4 * 15 * 16 + 2 * 12 * 64 + q * 12 * 64 + T; }
__device__ __forceinline__ const Tw* t16_scale(const uint32_t* s, int q, int T) { return reinterpret_cast<const Tw*>(s) + 4 * 15 * 16 + 4 * 12 * 64 + q * 16 * 64 + T; }

__device__ __forceinline__ const uint32_t* t16_lut1(const uint32_t* s, int q) { return s + kLut1Word + q * 256; }
__device__ __forceinline__ const U2* t16_lut24(const uint32_t* s, int q) { return reinterpret_cast<const U2*>(s + kLut1Word + 512) + q * 256; }
__device__ __forceinline__ const U2* t16_lut35(const uint32_t* s, int q) { return reinterpret_cast<const U2*>(s + kLut1Word + 512 + 1024) + q * 256; }

__device__ __forceinline__ void stage_tables16(uint32_t* s_tab, const Lane16Tables* g_tab)
{
    const uint4* src = reinterpret_cast<const uint4*>(g_tab);
    uint4* dst = reinterpret_cast<uint4*>(s_tab);
    for (int i = threadIdx.x; i < kLane16TableWords / 4; i += blockDim.x) dst[i] = __ldg(src + i);
}

// thread-constant exchange bases (word offsets into a padded buffer)
struct X16 {
    int wA, rB;      // A -> B : store T + 68 k            load 68 (T>>2) + (T&3) + 4 k
    int wB, rC;      // B -> C : store wB + 4 k + (k>>2)   load 17 T + k
    int rD;          // C -> D : store 17 T + k            load rD + 17 k
    int wD;          // D -> A : store wD + 16 k           load T + 64 k + 16 (k>>2)
};
__device__ __forceinline__ X16 make_x16(int T)
{
    X16 x;
    x.wA = T;
    x.rB = 68 * (T >> 2) + (T & 3);
    x.wB = 68 * (T >> 2) + (T & 3);           // padC(64 B + j + 4 k) = 68 B + j + 4 k + (k >> 2)
    x.rC = 17 * T;
    x.rD = 272 * (T >> 4) + (T & 15);         // padC(256 hi + lo + 16 k) = 272 hi + lo + 17 k
    x.wD = 272 * (T >> 4) + (T & 15);         // padDA(256 hi + lo + 16 k) = 272 hi + lo + 16 k
    return x;
}

// forward transform: layout A in, layout C out.  bufA / bufB: the team's two exchange
// buffers; `bar` = the gate's named barrier (all 128 threads of the gate call this).
// passes F2 and F3 (after the first pass left layout A in x)
__device__ __forceinline__ void fwd16_tail(uint32_t (&x)[16], uint32_t* bufA, uint32_t* bufB, const uint32_t* s_tab,
                                           const X16& a, int T, int q, uint32_t p, int bar, int barn)
{
#pragma unroll
    for (int k = 0; k < 16; k++) bufA[a.wA + 68 * k] = x[k];
    named_bar_sync(bar, barn);
#pragma unroll
    for (int k = 0; k < 16; k++) x[k] = bufA[a.rB + 4 * k];
    pass16_fwd<4, 8>(x, t16_fw2(s_tab, q, T >> 2), 16, p);
#pragma unroll
    for (int k = 0; k < 16; k++) bufB[a.wB + 4 * k + (k >> 2)] = x[k];
    named_bar_sync(bar, barn);
#pragma unroll
    for (int k = 0; k < 16; k++) x[k] = bufB[a.rC + k];
    pass16_fwd3(x, t16_fw3(s_tab, q, T), 64, p);
}
__device__ __forceinline__ void fwd16(uint32_t (&x)[16], uint32_t* bufA, uint32_t* bufB, const uint32_t* s_tab,
                                      const X16& a, int T, int q, uint32_t p, int bar, int barn = 128)
{
    pass16_fwd<4, 8>(x, c_fw1[q], 1, p);
    fwd16_tail(x, bufA, bufB, s_tab, a, T, q, p, bar, barn);
}
// the same transform of gadget digits: x[k] in [0, 256) stands for x[k] - 128 (pass16_fwd_digits)
__device__ __forceinline__ void fwd16_digits(uint32_t (&x)[16], uint32_t* bufA, uint32_t* bufB, const uint32_t* s_tab,
                                             const X16& a, int T, int q, uint32_t p, int bar, int barn = 128)
{
    pass16_fwd_digits(x, t16_lut1(s_tab, q), t16_lut24(s_tab, q), t16_lut35(s_tab, q), c_fw1[q], p);
    fwd16_tail(x, bufA, bufB, s_tab, a, T, q, p, bar, barn);
}

// two digit transforms at once (x0 through bufA, x1 through bufB): 16 independent butterflies per
// stage.  Both buffers are in use at every exchange, so each exchange's stores wait (barrier) for
// the previous exchange's loads -- four barriers, as for two single transforms.
__device__ __forceinline__ void fwd16_tail2(uint32_t (&x0)[16], uint32_t (&x1)[16], uint32_t* bufA, uint32_t* bufB,
                                            const uint32_t* s_tab, const X16& a, int T, int q, uint32_t p, int bar,
                                            bool lead, int barn)
{
    if (lead) named_bar_sync(bar, barn);        // not needed when a barrier already follows the buffers' last loads
#pragma unroll
    for (int k = 0; k < 16; k++) { bufA[a.wA + 68 * k] = x0[k]; bufB[a.wA + 68 * k] = x1[k]; }
    named_bar_sync(bar, barn);
#pragma unroll
    for (int k = 0; k < 16; k++) { x0[k] = bufA[a.rB + 4 * k]; x1[k] = bufB[a.rB + 4 * k]; }
    pass16_fwd<4, 8>(x0, t16_fw2(s_tab, q, T >> 2), 16, p);
    pass16_fwd<4, 8>(x1, t16_fw2(s_tab, q, T >> 2), 16, p);
    named_bar_sync(bar, barn);
#pragma unroll
    for (int k = 0; k < 16; k++) { bufA[a.wB + 4 * k + (k >> 2)] = x0[k]; bufB[a.wB + 4 * k + (k >> 2)] = x1[k]; }
    named_bar_sync(bar, barn);
#pragma unroll
    for (int k = 0; k < 16; k++) { x0[k] = bufA[a.rC + k]; x1[k] = bufB[a.rC + k]; }
    pass16_fwd3(x0, t16_fw3(s_tab, q, T), 64, p);
    pass16_fwd3(x1, t16_fw3(s_tab, q, T), 64, p);
}
__device__ __forceinline__ void fwd16_digits2(uint32_t (&x0)[16], uint32_t (&x1)[16], uint32_t* bufA, uint32_t* bufB,
                                              const uint32_t* s_tab, const X16& a, int T, int q, uint32_t p, int bar,
                                              bool lead = true, int barn = 128)
{
    pass16_fwd_digits(x0, t16_lut1(s_tab, q), t16_lut24(s_tab, q), t16_lut35(s_tab, q), c_fw1[q], p);
    pass16_fwd_digits(x1, t16_lut1(s_tab, q), t16_lut24(s_tab, q), t16_lut35(s_tab, q), c_fw1[q], p);
    fwd16_tail2(x0, x1, bufA, bufB, s_tab, a, T, q, p, bar, lead, barn);
}
// the same for two general inputs
__device__ __forceinline__ void fwd16_2(uint32_t (&x0)[16], uint32_t (&x1)[16], uint32_t* bufA, uint32_t* bufB,
                                        const uint32_t* s_tab, const X16& a, int T, int q, uint32_t p, int bar,
                                        bool lead = true, int barn = 128)
{
    pass16_fwd<4, 8>(x0, c_fw1[q], 1, p);
    pass16_fwd<4, 8>(x1, c_fw1[q], 1, p);
    fwd16_tail2(x0, x1, bufA, bufB, s_tab, a, T, q, p, bar, lead, barn);
}

// inverse transform incl. n^-1 psi^-e scaling: layout C in, layout A out, values in [0, 2p)
__device__ __forceinline__ void inv16(uint32_t (&x)[16], uint32_t* bufA, uint32_t* bufB, const uint32_t* s_tab,
                                      const X16& a, int T, int q, uint32_t p, int bar, int barn = 128)
{
    pass16_inv<true>(x, c_iv1[q], 1, p);
#pragma unroll
    for (int k = 0; k < 16; k++) bufA[a.rC + k] = x[k];
    named_bar_sync(bar, barn);
#pragma unroll
    for (int k = 0; k < 16; k++) x[k] = bufA[a.rD + 17 * k];
    pass16_inv<false>(x, t16_iv2(s_tab, q, T & 15), 16, p);
#pragma unroll
    for (int k = 0; k < 16; k++) bufB[a.wD + 16 * k] = x[k];
    named_bar_sync(bar, barn);
#pragma unroll
    for (int k = 0; k < 16; k++) x[k] = bufB[T + 64 * k + 16 * (k >> 2)];
    pass16_inv3(x, t16_iv3(s_tab, q, T), 64, p);
    const Tw* sc = t16_scale(s_tab, q, T);
#pragma unroll
    for (int k = 0; k < 16; k++) {
        const Tw f = sc[k * 64];
        x[k] = shoup_mul_p(x[k], f.w, f.ws, p);
    }
}

// two inverse transforms at once (x0 through bufA, x1 through bufB); the caller guarantees that
// nobody still reads either buffer when this starts
__device__ __forceinline__ void inv16_2(uint32_t (&x0)[16], uint32_t (&x1)[16], uint32_t* bufA, uint32_t* bufB,
                                        const uint32_t* s_tab, const X16& a, int T, int q, uint32_t p, int bar, int barn = 128)
{
    pass16_inv<true>(x0, c_iv1[q], 1, p);
    pass16_inv<true>(x1, c_iv1[q], 1, p);
#pragma unroll
    for (int k = 0; k < 16; k++) { bufA[a.rC + k] = x0[k]; bufB[a.rC + k] = x1[k]; }
    named_bar_sync(bar, barn);
#pragma unroll
    for (int k = 0; k < 16; k++) { x0[k] = bufA[a.rD + 17 * k]; x1[k] = bufB[a.rD + 17 * k]; }
    pass16_inv<false>(x0, t16_iv2(s_tab, q, T & 15), 16, p);
    pass16_inv<false>(x1, t16_iv2(s_tab, q, T & 15), 16, p);
    named_bar_sync(bar, barn);
#pragma unroll
    for (int k = 0; k < 16; k++) { bufA[a.wD + 16 * k] = x0[k]; bufB[a.wD + 16 * k] = x1[k]; }
    named_bar_sync(bar, barn);
#pragma unroll
    for (int k = 0; k < 16; k++) { x0[k] = bufA[T + 64 * k + 16 * (k >> 2)]; x1[k] = bufB[T + 64 * k + 16 * (k >> 2)]; }
    pass16_inv3(x0, t16_iv3(s_tab, q, T), 64, p);
    pass16_inv3(x1, t16_iv3(s_tab, q, T), 64, p);
    const Tw* sc = t16_scale(s_tab, q, T);
#pragma unroll
    for (int k = 0; k < 16; k++) {
        const Tw f = sc[k * 64];
        x0[k] = shoup_mul_p(x0[k], f.w, f.ws, p);
        x1[k] = shoup_mul_p(x1[k], f.w, f.ws, p);
    }
}

}  // namespace b2
