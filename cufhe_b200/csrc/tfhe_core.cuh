// tfhe_core.cuh -- scalar pieces of the gate bootstrap (host+device), normative
// integer semantics of SURVEY.md Appendix A.  Each function cites the reference
// device code it restates.
#pragma once
#include "modarith.cuh"

namespace b2 {

// gate op codes of the C ABI (include/b2fhe.h)
enum GateOp : int {
    OP_NAND = 0, OP_OR, OP_ORYN, OP_ORNY, OP_AND, OP_ANDYN, OP_ANDNY,
    OP_NOR, OP_XOR, OP_XNOR, OP_MUX, OP_NMUX, OP_NOT, OP_COPY, OP_NUM
};

// (casign, cbsign, offset in units of mu) of the ten two-input gates:
// reference src/bootstrap_gpu.cu:424-512 (lvl0 I/O) and :590-679 (lvl1 I/O)
B2_HD void gate_coeffs(int op, int& ca, int& cb, int& off)
{
    //               NAND OR ORYN ORNY AND ANDYN ANDNY NOR XOR XNOR
    const int CA[10] = {-1, 1, 1, -1, 1, 1, -1, -1, 2, -2};
    const int CB[10] = {-1, 1, -1, 1, 1, -1, 1, -1, 2, -2};
    const int OF[10] = {1, 1, 1, 1, -1, -1, -1, -1, 2, -2};
    ca = CA[op]; cb = CB[op]; off = OF[op];
}

// modSwitchFromTorus<lvl01param> on the mask part, with the reference's rounding
// offset 2^(16-2-10) = 16 (include/gatebootstrapping_gpu.cuh:10-15,336-341)
B2_HD uint32_t modswitch_a(uint32_t c16) { return ((c16 + 16u) & 0xFFFFu) >> 5; }
// the body term: no rounding offset (gatebootstrapping_gpu.cuh:327-331)
B2_HD uint32_t modswitch_b(uint32_t c16) { return 2u * kN - ((c16 & 0xFFFFu) >> 5); }

// RotatedTestVector (gatebootstrapping_gpu.cuh:29-52): b part of mu * X^bbar * (1+X+..)
B2_HD uint32_t testvec_b(uint32_t c, uint32_t bbar, uint32_t mu)
{
    if (bbar == 2u * kN) return mu;
    return ((c < (bbar & (kN - 1))) ^ (bbar >> kNbit)) ? 0u - mu : mu;
}

// sign of coefficient c of (X^abar * poly): true -> negate poly[(c - abar) & 1023]
// (gatebootstrapping_gpu.cuh:159-166)
B2_HD bool rot_neg(uint32_t c, uint32_t abar) { return ((c < (abar & (kN - 1))) ^ (abar >> kNbit)) != 0; }

// decomposition offsets: offsetgen<lvl1param>() = 0x80800000 plus the reference's
// rounding 1 << (32 - l*Bgbit - 1) = 0x8000 (gatebootstrapping_gpu.cuh:17-27,148-150)
constexpr uint32_t kDecompAdd = 0x80808000u;
// digit d (0 = most significant) in [-128, 127] (gatebootstrapping_gpu.cuh:172-179)
B2_HD int32_t digit(uint32_t t, int d) { return (int32_t)((t >> (24 - 8 * d)) & 0xFFu) - 128; }

// identity key switch constants (include/keyswitch_gpu.cuh:13-23,92-95)
constexpr uint32_t kIksAdd = 0xAAA80000u + (1u << 17);
// digit d (0 = most significant) of the key-switch decomposition, in {0,1,2,3};
// value = digit - 2 in {-2,-1,0,1} (keyswitch_gpu.cuh:112-131)
B2_HD uint32_t iks_digit(uint32_t t, int d) { return (t >> (30 - 2 * d)) & 3u; }
// b-term rounding u32 -> u16 (keyswitch_gpu.cuh:108-109)
B2_HD uint32_t iks_round_b(uint32_t b) { return ((b + (1u << 15)) >> 16) & 0xFFFFu; }

}  // namespace b2
