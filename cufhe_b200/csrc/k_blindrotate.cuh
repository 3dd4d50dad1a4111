// k_blindrotate.cuh -- work-item and parameter types shared by the blind-rotation kernels
// (k_blindrotate_s.cuh: throughput; k_blindrotate_cl4s.cuh: latency).
//
// They replace __BlindRotatePreAdd__/__BlindRotate__/Accumulate
// (reference include/gatebootstrapping_gpu.cuh:115-345) and __SampleExtractIndex__
// (src/bootstrap_gpu.cu:366-381), which run ONE gate per 512-thread block and re-read
// the whole 41.7 MB key per gate.
#pragma once
#include "dev_common.cuh"

namespace b2 {

// one blind-rotation work item: c = ca*in0[g0] + cb*in1[g1] + (0, off) over u16, result
// (a lvl1 TLWE, 1025 words) written to out[slot].
struct BrItem {
    uint32_t g0, g1;        // ciphertext indices into src0 / src1
    int8_t ca, cb;          // linear-combination signs
    int8_t src1;            // 0: second operand from in1, 1: from in2
    int8_t pad;
    int32_t off;            // offset added to the body term (u16 arithmetic)
    uint32_t slot;          // output slot
};

struct BrParams {
    const uint16_t* in0;    // [*][637] lvl0 ciphertexts
    const uint16_t* in1;    // may be null when cb == 0
    const uint16_t* in2;    // third operand array (Mux), may be null
    uint32_t* out;          // [*][1025] lvl1 TLWEs
    const BrItem* items;    // null: uniform batch (item i = gate i with ca/cb/off below)
    const uint32_t* bk;     // NTT-domain key [636][4 rows][2 primes][2 outs][1024] (bk16_slot order)
    int count;
    int ca, cb, off;        // uniform coefficients
    uint32_t mu;            // test-vector constant (lvl1param::mu)
    int trlwe;              // 0: out[slot] = SampleExtract(0) (1025 words); 1: out[slot] = the TRLWE (2 x 1024 words)
    int first;              // k_blind_rotate_s only: this launch covers items first .. count-1
    uint32_t dbg;           // always 0 in the product; tools/ablate_brs.cu uses it as an opaque zero
};

}  // namespace b2
