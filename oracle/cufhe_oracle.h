/*
 * cufhe_oracle.h -- CPU oracle for the cuFHE gate-bootstrapping hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product path (cufhe_b200/, include/)
 * may link, import or execute this.  Only tests/, __graft_entry__.smoke() and the
 * cpu_baseline / --impl reference legs of bench.py use it, as the checker.
 *
 * It is a plain-C restatement of the *integer semantics of cuFHE's device code*
 * (SURVEY.md Appendix A).  Each function cites the reference file:line it follows
 * (paths relative to /root/reference).
 *
 * Parity status: the reference ships no golden vectors for this path
 * (SURVEY.md 8c).  The oracle is pinned by
 *   (1) the reference's own schoolbook truth for the polynomial product
 *       (test/test_polynomial_mult_1024.cu:51-73) -- orc_polymul_schoolbook,
 *   (2) the truth tables of test/plain.h:10-69 through decryption,
 *   (3) decrypt-level agreement with TFHEpp built from the reference sources
 *       (oracle/_ref/tfhepp_check, see oracle/Makefile), and
 *   (4) byte equality with the reference cuFHE GPU build run on the B200 box
 *       (oracle/_ref/cufhe_ref_dump; tests/test_reference_gpu.py).
 */
#ifndef CUFHE_ORACLE_H
#define CUFHE_ORACLE_H
#include <stdint.h>
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif

/* TFHEpp default 128-bit parameters (thirdparties/TFHEpp/include/params/128bit.hpp:7-72,226-235) */
#define ORC_n      636      /* lvl0 dimension            */
#define ORC_N      1024     /* lvl1 ring degree          */
#define ORC_NBIT   10
#define ORC_L      2        /* gadget levels             */
#define ORC_BGBIT  8
#define ORC_T      7        /* IKS digits                */
#define ORC_BASEBIT 2
#define ORC_MU0    (1u << 13)   /* lvl0param::mu */
#define ORC_MU1    (1u << 29)   /* lvl1param::mu */
#define ORC_LVL0_WORDS (ORC_n + 1)   /* u16 words per lvl0 TLWE  */
#define ORC_LVL1_WORDS (ORC_N + 1)   /* u32 words per lvl1 TLWE  */
#define ORC_BK_WORDS   ((size_t)ORC_n * 4 * 2 * ORC_N)            /* u32 */
#define ORC_KSK_WORDS  ((size_t)ORC_N * ORC_T * 2 * ORC_LVL0_WORDS) /* u16 */

/* gate op codes (same numbering as include/b2fhe.h) */
enum {
    ORC_NAND = 0, ORC_OR, ORC_ORYN, ORC_ORNY, ORC_AND, ORC_ANDYN, ORC_ANDNY,
    ORC_NOR, ORC_XOR, ORC_XNOR, ORC_MUX, ORC_NMUX, ORC_NOT, ORC_COPY, ORC_NUM_OPS
};

/* ---- deterministic key / ciphertext generation (follows TFHEpp's algorithms,
 *      seeded because TFHEpp itself is unseeded) ---------------- */
void orc_keygen(uint64_t seed, int32_t *sk0 /*[636] in {0,1}*/,
                int32_t *sk1 /*[1024] in {-1,0,1}*/,
                uint32_t *bk /*[636][4][2][1024]*/,
                uint16_t *ksk /*[1024][7][2][637]*/);
void orc_encrypt_lvl0(uint64_t seed, uint64_t index, int bit, const int32_t *sk0, uint16_t *ct);
void orc_encrypt_lvl1(uint64_t seed, uint64_t index, int bit, const int32_t *sk1, uint32_t *ct);
int  orc_decrypt_lvl0(const uint16_t *ct, const int32_t *sk0);
int  orc_decrypt_lvl1(const uint32_t *ct, const int32_t *sk1);
int  orc_truth(int op, int a, int b, int c); /* test/plain.h:10-69; (a,b)=(in0,in1) or (inc,in1,in0) */

/* ---- the hot path, exact --------------------------------------------------- */
/* negacyclic product mod 2^32, schoolbook (test/test_polynomial_mult_1024.cu:51-73) */
void orc_polymul_schoolbook(uint32_t *out, const uint32_t *a, const uint32_t *b);
/* same product through the oracle's exact 64-bit NTT (centered lift) */
void orc_polymul_ntt(uint32_t *out, const uint32_t *a, const uint32_t *b);

/* one CMUX step acc += (X^abar - 1) acc [.] BK_i  (gatebootstrapping_gpu.cuh:115-285).
 * use_schoolbook != 0 selects the O(N^2) integer convolution instead of the NTT. */
void orc_cmux_step(uint32_t *acc /*[2][1024]*/, uint32_t abar,
                   const uint32_t *bk_i /*[4][2][1024]*/, int use_schoolbook);

/* CMUXNTT (src/bootstrap_gpu.cu:162-285): out = c0 + (c1 - c0) [.] TRGSW on TRLWEs [2][1024],
 * trgsw = raw TFHEpp::TRGSW<lvl1param> [4][2][1024] */
void orc_cmux(uint32_t *out, const uint32_t *trgsw, const uint32_t *c1, const uint32_t *c0, int use_schoolbook);

/* blind rotation of the lvl0 TLWE c (already the linear combination) with test
 * vector mu, followed by SampleExtract(0): out is a lvl1 TLWE.
 * (gatebootstrapping_gpu.cuh:287-345, bootstrap_gpu.cu:366-381) */
void orc_blind_rotate_extract(uint32_t *out, const uint16_t *c, uint32_t mu,
                              const uint32_t *bk, int use_schoolbook);
/* TRLWE-level entry points: blind rotation without SampleExtract
 * (BootstrapTLWE2TRLWE, src/bootstrap_gpu.cu:316-322,806-815), SampleExtract(0)+key switch
 * (SEIandKS, src/keyswitch_gpu.cu:26-40), and Refresh = both in sequence (the reference's
 * __SEIandBootstrap2TRLWE__ with its b-term bug fixed) */
void orc_blind_rotate_trlwe(uint32_t *out /*[2][1024]*/, const uint16_t *c, uint32_t mu, const uint32_t *bk);
void orc_sample_extract_keyswitch(uint16_t *out, const uint32_t *trlwe, const uint16_t *ksk);
void orc_refresh(uint32_t *out, const uint32_t *trlwe, uint32_t mu, const uint32_t *bk, const uint16_t *ksk);
/* identity key switch lvl1 -> lvl0 (keyswitch_gpu.cuh:83-134) */
void orc_keyswitch(uint16_t *out, const uint32_t *in, const uint16_t *ksk);

/* full gates.  level 0: lvl0 in/out, BR->SE->IKS (bootstrap_gpu.cu:402-421,424-588);
 *              level 1: lvl1 in/out, IKS->BR->SE (bootstrap_gpu.cu:383-400,590-780).
 * in2 is only read by MUX/NMUX (in0=inc, in1=in1, in2=in0 of the reference). */
void orc_gate(int op, int level, void *out, const void *in0, const void *in1,
              const void *in2, const uint32_t *bk, const uint16_t *ksk);
/* batch of independent gates, contiguous arrays, OpenMP over gates;
 * ops == NULL means every gate is `op`. */
void orc_gate_batch(int op, const uint8_t *ops, int level, void *out, const void *in0,
                    const void *in1, const void *in2, size_t count,
                    const uint32_t *bk, const uint16_t *ksk, int threads);

#ifdef __cplusplus
}
#endif
#endif
