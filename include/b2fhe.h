/*
 * b2fhe.h -- C ABI of the B200-native TFHE gate-bootstrapping engine.
 *
 * This is the drop-in boundary for cuFHE's bootstrapping path: plain
 * pointers and sizes, no C++/torch types.  The reference has no FFI layer; the
 * functions below replace its launch wrappers and key upload, which the C++ shim
 * include/cufhe_b200.hpp (same names/signatures as include/cufhe_gpu.cuh) calls.
 * Paths cited are relative to the reference repository.
 *
 * Conventions: every function returns 0 on success or a non-zero code (CUDA error
 * code, or B2FHE_E_*); b2fhe_last_error_string() describes the last failure of the
 * calling thread.  Nothing here ever falls back to a CPU path.
 *
 * Ciphertext formats are TFHEpp's raw arrays (default 128-bit parameters):
 *   level 0: TLWE<lvl0param> = 637 x uint16_t (1274 B), level 1: TLWE<lvl1param> =
 *   1025 x uint32_t (4100 B); batches are contiguous arrays of those.
 * Key formats are TFHEpp's raw arrays:
 *   bk  = BootstrappingKey<lvl01param>  [636][4][2][1024] uint32_t  (20,840,448 B)
 *   ksk = KeySwitchingKey<lvl10param>   [1024][7][2][637] uint16_t  (18,264,064 B)
 */
#ifndef B2FHE_H
#define B2FHE_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define B2FHE_E_NOT_INITIALIZED 10001
#define B2FHE_E_BAD_ARGUMENT    10002
#define B2FHE_E_NO_DEVICE       10003
#define B2FHE_E_BAD_FORMAT      10004   /* wire-format reader: not a blob this engine understands */
#define B2FHE_E_TOO_LARGE       10005   /* b2fhe_gate_batch_noalloc: batch beyond the fixed scratch */

/* gate op codes; (casign, cbsign, offset) per src/bootstrap_gpu.cu:424-512,590-679 */
enum b2fhe_op {
    B2FHE_NAND = 0, B2FHE_OR, B2FHE_ORYN, B2FHE_ORNY, B2FHE_AND, B2FHE_ANDYN, B2FHE_ANDNY,
    B2FHE_NOR, B2FHE_XOR, B2FHE_XNOR, B2FHE_MUX, B2FHE_NMUX, B2FHE_NOT, B2FHE_COPY, B2FHE_NUM_OPS
};

#define B2FHE_LVL0_WORDS 637
#define B2FHE_LVL1_WORDS 1025
#define B2FHE_BK_RAW_WORDS   ((size_t)636 * 4 * 2 * 1024)      /* uint32_t */
#define B2FHE_KSK_RAW_WORDS  ((size_t)1024 * 7 * 2 * 637)      /* uint16_t */
#define B2FHE_BK_NTT_BYTES   ((size_t)636 * 4 * 2 * 2 * 1024 * 4)   /* 41,680,896 */
#define B2FHE_TRGSW_RAW_WORDS ((size_t)4 * 2 * 1024)              /* TFHEpp::TRGSW<lvl1param>, uint32_t */
#define B2FHE_TRGSW_NTT_BYTES ((size_t)4 * 2 * 2 * 1024 * 4)       /* 65,536: NTT-domain TRGSW */
#define B2FHE_KSK_DEV_BYTES  ((size_t)1024 * 7 * 3 * 320 * 4)       /* 27,525,120 */

typedef void* b2fhe_stream_t;   /* a cudaStream_t (0 = default stream) */

/* ---- lifetime ----------------------------------------------------------------
 * b2fhe_init: replaces Initialize(EvalKey) = InitializeNTThandlers +
 * BootstrappingKeyToNTT + KeySwitchingKeyToDevice (src/cufhe_gates_gpu.cu:42-47,
 * src/bootstrap_gpu.cu:97-138, src/keyswitch_gpu.cu:6-16).  Host pointers.  Converts
 * the key on GPU 0 and replicates it to GPUs 1..ngpu-1 with peer copies. */
int b2fhe_init(int ngpu, const uint32_t* bk_raw, const uint16_t* ksk_raw);
/* one device only (one process per GPU): device = CUDA ordinal of this process */
/* Initialize() without keys (include/cufhe_gpu.cuh:56-57, InitializeNTThandlers src/bootstrap_gpu.cu:97-109):
 * twiddle tables + kernel attributes on devices 0..ngpu-1.  b2fhe_init* do this themselves. */
int b2fhe_setup_devices(int ngpu);
int b2fhe_init_device(int device, const uint32_t* bk_raw, const uint16_t* ksk_raw);
/* keys already converted and resident on `device` (after an NCCL broadcast of the
 * buffers produced on rank 0): adopts the pointers, does not copy or free them. */
int b2fhe_adopt_keys(int device, const void* bk_ntt_dev, const void* ksk_dev);
/* device pointers of the converted keys (to broadcast them), sizes B2FHE_*_BYTES */
int b2fhe_key_pointers(int device, void** bk_ntt_dev, void** ksk_dev);
/* replaces CleanUp / DeleteBootstrappingKeyNTT / DeleteKeySwitchingKey
 * (src/bootstrap_gpu.cu:145-155, src/keyswitch_gpu.cu:18-24) */
int b2fhe_cleanup(void);
int b2fhe_num_devices(void);
const char* b2fhe_last_error_string(void);

/* ---- gates on DEVICE buffers (the g* / device-resident path,
 * include/cufhe_gpu.cuh:229-252, launchers src/bootstrap_gpu.cu:834-1292) --------
 * out/in0/in1/in2: contiguous batches of `count` ciphertexts of `level` on `device`.
 * Two-input gates read in0,in1; MUX/NMUX read in0=inc, in1=in1, in2=in0 (reference
 * argument order Mux(out, inc, in1, in0)); NOT/COPY read in0.  out may alias an input.
 * Asynchronous on `stream`. */
int b2fhe_gate_batch(int device, b2fhe_stream_t stream, int op, int level, void* out,
                     const void* in0, const void* in1, const void* in2, size_t count);
/* per-gate op codes (host array of `count` bytes); same buffers */
int b2fhe_gate_batch_mixed(int device, b2fhe_stream_t stream, const uint8_t* ops_host, int level,
                           void* out, const void* in0, const void* in1, const void* in2, size_t count);
/* ---- gates on HOST buffers (the copying API path: CtxtCopyH2D + gate + CtxtCopyD2H,
 * src/cufhe_gates_gpu.cu:148-158).  Copies inputs H2D, runs, copies the result D2H,
 * all on `stream`; results valid after b2fhe_synchronize / stream sync. */
int b2fhe_gate_batch_host(int device, b2fhe_stream_t stream, int op, const uint8_t* ops_host, int level,
                          void* out_host, const void* in0_host, const void* in1_host,
                          const void* in2_host, size_t count);
int b2fhe_synchronize(void);   /* Synchronize(): flushes the deferred queues, then waits for all
                                  devices (include/cufhe_gpu.cuh:68-74) */
/* gates over SCATTERED device buffers (one pointer per gate; host arrays of device pointers;
 * in1/in2 entries may be null where the op does not read them): gather -> batch -> scatter */
int b2fhe_gate_batch_ptrs(int device, b2fhe_stream_t stream, const uint8_t* ops_host, int level,
                          void* const* out, const void* const* in0, const void* const* in1,
                          const void* const* in2, size_t count);
/* ---- deferred per-gate queue: how the legacy one-gate-per-call API (src/cufhe_gates_gpu.cu:
 * 148-665) gets batch-major execution.  b2fhe_defer_gate queues ONE gate on device buffers
 * (work already issued on `user_stream`, e.g. the H2D of its operands, is ordered before it);
 * the queue runs at b2fhe_defer_flush / b2fhe_defer_fence / b2fhe_defer_query /
 * b2fhe_synchronize as one batched launch sequence per dependency level (a gate that reads a
 * queued gate's output runs one level later; program order is preserved for write-after-read
 * and write-after-write).  b2fhe_defer_d2h queues a device-to-host copy after the gates. */
int b2fhe_defer_gate(int device, b2fhe_stream_t user_stream, int op, int level, void* out,
                     const void* in0, const void* in1, const void* in2);
int b2fhe_defer_h2d(int device, void* dev_dst, const void* host_src, size_t bytes);   /* operand
    upload, done when the queue runs; skipped if a queued or running gate produces dev_dst */
int b2fhe_defer_d2h(int device, void* host_dst, const void* dev_src, size_t bytes);
int b2fhe_defer_flush(int device);                          /* asynchronous */
int b2fhe_defer_fence(int device, b2fhe_stream_t stream);   /* flush; `stream` waits for it */
int b2fhe_defer_fence_buffer(int device, b2fhe_stream_t stream, const void* buffer); /* the fence, only
                                                               if a queued gate reads or writes `buffer` */
int b2fhe_defer_query(int device, int* busy);               /* flush; *busy = 1 while running */
/* the same for ONE user stream (StreamQuery(st), include/cufhe_gpu.cuh:68-74): flushes only when `stream`
 * itself has queued work; busy = flushed work still running */
int b2fhe_defer_query_stream(int device, b2fhe_stream_t stream, int* busy);
int b2fhe_defer_pending(int device);                        /* queued operations */

/* ---- stages (next-tier entry points) ---------------------------------
 * blind rotation + SampleExtract(0) of c = in (lvl0, device) -> lvl1 TLWE (device):
 * GateBootstrappingTLWE2TLWE without key switch (gatebootstrapping_gpu.cuh:287-314) */
int b2fhe_blind_rotate_batch(int device, b2fhe_stream_t stream, uint32_t* out_lvl1,
                             const uint16_t* in_lvl0, uint32_t mu, size_t count);
/* identity key switch lvl1 -> lvl0 (SampleExtractAndKeySwitch tail, src/keyswitch_gpu.cu:26-40) */
int b2fhe_keyswitch_batch(int device, b2fhe_stream_t stream, uint16_t* out_lvl0,
                          const uint32_t* in_lvl1, size_t count);

/* A gate batch that ALLOCATES NOTHING.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving
 * loops, harnesses that forbid it).  Same result as b2fhe_gate_batch(op, level 0) for the ten
 * two-input gates (op < B2FHE_MUX), device pointers, out may alias an input; the lvl1 scratch between
 * blind rotation and key switch lives in a fixed device array of B2FHE_NOALLOC_MAX_GATES x 4100 B
 * (8.4 MB, held by every GPU that loads the library until exit).  count > B2FHE_NOALLOC_MAX_GATES
 * returns B2FHE_E_TOO_LARGE.  Calls no cudaMalloc / cudaFree / allocating routine: two kernel
 * launches on `stream`.  All calls on a device share the array: back-to-back calls on one stream are
 * ordered, calls that OVERLAP on the GPU (different streams) race.  Needs b2fhe_init* first. */
#define B2FHE_NOALLOC_MAX_GATES 2048
size_t b2fhe_noalloc_max_gates(void);
int b2fhe_gate_batch_noalloc(int device, b2fhe_stream_t stream, int op, uint16_t* out_lvl0,
                             const uint16_t* in0_lvl0, const uint16_t* in1_lvl0, size_t count);

/* Bootstrap (src/bootstrap_gpu.cu:290-301,782-789): blind rotation with a RUN-TIME test-vector
 * constant mu, SampleExtract(0), identity key switch: lvl0 TLWE -> lvl0 TLWE.  out may alias in. */
int b2fhe_bootstrap_batch(int device, b2fhe_stream_t stream, uint16_t* out_lvl0,
                          const uint16_t* in_lvl0, uint32_t mu, size_t count);

/* ---- TRLWE-level entry points.  A TRLWE<lvl1param> is [a(X)][b(X)] =
 * 2 x 1024 uint32_t (8192 B); batches are contiguous.  Device pointers.
 * BootstrapTLWE2TRLWE / GateBootstrappingTLWE2TRLWElvl01NTT (src/bootstrap_gpu.cu:316-322,
 * 806-815; src/cufhe_gates_gpu.cu:88-108): blind rotation WITHOUT SampleExtract */
int b2fhe_bootstrap_tlwe2trlwe_batch(int device, b2fhe_stream_t stream, uint32_t* out_trlwe,
                                     const uint16_t* in_lvl0, uint32_t mu, size_t count);
/* SEIandKS / SampleExtractAndKeySwitch (src/keyswitch_gpu.cu:26-40, src/cufhe_gates_gpu.cu:
 * 128-146): SampleExtract(0) fused into the identity key switch, TRLWE -> lvl0 TLWE */
int b2fhe_sample_extract_keyswitch_batch(int device, b2fhe_stream_t stream, uint16_t* out_lvl0,
                                         const uint32_t* in_trlwe, size_t count);
/* Refresh (src/cufhe_gates_gpu.cu:110-126) = the two above in sequence.  NOTE: the reference's
 * __SEIandBootstrap2TRLWE__ reads the body term from an uninitialised buffer
 * (src/bootstrap_gpu.cu:342-348); this computes the intended result. */
int b2fhe_refresh_batch(int device, b2fhe_stream_t stream, uint32_t* out_trlwe,
                        const uint32_t* in_trlwe, uint32_t mu, size_t count);

/* ---- CMUX with a caller-supplied TRGSW.  Device pointers.
 * TRGSW2NTT (src/bootstrap_gpu.cu:43-95): raw TRGSW<lvl1param> [count][4][2][1024] u32 ->
 * this engine's NTT-domain form, B2FHE_TRGSW_NTT_BYTES each (opaque; two 27-bit primes,
 * Montgomery residues -- NOT the reference's single-prime FFP words). */
int b2fhe_trgsw2ntt_batch(int device, b2fhe_stream_t stream, uint32_t* out_ntt, const uint32_t* in_raw, size_t count);
/* CMUXNTT (src/bootstrap_gpu.cu:162-285, src/cufhe_gates_gpu.cu:69-85): out = c0 + (c1 - c0) [.] cs
 * on TRLWEs; cs_per_item != 0: item i uses TRGSW i, else every item uses TRGSW 0.
 * out may alias c0 or c1. */
int b2fhe_cmux_ntt_batch(int device, b2fhe_stream_t stream, uint32_t* out, const uint32_t* cs_ntt,
                         const uint32_t* c1, const uint32_t* c0, size_t count, int cs_per_item);

/* ---- NTT microbenchmark entry points (test/test_polynomial_mult_1024.cu:76-99) ------
 * All device pointers.  ntt: in[count][1024] u32 -> out[count][2][1024] residues mod the
 * two NTT primes (engine-internal order).  intt: inverse + centred CRT, mod 2^32.
 * polymul: out = a (*) b negacyclic mod 2^32, exact when |sum a_i b_j| < 2^52 with a read
 * as unsigned and b as SIGNED int32 (e.g. |b| <= 2^9); beyond that bound the words are wrong without
 * an error (tests/test_gpu_parity.py pins the bound) -- use b2fhe_polymul1024_exact_batch. */
int b2fhe_ntt1024_batch(int device, b2fhe_stream_t stream, uint32_t* out, const uint32_t* in, size_t count);
int b2fhe_intt1024_batch(int device, b2fhe_stream_t stream, uint32_t* out, const uint32_t* in, size_t count);
int b2fhe_polymul1024_batch(int device, b2fhe_stream_t stream, uint32_t* out, const uint32_t* a,
                            const uint32_t* b, size_t count);
/* the same product, exact for EVERY input (a: any u32, b: any int32; |sum| < 2^73): three NTT primes,
 * P ~ 2^80.  Covers the reference test's domain a < 2^32, b < 2^18 (test/test_polynomial_mult_1024.cu:
 * 38-39,161-163; its own tolerance is <= 2, here 0), which the two-prime product above cannot. */
int b2fhe_polymul1024_exact_batch(int device, b2fhe_stream_t stream, uint32_t* out, const uint32_t* a,
                                  const uint32_t* b, size_t count);

/* ---- wire formats.  Host memory, little-endian hosts. ----------------
 * TFHEpp's cereal portable_binary archives, default 128-bit parameter set:
 * EvalKey (TFHEpp/include/cloudkey.hpp:103-107; tutorial/client.cpp writes cloud.key) holding
 * BootstrappingKey<lvl01param> + KeySwitchingKey<lvl10param> -- exactly what the reference's
 * Initialize(EvalKey) consumes.  The writer emits that minimal key; the reader also accepts
 * archives that additionally hold the FFT / NTT forms of the same key and skips them. */
size_t b2fhe_wire_evalkey_size(void);
int b2fhe_wire_evalkey_write(void* blob, size_t cap, const uint32_t* bk_raw, const uint16_t* ksk_raw,
                             size_t* written);
int b2fhe_wire_evalkey_parse(const void* blob, size_t len, size_t* bk_offset, size_t* ksk_offset);
int b2fhe_init_from_evalkey(int ngpu, const void* blob, size_t len);   /* parse + b2fhe_init */
/* std::vector<TLWE<P>> archives (tutorial/client.cpp cloud.data, cloud.cpp result.data) */
size_t b2fhe_wire_tlwes_size(int level, size_t count);
int b2fhe_wire_tlwes_write(void* blob, size_t cap, int level, const void* cts, size_t count, size_t* written);
int b2fhe_wire_tlwes_parse(const void* blob, size_t len, int level, size_t* count, size_t* offset);
/* cache of this engine's CONVERTED keys (NTT-domain bk + device ksk, checksummed, versioned):
 * b2fhe_init_from_key_cache skips the conversion of b2fhe_init */
size_t b2fhe_key_cache_size(void);
int b2fhe_key_cache_export(int device, void* blob, size_t cap, size_t* written);
int b2fhe_init_from_key_cache(int ngpu, const void* blob, size_t len);

/* tuning / introspection */
int b2fhe_set_gates_per_cta(int gates);          /* 0 = default */
/* blind-rotation kernel: 0 = by batch size (k_blind_rotate_cl4s up to 70 gates while latency mode is on,
 * else k_blind_rotate_s), 5 = k_blind_rotate_cl4s (one gate per cluster of 4 CTAs x 128 threads, hand-overs
 * by st.async + mbarrier), 40 = k_blind_rotate_s (128 threads x 16 coefficients per gate, signed lazy
 * residues, four gates per CTA), 48 = its timing experiment that reads key row 0 at every step (WRONG
 * results, tools only), 6 / 41 = schedule-fuzzing builds of 5 / 40 (every warp sleeps a random 0..1 us around
 * every barrier, __syncwarp, mbarrier wait and hand-over; tests only).  0 / 5 / 6 / 40 / 41 produce identical bits.  -1 / -2 switch off / on the separate launch
 * of a last partial wave (on by default; does not change the variant).  -3 / -4 switch off / on the
 * bulk-copy input prefetch of b2fhe_ntt1024_batch / b2fhe_intt1024_batch and the larger CTAs of the polynomial
 * products (five / three products per CTA instead of two) -- on by default; A/B only, same bits. */
int b2fhe_set_br_variant(int variant);
/* 1 (default): small batches use the latency kernels; 0: every batch uses the throughput kernel (right for
 * many small calls in flight on many streams).  The choice depends on this setting and the batch size only. */
int b2fhe_set_latency_mode(int on);
/* key-switch kernel: 0 = by batch size (k_keyswitch_lat up to 32 gates per SM, else
 * k_keyswitch_step<16, 2>), 1 = k_keyswitch_lat (16 or 32 warps per gate), 3 = its round-1 form (A/B), 2 = k_keyswitch (one
 * free-running warp per gate), 10..18 = k_keyswitch_step<W, SYNCJ> with W = 8 / 16 / 32 gates per
 * CTA walking the mask in step, a barrier every SYNCJ = 1 / 2 / 4 coefficients (10 + 3 iW + iS),
 * 19 / 20 / 21 = <32, 8> / <32, 16> / <16, 8>.  All produce identical bits. */
int b2fhe_set_ks_variant(int variant);
uint64_t b2fhe_kernel_launches(void);             /* kernels launched by this library so far */
/* measured integer-pipe ceiling of `device` (register-only probe kernel): mode 0 = 1e9
 * 32-bit IMAD lane-ops/s, mode 1 = 1e9 Montgomery butterflies/s (the NTT's instruction mix) */
int b2fhe_int_peak(int device, int mode, double* giga_ops_per_s);

#ifdef __cplusplus
}
#endif
#endif
