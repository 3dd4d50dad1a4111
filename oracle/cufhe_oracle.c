/*
 * cufhe_oracle.c -- CPU oracle (TEST INFRASTRUCTURE ONLY; see cufhe_oracle.h).
 *
 * Plain-C restatement of the integer semantics of cuFHE's gate bootstrap.
 * The arithmetic is deliberately different from the CUDA product: a single
 * 64-bit Goldilocks-prime NTT (or schoolbook), while the product uses two
 * 27-bit primes + CRT.  Both must equal the exact integer convolution
 * (SURVEY.md F2), so agreement is a real check.
 */
#include "cufhe_oracle.h"
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef unsigned __int128 u128;

/* ------------------------------------------------------------------ PRNG --- */
static inline uint64_t splitmix64(uint64_t *s)
{
    uint64_t z = (*s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
/* integer-only approximate Gaussian, standard deviation sigma_q16 / 2^16 (in
 * units of the torus LSB): Irwin-Hall sum of 12 uniforms.  TFHEpp draws
 * ModularGaussian from std::normal_distribution (utils.hpp); the distribution
 * only matters for noise margins, not for the deterministic hot path. */
static inline int64_t gauss_fx(uint64_t *s, uint64_t sigma_q16)
{
    int64_t acc = 0;
    for (int i = 0; i < 6; i++) {
        uint64_t r = splitmix64(s);
        acc += (int64_t)(r & 0xffffffffu) + (int64_t)(r >> 32);
    }
    acc -= 6ll * 0xffffffffll;             /* mean 0, variance 2^64 */
    /* acc / 2^32 ~ N(0,1);  result = round(acc * sigma / 2^48) */
    __int128 v = (__int128)acc * (__int128)sigma_q16;
    v += ((__int128)1 << 47);
    return (int64_t)(v >> 48);
}

/* ------------------------------------------- Goldilocks field 2^64-2^32+1 --- */
#define GP 0xFFFFFFFF00000001ull
static inline uint64_t g_add(uint64_t a, uint64_t b)
{
    uint64_t s = a + b;
    if (s < a || s >= GP) s -= GP;
    return s;
}
static inline uint64_t g_sub(uint64_t a, uint64_t b) { return a >= b ? a - b : a + (GP - b); }
static inline uint64_t g_mul(uint64_t a, uint64_t b)
{
    u128 x = (u128)a * b;
    uint64_t lo = (uint64_t)x, hi = (uint64_t)(x >> 64);
    uint64_t hh = hi >> 32, hl = hi & 0xffffffffu;
    /* 2^64 = 2^32 - 1, 2^96 = -1 (mod p) */
    uint64_t t = lo - hh;
    if (lo < hh) t -= 0xffffffffu;          /* borrow: + p */
    uint64_t m = hl * 0xffffffffu;
    uint64_t r = t + m;
    if (r < t) r += 0xffffffffu;            /* carry: 2^64 = 2^32-1 */
    if (r >= GP) r -= GP;
    return r;
}
static uint64_t g_pow(uint64_t b, uint64_t e)
{
    uint64_t r = 1;
    while (e) { if (e & 1) r = g_mul(r, b); b = g_mul(b, b); e >>= 1; }
    return r;
}

static uint64_t g_psi[ORC_N], g_ipsi[ORC_N];   /* psi^brv(i), psi^-brv(i) */
static uint64_t g_ninv;
static int g_tables_ready = 0;
static unsigned brv10(unsigned x)
{
    unsigned r = 0;
    for (int i = 0; i < ORC_NBIT; i++) r |= ((x >> i) & 1u) << (ORC_NBIT - 1 - i);
    return r;
}
static void g_init(void)
{
    if (g_tables_ready) return;
#ifdef _OPENMP
#pragma omp critical(orc_tables)
#endif
    {
        if (!g_tables_ready) {
            uint64_t psi = g_pow(7, (GP - 1) / (2 * ORC_N)); /* 7 generates F_p^* */
            uint64_t ipsi = g_pow(psi, GP - 2);
            for (unsigned i = 0; i < ORC_N; i++) {
                g_psi[i] = g_pow(psi, brv10(i));
                g_ipsi[i] = g_pow(ipsi, brv10(i));
            }
            g_ninv = g_pow(ORC_N, GP - 2);
            g_tables_ready = 1;
        }
    }
}
/* psi-merged negacyclic NTT (Longa-Naehrig): CT forward natural->bit-reversed,
 * GS inverse bit-reversed->natural.  Same structure as the reference's
 * SmallForwardNTT_1024/SmallInverseNTT_1024 (include/ntt_gpu/ntt_gpuntt.cuh:232-392)
 * but over a different prime -- results after the centred lift are identical (F2). */
static void g_ntt(uint64_t *a)
{
    unsigned t = ORC_N;
    for (unsigned m = 1; m < ORC_N; m <<= 1) {
        t >>= 1;
        for (unsigned i = 0; i < m; i++) {
            uint64_t w = g_psi[m + i];
            for (unsigned j = 2 * i * t; j < 2 * i * t + t; j++) {
                uint64_t u = a[j], v = g_mul(a[j + t], w);
                a[j] = g_add(u, v);
                a[j + t] = g_sub(u, v);
            }
        }
    }
}
static void g_intt(uint64_t *a)
{
    unsigned t = 1;
    for (unsigned m = ORC_N; m > 1; m >>= 1) {
        unsigned h = m >> 1, j1 = 0;
        for (unsigned i = 0; i < h; i++) {
            uint64_t w = g_ipsi[h + i];
            for (unsigned j = j1; j < j1 + t; j++) {
                uint64_t u = a[j], v = a[j + t];
                a[j] = g_add(u, v);
                a[j + t] = g_mul(g_sub(u, v), w);
            }
            j1 += 2 * t;
        }
        t <<= 1;
    }
    for (unsigned j = 0; j < ORC_N; j++) a[j] = g_mul(a[j], g_ninv);
}
/* lift rule of the reference: v > p/2 ? v - p : v, then truncate to 32 bits
 * (include/gatebootstrapping_gpu.cuh:258,270-281) */
static inline uint32_t g_lift32(uint64_t v)
{
    return (v > GP / 2) ? (uint32_t)(v - GP) : (uint32_t)v;
}
/* signed small integer -> residue, as FFP(int32_t) (include/ntt_gpu/ntt_gpuntt.cuh:57-73) */
static inline uint64_t g_from_i64(int64_t v) { return v < 0 ? GP - (uint64_t)(-v) : (uint64_t)v; }

void orc_polymul_schoolbook(uint32_t *out, const uint32_t *a, const uint32_t *b)
{
    /* test/test_polynomial_mult_1024.cu:51-73: wrap-around with sign flip, mod 2^32 */
    uint32_t r[ORC_N];
    memset(r, 0, sizeof r);
    for (unsigned i = 0; i < ORC_N; i++)
        for (unsigned j = 0; j < ORC_N; j++) {
            uint32_t p = a[i] * b[j];
            unsigned k = i + j;
            if (k < ORC_N) r[k] += p; else r[k - ORC_N] -= p;
        }
    memcpy(out, r, sizeof r);
}
void orc_polymul_ntt(uint32_t *out, const uint32_t *a, const uint32_t *b)
{
    /* exact when |sum| < p/2: a is treated as unsigned u32, b as SIGNED int32
     * (the workload: b = digits or small values); for b < 2^18 unsigned as in the
     * reference test (test_polynomial_mult_1024.cu:161-163) signed==unsigned. */
    g_init();
    uint64_t fa[ORC_N], fb[ORC_N];
    for (unsigned i = 0; i < ORC_N; i++) { fa[i] = a[i]; fb[i] = g_from_i64((int32_t)b[i]); }
    g_ntt(fa); g_ntt(fb);
    for (unsigned i = 0; i < ORC_N; i++) fa[i] = g_mul(fa[i], fb[i]);
    g_intt(fa);
    for (unsigned i = 0; i < ORC_N; i++) out[i] = g_lift32(fa[i]);
}

/* ------------------------------------------------------------ hot path --- */
#define DECOMP_OFFSET 0x80800000u  /* offsetgen<lvl1param>(): gatebootstrapping_gpu.cuh:17-27 */
#define DECOMP_ROUND  0x00008000u  /* 1 << (32 - l*Bgbit - 1): gatebootstrapping_gpu.cuh:148-150 */

/* digit d (0 = most significant) of t, in [-128,127] (gatebootstrapping_gpu.cuh:172-179) */
static inline int32_t gadget_digit(uint32_t t, int d)
{
    return (int32_t)((t >> (32 - (d + 1) * ORC_BGBIT)) & 0xFFu) - 128;
}
/* r = (X^abar - 1) * acc_j, then + offsets (gatebootstrapping_gpu.cuh:159-170) */
static inline uint32_t rotated_minus_one(const uint32_t *poly, unsigned c, uint32_t abar)
{
    uint32_t temp = poly[(c - abar) & (ORC_N - 1)];
    if ((c < (abar & (ORC_N - 1))) ^ (abar >> ORC_NBIT)) temp = 0u - temp;
    temp -= poly[c];
    return temp + DECOMP_OFFSET + DECOMP_ROUND;
}

static void ext_product_accumulate(uint32_t *acc, int32_t (*dig)[ORC_N], const uint32_t *bk_i, int use_schoolbook);

void orc_cmux_step(uint32_t *acc, uint32_t abar, const uint32_t *bk_i, int use_schoolbook)
{
    /* Accumulate<P> (gatebootstrapping_gpu.cuh:115-285): rows j*l+digit, outputs out_k,
     * BK index ((k+1)*row + out_k) * N + coeff (:217) */
    static _Thread_local int32_t dig[4][ORC_N];
    for (int j = 0; j < 2; j++)
        for (unsigned c = 0; c < ORC_N; c++) {
            uint32_t t = rotated_minus_one(acc + j * ORC_N, c, abar);
            dig[2 * j + 0][c] = gadget_digit(t, 0);
            dig[2 * j + 1][c] = gadget_digit(t, 1);
        }
    ext_product_accumulate(acc, dig, bk_i, use_schoolbook);
}

/* CMUXNTT (src/bootstrap_gpu.cu:162-285, src/cufhe_gates_gpu.cu:69-85):
 * out = c0 + (c1 - c0) [.] TRGSW, decomposition with the same offsets as the blind rotation
 * (TRLWESubAndDecomposition :162-194); trgsw = raw TRGSW<lvl1param> [4 rows][2][1024] */
void orc_cmux(uint32_t *out, const uint32_t *trgsw, const uint32_t *c1, const uint32_t *c0, int use_schoolbook)
{
    static _Thread_local int32_t dig[4][ORC_N];
    for (int j = 0; j < 2; j++)
        for (unsigned c = 0; c < ORC_N; c++) {
            uint32_t t = c1[j * ORC_N + c] - c0[j * ORC_N + c] + DECOMP_OFFSET + DECOMP_ROUND;
            dig[2 * j + 0][c] = gadget_digit(t, 0);
            dig[2 * j + 1][c] = gadget_digit(t, 1);
        }
    uint32_t acc[2 * ORC_N];
    memcpy(acc, c0, sizeof acc);
    ext_product_accumulate(acc, dig, trgsw, use_schoolbook);
    memcpy(out, acc, sizeof acc);
}

/* acc += sum_row dig[row] (*) bk_i[row][out]  (negacyclic, exact, mod 2^32) */
static void ext_product_accumulate(uint32_t *acc, int32_t (*dig)[ORC_N], const uint32_t *bk_i, int use_schoolbook)
{
    if (use_schoolbook) {
        /* exact integer convolution; wrapping u32 accumulation == (sum mod 2^32) */
        for (int o = 0; o < 2; o++) {
            uint32_t s[ORC_N];
            memset(s, 0, sizeof s);
            for (int row = 0; row < 4; row++) {
                const uint32_t *b = bk_i + (size_t)(2 * row + o) * ORC_N;
                for (unsigned x = 0; x < ORC_N; x++) {
                    uint32_t d = (uint32_t)dig[row][x];
                    if (!d) continue;
                    for (unsigned y = 0; y < ORC_N; y++) {
                        uint32_t p = d * b[y];
                        unsigned k = x + y;
                        if (k < ORC_N) s[k] += p; else s[k - ORC_N] -= p;
                    }
                }
            }
            for (unsigned c = 0; c < ORC_N; c++) acc[o * ORC_N + c] += s[c];
        }
        return;
    }
    g_init();
    uint64_t sum[2][ORC_N];
    memset(sum, 0, sizeof sum);
    uint64_t w[ORC_N], b[ORC_N];
    for (int row = 0; row < 4; row++) {
        for (unsigned c = 0; c < ORC_N; c++) w[c] = g_from_i64(dig[row][c]);
        g_ntt(w);
        for (int o = 0; o < 2; o++) {
            const uint32_t *src = bk_i + (size_t)(2 * row + o) * ORC_N;
            for (unsigned c = 0; c < ORC_N; c++) b[c] = src[c];   /* FFP(uint32_t): unsigned */
            g_ntt(b);
            for (unsigned c = 0; c < ORC_N; c++) sum[o][c] = g_add(sum[o][c], g_mul(w[c], b[c]));
        }
    }
    for (int o = 0; o < 2; o++) {
        g_intt(sum[o]);
        for (unsigned c = 0; c < ORC_N; c++) acc[o * ORC_N + c] += g_lift32(sum[o][c]);
    }
}

/* NTT-domain copy of the bootstrapping key, built lazily per key pointer so that
 * batches do not re-transform BK rows (the reference does the same once at
 * Initialize: __TRGSW2NTT__, src/bootstrap_gpu.cu:43-70,111-138). */
static const uint32_t *g_bk_src = NULL;
static uint64_t *g_bk_ntt = NULL;
static uint64_t g_bk_fingerprint = 0;
static const uint64_t *bk_ntt_get(const uint32_t *bk)
{
    g_init();
    uint64_t fp = 0;
    for (size_t i = 0; i < ORC_BK_WORDS; i += 4099) fp = fp * 0x100000001B3ull + bk[i];
    if (g_bk_src == bk && g_bk_fingerprint == fp && g_bk_ntt) return g_bk_ntt;
    if (!g_bk_ntt) g_bk_ntt = (uint64_t *)malloc(ORC_BK_WORDS * sizeof(uint64_t));
    size_t npoly = ORC_BK_WORDS / ORC_N;
#ifdef _OPENMP
#pragma omp parallel for schedule(static)
#endif
    for (size_t p = 0; p < npoly; p++) {
        uint64_t *dst = g_bk_ntt + p * ORC_N;
        for (unsigned c = 0; c < ORC_N; c++) dst[c] = bk[p * ORC_N + c];
        g_ntt(dst);
    }
    g_bk_src = bk; g_bk_fingerprint = fp;
    return g_bk_ntt;
}
static void cmux_step_pre(uint32_t *acc, uint32_t abar, const uint64_t *bkn_i)
{
    uint64_t sum[2][ORC_N], w[ORC_N];
    uint32_t t[2][ORC_N];
    memset(sum, 0, sizeof sum);
    for (int j = 0; j < 2; j++)
        for (unsigned c = 0; c < ORC_N; c++) t[j][c] = rotated_minus_one(acc + j * ORC_N, c, abar);
    for (int row = 0; row < 4; row++) {
        for (unsigned c = 0; c < ORC_N; c++) w[c] = g_from_i64(gadget_digit(t[row >> 1][c], row & 1));
        g_ntt(w);
        for (int o = 0; o < 2; o++) {
            const uint64_t *b = bkn_i + (size_t)(2 * row + o) * ORC_N;
            for (unsigned c = 0; c < ORC_N; c++) sum[o][c] = g_add(sum[o][c], g_mul(w[c], b[c]));
        }
    }
    for (int o = 0; o < 2; o++) {
        g_intt(sum[o]);
        for (unsigned c = 0; c < ORC_N; c++) acc[o * ORC_N + c] += g_lift32(sum[o][c]);
    }
}

static void blind_rotate_trlwe_impl(uint32_t *acc /*[2][1024]*/, const uint16_t *c, uint32_t mu,
                                    const uint32_t *bk, const uint64_t *bkn, int use_schoolbook)
{
    /* modSwitchFromTorus: u16 phase >> (16-1-10) (gatebootstrapping_gpu.cuh:10-15);
     * NO rounding offset on b (:296-299,327-331) */
    uint32_t bbar = 2 * ORC_N - ((uint32_t)c[ORC_n] >> 5);
    /* RotatedTestVector (gatebootstrapping_gpu.cuh:29-52) */
    for (unsigned i = 0; i < ORC_N; i++) {
        acc[i] = 0;
        if (bbar == 2 * ORC_N) acc[ORC_N + i] = mu;
        else acc[ORC_N + i] = ((i < (bbar & (ORC_N - 1))) ^ (bbar >> ORC_NBIT)) ? 0u - mu : mu;
    }
    for (unsigned i = 0; i < ORC_n; i++) {
        /* roundoffset 1 << (16-2-10) = 16, truncated to u16 before the shift (:303-307) */
        uint32_t abar = (uint32_t)((uint16_t)(c[i] + 16u)) >> 5;
        if (bkn) cmux_step_pre(acc, abar, bkn + (size_t)i * 8 * ORC_N);
        else orc_cmux_step(acc, abar, bk + (size_t)i * 8 * ORC_N, use_schoolbook);
    }
}
static void sample_extract0(uint32_t *out, const uint32_t *acc)
{
    /* __SampleExtractIndex__<P,0> (src/bootstrap_gpu.cu:366-381) */
    out[0] = acc[0];
    for (unsigned i = 1; i < ORC_N; i++) out[i] = 0u - acc[ORC_N - i];
    out[ORC_N] = acc[ORC_N];
}
static void blind_rotate_extract_impl(uint32_t *out, const uint16_t *c, uint32_t mu,
                                      const uint32_t *bk, const uint64_t *bkn, int use_schoolbook)
{
    uint32_t acc[2 * ORC_N];
    blind_rotate_trlwe_impl(acc, c, mu, bk, bkn, use_schoolbook);
    sample_extract0(out, acc);
}
/* BootstrapTLWE2TRLWE = __BlindRotateGlobal__ (src/bootstrap_gpu.cu:316-322,806-815):
 * blind rotation without SampleExtract, out = TRLWE [a(X)][b(X)] */
void orc_blind_rotate_trlwe(uint32_t *out, const uint16_t *c, uint32_t mu, const uint32_t *bk)
{
    blind_rotate_trlwe_impl(out, c, mu, bk, bk_ntt_get(bk), 0);
}
/* SEIandKS = KeySwitch<lvl10param> on a TRLWE (include/keyswitch_gpu.cuh:25-78,
 * src/keyswitch_gpu.cu:26-40): SampleExtract(0) fused into the key switch */
void orc_sample_extract_keyswitch(uint16_t *out, const uint32_t *trlwe, const uint16_t *ksk)
{
    uint32_t t[ORC_LVL1_WORDS];
    sample_extract0(t, trlwe);
    orc_keyswitch(out, t, ksk);
}
/* Refresh = SEIandBootstrap2TRLWE with the reference's b-term bug fixed (it
 * reads the body from an uninitialised buffer, src/bootstrap_gpu.cu:342-348): key switch of
 * SampleExtract(0), then blind rotation to a fresh TRLWE */
void orc_refresh(uint32_t *out, const uint32_t *trlwe, uint32_t mu, const uint32_t *bk, const uint16_t *ksk)
{
    uint16_t c[ORC_LVL0_WORDS];
    orc_sample_extract_keyswitch(c, trlwe, ksk);
    orc_blind_rotate_trlwe(out, c, mu, bk);
}
void orc_blind_rotate_extract(uint32_t *out, const uint16_t *c, uint32_t mu,
                              const uint32_t *bk, int use_schoolbook)
{
    blind_rotate_extract_impl(out, c, mu, bk, use_schoolbook ? NULL : bk_ntt_get(bk), use_schoolbook);
}

void orc_keyswitch(uint16_t *out, const uint32_t *in, const uint16_t *ksk)
{
    /* KeySwitchFromTLWE<lvl10param> (include/keyswitch_gpu.cuh:83-134) */
    const uint32_t decompoffset = 0xAAA80000u; /* iksoffsetgen: sum_{i=1..7} 2 * 2^(32-2i) (:13-23) */
    const uint32_t roundoffset = 1u << (32 - (1 + ORC_BASEBIT * ORC_T)); /* 2^17 (:92-95) */
    uint16_t res[ORC_LVL0_WORDS];
    memset(res, 0, sizeof res);
    res[ORC_n] = (uint16_t)((in[ORC_N] + (1u << 15)) >> 16);            /* (:108-109) */
    for (unsigned j = 0; j < ORC_N; j++) {
        uint32_t tmp = in[j] + decompoffset + roundoffset;
        for (int k = 0; k < ORC_T; k++) {
            int32_t val = (int32_t)((tmp >> (32 - (k + 1) * ORC_BASEBIT)) & 3u) - 2;
            if (val == 0) continue;
            int idx = (val < 0 ? -val : val) - 1;
            const uint16_t *row = ksk + (((size_t)j * ORC_T + k) * 2 + idx) * ORC_LVL0_WORDS;
            if (val > 0) for (unsigned i = 0; i < ORC_LVL0_WORDS; i++) res[i] -= row[i];
            else         for (unsigned i = 0; i < ORC_LVL0_WORDS; i++) res[i] += row[i];
        }
    }
    memcpy(out, res, sizeof res);
}

/* ----------------------------------------------------------------- gates --- */
/* (casign, cbsign, offset/mu) table: src/bootstrap_gpu.cu:424-512 (lvl0) and :590-679 (lvl1) */
static const int8_t GATE_CA[10]  = {-1, 1, 1,-1, 1, 1,-1,-1, 2,-2};
static const int8_t GATE_CB[10]  = {-1, 1,-1, 1, 1,-1, 1,-1, 2,-2};
static const int8_t GATE_OFF[10] = { 1, 1, 1, 1,-1,-1,-1,-1, 2,-2};
/*            op:  NAND OR ORYN ORNY AND ANDYN ANDNY NOR XOR XNOR */

static void lincomb0(uint16_t *c, int ca, const uint16_t *x0, int cb, const uint16_t *x1, int off)
{
    /* int arithmetic then truncation to u16 (gatebootstrapping_gpu.cuh:327-341) */
    for (unsigned i = 0; i < ORC_n; i++) c[i] = (uint16_t)(ca * (int)x0[i] + cb * (int)x1[i]);
    c[ORC_n] = (uint16_t)(off + ca * (int)x0[ORC_n] + cb * (int)x1[ORC_n]);
}
static void lincomb1(uint32_t *c, int ca, const uint32_t *x0, int cb, const uint32_t *x1, uint32_t off)
{
    /* u32 wraparound (keyswitch_gpu.cuh:158,168) */
    for (unsigned i = 0; i < ORC_N; i++) c[i] = (uint32_t)ca * x0[i] + (uint32_t)cb * x1[i];
    c[ORC_N] = (uint32_t)ca * x0[ORC_N] + (uint32_t)cb * x1[ORC_N] + off;
}

static void gate_impl(int op, int level, void *out, const void *in0, const void *in1,
                      const void *in2, const uint32_t *bk, const uint64_t *bkn, const uint16_t *ksk)
{
    if (op == ORC_NOT || op == ORC_COPY) {          /* src/bootstrap_gpu.cu:681-703 */
        if (level == 0) {
            const uint16_t *x = (const uint16_t *)in0; uint16_t *o = (uint16_t *)out;
            for (unsigned i = 0; i < ORC_LVL0_WORDS; i++) o[i] = (op == ORC_NOT) ? (uint16_t)(0 - x[i]) : x[i];
        } else {
            const uint32_t *x = (const uint32_t *)in0; uint32_t *o = (uint32_t *)out;
            for (unsigned i = 0; i < ORC_LVL1_WORDS; i++) o[i] = (op == ORC_NOT) ? 0u - x[i] : x[i];
        }
        return;
    }
    uint16_t c0[ORC_LVL0_WORDS];
    uint32_t e[ORC_LVL1_WORDS], e2[ORC_LVL1_WORDS], t1[ORC_LVL1_WORDS];
    if (op < 10) {
        int ca = GATE_CA[op], cb = GATE_CB[op], so = GATE_OFF[op];
        if (level == 0) {    /* __HomGate__<brP,mu,iksP,...>: src/bootstrap_gpu.cu:402-421 */
            lincomb0(c0, ca, (const uint16_t *)in0, cb, (const uint16_t *)in1, so * (int)ORC_MU0);
            blind_rotate_extract_impl(e, c0, ORC_MU1, bk, bkn, 0);
            orc_keyswitch((uint16_t *)out, e, ksk);
        } else {             /* __HomGate__<iksP,brP,mu,...>: src/bootstrap_gpu.cu:383-400 */
            lincomb1(t1, ca, (const uint32_t *)in0, cb, (const uint32_t *)in1, (uint32_t)so * ORC_MU1);
            orc_keyswitch(c0, t1, ksk);              /* IdentityKeySwitchPreAdd: keyswitch_gpu.cuh:136-188 */
            blind_rotate_extract_impl((uint32_t *)out, c0, ORC_MU1, bk, bkn, 0);
        }
        return;
    }
    /* MUX / NMUX: in0 = inc, in1 = in1, in2 = in0 of the reference signature */
    int neg = (op == ORC_NMUX);
    if (level == 0) {        /* src/bootstrap_gpu.cu:515-588 */
        lincomb0(c0, 1, (const uint16_t *)in0, 1, (const uint16_t *)in1, -(int)ORC_MU0);
        blind_rotate_extract_impl(e, c0, ORC_MU1, bk, bkn, 0);
        lincomb0(c0, -1, (const uint16_t *)in0, 1, (const uint16_t *)in2, -(int)ORC_MU0);
        blind_rotate_extract_impl(e2, c0, ORC_MU1, bk, bkn, 0);
        /* SampleExtract is linear: SE(t1+t0 (+mu on b[0])) == SE(t1)+SE(t0) (+mu on b) */
        for (unsigned i = 0; i < ORC_LVL1_WORDS; i++) e[i] = neg ? 0u - e[i] - e2[i] : e[i] + e2[i];
        e[ORC_N] += neg ? 0u - ORC_MU1 : ORC_MU1;
        orc_keyswitch((uint16_t *)out, e, ksk);
    } else {                 /* src/bootstrap_gpu.cu:706-780 */
        uint32_t *o = (uint32_t *)out;
        lincomb1(t1, 1, (const uint32_t *)in0, 1, (const uint32_t *)in1, 0u - ORC_MU1);
        orc_keyswitch(c0, t1, ksk);
        blind_rotate_extract_impl(e, c0, ORC_MU1, bk, bkn, 0);
        lincomb1(t1, -1, (const uint32_t *)in0, 1, (const uint32_t *)in2, 0u - ORC_MU1);
        orc_keyswitch(c0, t1, ksk);
        blind_rotate_extract_impl(e2, c0, ORC_MU1, bk, bkn, 0);
        for (unsigned i = 0; i < ORC_LVL1_WORDS; i++) o[i] = neg ? 0u - e[i] - e2[i] : e[i] + e2[i];
        o[ORC_N] += neg ? 0u - ORC_MU1 : ORC_MU1;
    }
}
void orc_gate(int op, int level, void *out, const void *in0, const void *in1,
              const void *in2, const uint32_t *bk, const uint16_t *ksk)
{
    gate_impl(op, level, out, in0, in1, in2, bk, bk_ntt_get(bk), ksk);
}
void orc_gate_batch(int op, const uint8_t *ops, int level, void *out, const void *in0,
                    const void *in1, const void *in2, size_t count,
                    const uint32_t *bk, const uint16_t *ksk, int threads)
{
    const uint64_t *bkn = bk_ntt_get(bk);
    size_t stride = level == 0 ? ORC_LVL0_WORDS * 2 : ORC_LVL1_WORDS * 4;
    (void)threads;
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads);
#pragma omp parallel for schedule(dynamic, 1)
#endif
    for (size_t g = 0; g < count; g++) {
        int o = ops ? ops[g] : op;
        gate_impl(o, level, (char *)out + g * stride, (const char *)in0 + g * stride,
                  in1 ? (const char *)in1 + g * stride : NULL,
                  in2 ? (const char *)in2 + g * stride : NULL, bk, bkn, ksk);
    }
}

int orc_truth(int op, int a, int b, int c)
{
    switch (op) {   /* test/plain.h:10-69 */
    case ORC_NAND: return 1 - a * b;
    case ORC_OR: return (a + b) > 0;
    case ORC_ORYN: return (a + (1 - b)) > 0;
    case ORC_ORNY: return ((1 - a) + b) > 0;
    case ORC_AND: return a * b;
    case ORC_ANDYN: return a * (1 - b);
    case ORC_ANDNY: return (1 - a) * b;
    case ORC_NOR: return (a + b) == 0;
    case ORC_XOR: return (a + b) & 1;
    case ORC_XNOR: return (~(a ^ b)) & 1;
    case ORC_MUX: return a ? b : c;
    case ORC_NMUX: return (~(a ? b : c)) & 1;
    case ORC_NOT: return (~a) & 1;
    case ORC_COPY: return a;
    }
    return -1;
}

/* ------------------------------------------------- keys and ciphertexts --- */
#define SIGMA0_Q16 397333ull          /* lvl0param::alpha * 2^16 * 2^16 (128bit.hpp:16) */
#define SIGMA1_Q16 (128ull << 16)     /* lvl1param::alpha = 2^-25 -> 128 LSB (128bit.hpp:58) */
static inline uint64_t stream_seed(uint64_t seed, uint64_t a, uint64_t b)
{
    uint64_t s = seed ^ (a * 0xD6E8FEB86659FD93ull) ^ (b * 0xA0761D6478BD642Full) ^ 0x1234567ull;
    splitmix64(&s);
    return s;
}
void orc_keygen(uint64_t seed, int32_t *sk0, int32_t *sk1, uint32_t *bk, uint16_t *ksk)
{
    uint64_t s = stream_seed(seed, 1, 0);
    for (unsigned i = 0; i < ORC_n; i++) sk0[i] = (int32_t)(splitmix64(&s) & 1);         /* binary  (128bit.hpp:8-9)  */
    for (unsigned i = 0; i < ORC_N; i++) sk1[i] = (int32_t)(splitmix64(&s) % 3) - 1;     /* ternary (128bit.hpp:43-44) */
    g_init();
    /* bkgen (TFHEpp evalkeygens.hpp:12-45) -> trgswSymEncrypt = 4 x trlweSymEncryptZero
     * (trlwe.hpp:9-20) + gadget h[i] = 2^(32-8(i+1)) (trgsw.hpp:689-765) */
#ifdef _OPENMP
#pragma omp parallel for schedule(static)
#endif
    for (unsigned i = 0; i < ORC_n; i++) {
        uint32_t as[ORC_N];
        for (int row = 0; row < 4; row++) {
            uint64_t st = stream_seed(seed, 2 + i, row);
            uint32_t *a = bk + (((size_t)i * 4 + row) * 2 + 0) * ORC_N;
            uint32_t *b = bk + (((size_t)i * 4 + row) * 2 + 1) * ORC_N;
            for (unsigned c = 0; c < ORC_N; c++) a[c] = (uint32_t)splitmix64(&st);
            orc_polymul_ntt(as, a, (const uint32_t *)sk1);
            for (unsigned c = 0; c < ORC_N; c++) b[c] = as[c] + (uint32_t)gauss_fx(&st, SIGMA1_Q16);
            uint32_t h = 1u << (32 - ((row & 1) + 1) * ORC_BGBIT);
            if (sk0[i] == 1) { if (row < 2) a[0] += h; else b[0] += h; }
        }
    }
    /* ikskgen (TFHEpp evalkeygens.hpp:188-203) */
#ifdef _OPENMP
#pragma omp parallel for schedule(static)
#endif
    for (unsigned i = 0; i < ORC_N; i++)
        for (int j = 0; j < ORC_T; j++)
            for (int k = 0; k < 2; k++) {
                uint64_t st = stream_seed(seed, 100000 + i, j * 2 + k);
                uint16_t *ct = ksk + (((size_t)i * ORC_T + j) * 2 + k) * ORC_LVL0_WORDS;
                uint16_t m = (uint16_t)(sk1[i] * (k + 1) * (1 << (16 - (j + 1) * ORC_BASEBIT)));
                uint16_t b = (uint16_t)(m + (uint16_t)gauss_fx(&st, SIGMA0_Q16));
                for (unsigned c = 0; c < ORC_n; c++) {
                    ct[c] = (uint16_t)splitmix64(&st);
                    b = (uint16_t)(b + ct[c] * (uint16_t)sk0[c]);
                }
                ct[ORC_n] = b;
            }
}
/* tlweSymEncrypt (TFHEpp tlwe.hpp:10-21); message = bit ? mu : -mu (test/test_util.h:23-27) */
void orc_encrypt_lvl0(uint64_t seed, uint64_t index, int bit, const int32_t *sk0, uint16_t *ct)
{
    uint64_t st = stream_seed(seed, 0x10000000ull + index, 7);
    uint16_t b = (uint16_t)((bit ? ORC_MU0 : 0u - ORC_MU0) + (uint16_t)gauss_fx(&st, SIGMA0_Q16));
    for (unsigned c = 0; c < ORC_n; c++) {
        ct[c] = (uint16_t)splitmix64(&st);
        b = (uint16_t)(b + ct[c] * (uint16_t)sk0[c]);
    }
    ct[ORC_n] = b;
}
void orc_encrypt_lvl1(uint64_t seed, uint64_t index, int bit, const int32_t *sk1, uint32_t *ct)
{
    uint64_t st = stream_seed(seed, 0x20000000ull + index, 9);
    uint32_t b = (bit ? ORC_MU1 : 0u - ORC_MU1) + (uint32_t)gauss_fx(&st, SIGMA1_Q16);
    for (unsigned c = 0; c < ORC_N; c++) {
        ct[c] = (uint32_t)splitmix64(&st);
        b += ct[c] * (uint32_t)sk1[c];
    }
    ct[ORC_N] = b;
}
/* tlweSymDecrypt (TFHEpp tlwe.hpp:86-103): signed(phase) > 0 */
int orc_decrypt_lvl0(const uint16_t *ct, const int32_t *sk0)
{
    uint16_t ph = ct[ORC_n];
    for (unsigned c = 0; c < ORC_n; c++) ph = (uint16_t)(ph - ct[c] * (uint16_t)sk0[c]);
    return (int16_t)ph > 0;
}
int orc_decrypt_lvl1(const uint32_t *ct, const int32_t *sk1)
{
    uint32_t ph = ct[ORC_N];
    for (unsigned c = 0; c < ORC_N; c++) ph -= ct[c] * (uint32_t)sk1[c];
    return (int32_t)ph > 0;
}
