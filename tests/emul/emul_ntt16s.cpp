// CPU emulation of the signed / Solinas 64-thread x 16-coefficient team dataflow
// (cufhe_b200/csrc/ntt16s.cuh) used by k_blind_rotate_s: the same host+device pass functions and
// exchange maps, thread by thread, checked against the oracle's exact CMUX step and the schoolbook
// product.  Also checks (a) that every exchange store/load is shared-memory bank-conflict free
// under the 32-bit / 128-bit access rules, (b) the lazy-residue bounds by interval propagation
// over the real butterfly structure, (c) the Solinas and IMAD forms of q*p are bit-identical.
// Test infrastructure only.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <vector>
#include "../../cufhe_b200/csrc/ntt16s.cuh"
#include "../../cufhe_b200/csrc/tfhe_core.cuh"
#include "../../oracle/cufhe_oracle.h"

using namespace b2;
static TeamSTables3 L;
static TeamSConst C;
static int g_conflicts = 0;
static int64_t g_maxabs_fwd = 0, g_maxabs_inv = 0;

typedef int32_t Team[64][16];

// ---- shared-memory access model ------------------------------------------------------------
// one instruction of a 64-thread team = two warps; `words` = 1 (32-bit: whole warp at once),
// 4 (128-bit: quarter-warps of 8 lanes).  Lanes hitting the same address broadcast.
static void check_access(const int (&addr)[64], int words)
{
    const int group = words == 4 ? 8 : words == 2 ? 16 : 32;
    for (int base = 0; base < 64; base += group) {
        int owner[32];
        for (int b = 0; b < 32; b++) owner[b] = -1;
        for (int l = 0; l < group; l++) {
            const int a = addr[base + l];
            if (a < 0) continue;
            if (a % words != 0 || a + words > kXsWords) { g_conflicts += 1000; continue; }
            for (int w = 0; w < words; w++) {
                const int bank = (a + w) & 31;
                if (owner[bank] >= 0 && owner[bank] != a + w) g_conflicts++;
                owner[bank] = a + w;
            }
        }
    }
}
static void st128(int32_t* buf, const int (&addr)[64], const V4 (&val)[64])
{
    check_access(addr, 4);
    for (int t = 0; t < 64; t++) memcpy(buf + addr[t], &val[t], 16);
}
static void own_ld128(const int32_t* buf, Team x, int v, const int (&regs)[4])
{
    int addr[64];
    for (int t = 0; t < 64; t++) addr[t] = chunk_of(t) + 4 * v;
    check_access(addr, 4);
    for (int t = 0; t < 64; t++)
        for (int i = 0; i < 4; i++) x[t][regs[i]] = buf[addr[t] + i];
}

static void xchg_AD(Team x)
{
    static int32_t buf[kXsWords];
    for (int kh = 0; kh < 4; kh++) {
        int addr[64]; V4 val[64];
        for (int t = 0; t < 64; t++) { addr[t] = xs_AD_w(t, kh); val[t] = V4{x[t][4 * kh], x[t][4 * kh + 1], x[t][4 * kh + 2], x[t][4 * kh + 3]}; }
        st128(buf, addr, val);
    }
    for (int th = 0; th < 4; th++) { const int regs[4] = {th, th + 4, th + 8, th + 12}; own_ld128(buf, x, th, regs); }
}
static void xchg_DC(Team x)      // D -> C (forward) and C -> D (inverse): the same 16 x 16 transposes
{
    static int32_t buf[kXsWords];
    for (int k = 0; k < 16; k++) {
        int addr[64];
        for (int t = 0; t < 64; t++) addr[t] = xs_DC_w(t, k);
        check_access(addr, 1);
        for (int t = 0; t < 64; t++) buf[addr[t]] = x[t][k];
    }
    for (int v = 0; v < 4; v++) { const int regs[4] = {4 * v, 4 * v + 1, 4 * v + 2, 4 * v + 3}; own_ld128(buf, x, v, regs); }
}
static void xchg_DA(Team x)
{
    static int32_t buf[kXsWords];
    for (int kl = 0; kl < 4; kl++) {
        int addr[64]; V4 val[64];
        for (int t = 0; t < 64; t++) { addr[t] = xs_DA_w(t, kl); val[t] = V4{x[t][kl], x[t][kl + 4], x[t][kl + 8], x[t][kl + 12]}; }
        st128(buf, addr, val);
    }
    for (int h = 0; h < 4; h++) { const int regs[4] = {4 * h, 4 * h + 1, 4 * h + 2, 4 * h + 3}; own_ld128(buf, x, h, regs); }
}

static void track(const Team x, int64_t& m)
{
    for (int t = 0; t < 64; t++)
        for (int k = 0; k < 16; k++) m = std::max<int64_t>(m, std::llabs((long long)x[t][k]));
}

static void fwd_tail(Team x, int q)   // after stages 0, 1 (layout A): layout C out
{
    const int32_t p = (int32_t)kP[q];
    int32_t dummy[16];
    xchg_AD(x);
    for (int t = 0; t < 64; t++) spass_fwd<1>(x[t], dummy, [&](int v) { return L.pr[q].f2[t >> 4][v]; }, p);
    track(x, g_maxabs_fwd);
    xchg_DC(x);
    for (int t = 0; t < 64; t++) spass_fwd<1>(x[t], dummy, [&](int v) { return L.pr[q].perT[kVecF3 + v][t]; }, p);
    track(x, g_maxabs_fwd);
}
static void fwd_digits(Team x, int q)   // raw digits in [0, 256), layout A
{
    for (int t = 0; t < 64; t++) {
        for (int k = 0; k < 4; k++) x[t][k] -= 128;          // slot k holds the signed digit, the others the raw byte (table handle)
        spass_digits(x[t], [&](int32_t d) { return L.pr[q].lut0[d]; }, [&](int32_t d) { return L.pr[q].lut12[0][d]; },
                     [&](int32_t d) { return L.pr[q].lut12[1][d]; });
    }
    fwd_tail(x, q);
}
static void fwd_general(Team x, int q)  // centred residues, layout A
{
    for (int t = 0; t < 64; t++) spass_fwd01(x[t], C.fw01[q][0], C.fw01[q][1], C.fw01[q][2], (int32_t)kP[q]);
    fwd_tail(x, q);
}
static void inv(Team x, int q)   // layout C in, layout A out, [-p/4, 5p/4)
{
    const int32_t p = (int32_t)kP[q];
    int32_t dummy[16];
    for (int t = 0; t < 64; t++) spass_inv1<1>(x[t], dummy, [&](int i) { return C.iv1[q][i]; }, p);
    track(x, g_maxabs_inv);
    xchg_DC(x);
    for (int t = 0; t < 64; t++) spass_inv2<1>(x[t], dummy, [&](int v) { return L.pr[q].i2[v][t & 15]; }, p);
    track(x, g_maxabs_inv);
    xchg_DA(x);
    for (int t = 0; t < 64; t++)
        spass_inv3<1>(x[t], dummy, [&](int v) { return L.pr[q].perT[kVecI3 + v][t]; }, [&](int v) { return L.pr[q].perT[kVecSc + v][t]; }, p);
}

// key polynomial -> NTT domain, Montgomery form in [0, p), bk16_slot order
static void key_to_ntt(uint32_t* dst, const uint32_t* src, int q)
{
    Team x;
    const uint32_t p = kP[q];
    for (int t = 0; t < 64; t++)
        for (int k = 0; k < 16; k++) x[t][k] = h_centre(src[eA(t, k)] % p, p);
    fwd_general(x, q);
    for (int t = 0; t < 64; t++)
        for (int k = 0; k < 16; k++) {
            int32_t m = smontmul1(x[t][k], (uint32_t)C.r2[q] + (C.r2[q] < 0 ? p : 0), (int32_t)p, kPinv[q]);
            m %= (int32_t)p;
            dst[bk16_slot(t, k)] = (uint32_t)(m < 0 ? m + (int32_t)p : m);
        }
}

static int check_polymul(std::mt19937_64& rng)
{
    uint32_t a[kN], b[kN], ref[kN], out[kN];
    for (int i = 0; i < kN; i++) { a[i] = (uint32_t)rng(); b[i] = (uint32_t)((int)(rng() & 255) - 128); }
    orc_polymul_schoolbook(ref, a, b);
    static uint32_t an[2][kN];
    Team x[2];
    for (int q = 0; q < 2; q++) {
        key_to_ntt(an[q], a, q);
        for (int t = 0; t < 64; t++)
            for (int k = 0; k < 16; k++) x[q][t][k] = (int32_t)b[eA(t, k)] + 128;        // raw digit
        fwd_digits(x[q], q);
        for (int t = 0; t < 64; t++)
            for (int k = 0; k < 16; k++)
                x[q][t][k] = smontmul1(x[q][t][k], an[q][bk16_slot(t, k)], (int32_t)kP[q], kPinv[q]);
        inv(x[q], q);
    }
    for (int t = 0; t < 64; t++)
        for (int k = 0; k < 16; k++) out[eA(t, k)] = scrt_lift32(x[0][t][k], x[1][t][k], C.crt.w, C.crt.ws);
    return memcmp(out, ref, sizeof ref) != 0;
}

// the exact product of the microbenchmark entry point (k_polymul_s<3>): a = any u32, b = any int32
static int check_polymul3(std::mt19937_64& rng, int mode)
{
    static uint32_t a[kN], b[kN], ref[kN], out[kN];
    for (int i = 0; i < kN; i++) {
        a[i] = mode == 2 ? 0xFFFFFFFFu : (uint32_t)rng();
        b[i] = mode == 0 ? (uint32_t)(rng() & 0x3FFFF) : mode == 2 ? 0x80000000u : (uint32_t)rng();      // 18-bit (the reference test's domain) / any
    }
    orc_polymul_schoolbook(ref, a, b);
    static Team x[3], y[3];
    for (int q = 0; q < 3; q++) {
        const int32_t p = (int32_t)kP[q];
        for (int t = 0; t < 64; t++)
            for (int k = 0; k < 16; k++) {
                const uint32_t v = a[eA(t, k)];
                const int32_t w = (int32_t)b[eA(t, k)];
                x[q][t][k] = (int32_t)(v - (v >> 27) * kP[q]);                                  // lazy_u32: [0, 2p)
                y[q][t][k] = (int32_t)((uint32_t)w - (uint32_t)(w >> 27) * kP[q]);              // lazy_s32
            }
        fwd_general(x[q], q);
        fwd_general(y[q], q);
        for (int t = 0; t < 64; t++)
            for (int k = 0; k < 16; k++) {
                const int32_t z = smontmul1(x[q][t][k], (uint32_t)sreduce(y[q][t][k], p), p, kPinv[q]);
                x[q][t][k] = smontmul1(z, (uint32_t)C.r2[q], p, kPinv[q]);
            }
        track(x[q], g_maxabs_inv);
        inv(x[q], q);
    }
    for (int t = 0; t < 64; t++)
        for (int k = 0; k < 16; k++) out[eA(t, k)] = scrt3_lift32(x[0][t][k], x[1][t][k], x[2][t][k], C.crt, C.crt02, C.crt12, C.p01);
    return memcmp(out, ref, sizeof ref) != 0;
}

// one CMUX step exactly as k_blind_rotate_s runs it: digits of acc_j transformed in pairs, row pairs
// share one Montgomery reduction, both outputs inverse-transformed, CRT, accumulate
static int check_cmux(std::mt19937_64& rng, int worst)
{
    static uint32_t bk[8 * kN], bkn[2][8 * kN], acc[2 * kN], ref[2 * kN];
    for (auto& v : bk) v = worst ? 0xFFFFFFFFu - (uint32_t)(rng() & 1) : (uint32_t)rng();
    for (auto& v : acc) v = (uint32_t)rng();
    uint32_t abar = (uint32_t)(rng() % 2048);
    memcpy(ref, acc, sizeof acc);
    orc_cmux_step(ref, abar, bk, 0);
    for (int pl = 0; pl < 8; pl++)
        for (int q = 0; q < 2; q++) key_to_ntt(bkn[q] + pl * kN, bk + pl * kN, q);
    static Team s[2][2];     // [prime][out]
    for (int j = 0; j < 2; j++) {
        int32_t dig[2][kN];
        for (uint32_t c = 0; c < (uint32_t)kN; c++) {
            uint32_t rot = acc[j * kN + ((c - abar) & (kN - 1))];
            if (rot_neg(c, abar)) rot = 0u - rot;
            const uint32_t t = rot - acc[j * kN + c] + kDecompAdd;
            dig[0][c] = (int32_t)(t >> 24);
            dig[1][c] = (int32_t)((t >> 16) & 0xFF);
        }
        for (int q = 0; q < 2; q++) {
            Team x0, x1;
            for (int t = 0; t < 64; t++)
                for (int k = 0; k < 16; k++) { x0[t][k] = dig[0][eA(t, k)]; x1[t][k] = dig[1][eA(t, k)]; }
            fwd_digits(x0, q);
            fwd_digits(x1, q);
            const uint32_t pinv = kPinv[q];
            for (int o = 0; o < 2; o++)
                for (int t = 0; t < 64; t++)
                    for (int k = 0; k < 16; k++) {
                        const int32_t r = smontmul2(x0[t][k], bkn[q][(2 * (2 * j) + o) * kN + bk16_slot(t, k)], x1[t][k],
                                                    bkn[q][(2 * (2 * j + 1) + o) * kN + bk16_slot(t, k)], (int32_t)kP[q], pinv);
                        s[q][o][t][k] = j ? s[q][o][t][k] + r : r;
                    }
        }
    }
    for (int o = 0; o < 2; o++) {
        for (int q = 0; q < 2; q++) { track(s[q][o], g_maxabs_inv); inv(s[q][o], q); }
        for (int t = 0; t < 64; t++)
            for (int k = 0; k < 16; k++) acc[o * kN + eA(t, k)] += scrt_lift32(s[0][o][t][k], s[1][o][t][k], C.crt.w, C.crt.ws);
    }
    return memcmp(acc, ref, sizeof ref) != 0;
}

// ---- analytic bounds: interval propagation over the real butterfly structure (units of p) ----
struct Iv { double lo, hi; };
static double iv_max(const std::vector<Iv>& v)
{
    double m = 0;
    for (auto& i : v) m = std::max(m, std::max(-i.lo, i.hi));
    return m;
}
static double bound_forward()
{
    std::vector<Iv> x(kN, Iv{-1.5 - 1e-6, 1.5 + 1e-6});            // after the table stages: 3p/2 + 128
    for (int st = 2; st < 10; st++) {
        const int t = 512 >> st;
        for (int e = 0; e < kN; e++)
            if ((e & t) == 0) {
                const Iv u = x[e];
                x[e] = Iv{u.lo - 0.25, u.hi + 1.25};
                x[e + t] = Iv{u.lo - 1.25, u.hi + 0.25};
            }
    }
    return iv_max(x);
}
static double bound_inverse(double b0, int q)
{
    const double rho = (67108864.0 + 16.0 * std::fabs(134217728.0 - (double)kP[q])) / (double)kP[q];   // |sreduce|
    std::vector<Iv> x(kN, Iv{-b0, b0});
    double m = b0;
    for (int st = 0; st < 10; st++) {
        const int t = 1 << st;
        for (int e = 0; e < kN; e++)
            if ((e & t) == 0) {
                const bool unit = st < 4 && (e & (t - 1)) == 0;        // pass I1 only: plain reduction
                const Iv u = x[e];
                if (unit) { x[e] = Iv{u.lo - rho, u.hi + rho}; x[e + t] = Iv{u.lo - rho, u.hi + rho}; }
                else { x[e] = Iv{u.lo - 0.25, u.hi + 1.25}; x[e + t] = Iv{u.lo - 1.25, u.hi + 0.25}; }
            }
        m = std::max(m, iv_max(x));
    }
    return m;
}

static int check_primitives(std::mt19937_64& rng)
{
    int bad = 0;
    for (int it = 0; it < 400000; it++) {
        const int q = it & 1;
        const int32_t p = (int32_t)kP[q];
        const Tws t = h_tws((uint32_t)(rng() % kP[q]), kP[q]);
        int32_t v = it < 8 ? (it & 2 ? INT32_MIN : INT32_MAX) : (int32_t)rng();
        const int32_t vr = v / 17 * 16;          // sreduce is specified for |v| < 2^31 - 2^26
        const int32_t r0 = sshoup(v, t.w, t.ws, p), r1 = sshoup_solinas(v, t.w, t.ws, kSolK[q]);
        const int64_t want = (((int64_t)v * t.w) % p + p) % p;
        if (r0 != r1) bad++;
        if (((int64_t)r0 % p + p) % p != want) bad++;
        if (4 * (int64_t)r0 < -(int64_t)p || 4 * (int64_t)r0 >= 5 * (int64_t)p) bad++;
        const int32_t rr = sreduce(vr, p);
        if (((int64_t)rr - vr) % p != 0 || std::llabs((long long)rr) > (1ll << 26) + 16ll * ((1 << kSolK[q]) - 1)) bad++;
        // pair product: |a| < 2^31, 0 <= b < p
        const int32_t a0 = (int32_t)rng(), a1 = (int32_t)rng();
        const uint32_t b0 = (uint32_t)(rng() % kP[q]), b1 = (uint32_t)(rng() % kP[q]);
        const int32_t m2 = smontmul2(a0, b0, a1, b1, p, kPinv[q]);
        const __int128 sum = (__int128)a0 * b0 + (__int128)a1 * b1;
        // m2 * 2^32 == sum (mod p)
        if ((int64_t)((((__int128)m2 << 32) - sum) % p) != 0) bad++;
        if (std::llabs((long long)m2) >= (long long)(std::llabs((long long)(sum >> 32)) + p / 2 + 2)) bad++;
    }
    return bad;
}

int main(int argc, char** argv)
{
    make_tables16s(L, C);
    std::mt19937_64 rng(20261105);
    const int trials = argc > 1 ? atoi(argv[1]) : 12;
    const int bad_p = check_primitives(rng);
    printf("primitive mismatches (SOL == IMAD form, residues, ranges): %d\n", bad_p);
    int bad_m = 0, bad_c = 0;
    for (int i = 0; i < trials; i++) bad_m += check_polymul(rng);
    printf("polymul mismatches: %d / %d\n", bad_m, trials);
    int bad_3 = 0;
    for (int i = 0; i < trials; i++) bad_3 += check_polymul3(rng, i % 3);
    printf("three-prime exact polymul mismatches: %d / %d\n", bad_3, trials);
    for (int i = 0; i < trials; i++) bad_c += check_cmux(rng, i % 4 == 3);
    printf("cmux mismatches: %d / %d\n", bad_c, trials);
    printf("bank conflicts: %d\n", g_conflicts);
    const double bf = bound_forward();
    // key products: two pair reductions, each |x0 b0 + x1 b1| / 2^32 + p/2 with |x| < bf p, b < p < 2^27
    const double b0 = 2.0 * (2.0 * bf / 32.0 + 0.5);
    const double bi0 = bound_inverse(b0, 0), bi1 = bound_inverse(b0, 1), bi2 = bound_inverse(b0, 2);
    const double lim0 = 2147483648.0 / kP[0], lim1 = 2147483648.0 / kP[1], lim2 = 2147483648.0 / kP[2];
    const double bg = 2.0 + 10 * 1.25;       // general inputs (stand-alone transforms): [0, 2p) + ten stages
    printf("bounds (units of p): forward %.3f (general inputs %.3f), key products %.3f, inverse %.3f / %.3f / %.3f, int32 limit %.4f / %.4f / %.4f\n", bf, bg, b0, bi0, bi1, bi2, lim0, lim1, lim2);
    printf("observed max |x| / p: forward %.3f, inverse %.3f\n", (double)g_maxabs_fwd / kP[1], (double)g_maxabs_inv / kP[1]);
    const bool bounds_ok = bf + 0.01 < lim0 && bi0 + 0.01 < lim0 && bi1 + 0.01 < lim1 && bi2 + 0.01 < lim2 && bg + 0.01 < lim2;
    return (bad_p || bad_m || bad_3 || bad_c || g_conflicts || !bounds_ok) ? 1 : 0;
}
