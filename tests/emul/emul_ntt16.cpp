// CPU emulation of the 64-thread x 16-coefficient team dataflow (cufhe_b200/csrc/ntt16.cuh)
// used by k_blind_rotate16: same host+device pass functions, same padded exchange slots,
// thread by thread, checked against the oracle.  Also checks that every exchange
// store/load is shared-memory bank-conflict free.  Test infrastructure only.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include "../../cufhe_b200/csrc/ntt16.cuh"
#include "../../cufhe_b200/csrc/tfhe_core.cuh"
#include "../../oracle/cufhe_oracle.h"

using namespace b2;
static NttTables T;
static Lane16Tables L;
static int g_conflicts = 0;

typedef uint32_t Team[64][16];

// one exchange: every thread stores x[k] at pad(eFrom(T,k)), then loads pad(eTo(T,k))
template <class EF, class ET, class PAD>
static void exchange(Team x, EF eFrom, ET eTo, PAD pad)
{
    static uint32_t buf[kXbufWords];
    for (int k = 0; k < 16; k++)
        for (int w = 0; w < 2; w++) {
            unsigned banksW = 0, banksR = 0;
            for (int l = 0; l < 32; l++) {
                const int t = 32 * w + l;
                const int sw = pad(eFrom(t, k)), sr = pad(eTo(t, k));
                if (sw >= kXbufWords || sr >= kXbufWords) { g_conflicts += 1000; continue; }
                if (banksW & (1u << (sw & 31))) g_conflicts++;
                if (banksR & (1u << (sr & 31))) g_conflicts++;
                banksW |= 1u << (sw & 31);
                banksR |= 1u << (sr & 31);
            }
        }
    for (int t = 0; t < 64; t++)
        for (int k = 0; k < 16; k++) buf[pad(eFrom(t, k))] = x[t][k];
    for (int t = 0; t < 64; t++)
        for (int k = 0; k < 16; k++) x[t][k] = buf[pad(eTo(t, k))];
}

static void fwd(Team x, int q)   // layout A in, layout C out
{
    const uint32_t p = kP[q];
    for (int t = 0; t < 64; t++) pass16_fwd<4, 8>(x[t], T.fw1[q], 1, p);
    exchange(x, eA, eB, padAB);
    for (int t = 0; t < 64; t++) pass16_fwd<4, 8>(x[t], &L.fw2[q][0][t >> 2], 16, p);
    exchange(x, eB, eC, padC);
    for (int t = 0; t < 64; t++) pass16_fwd3(x[t], &L.fw3[q][0][t], 64, p);
}
static void inv(Team x, int q)   // layout C in, layout A out (scaled, [0, 2p))
{
    const uint32_t p = kP[q];
    for (int t = 0; t < 64; t++) pass16_inv<true>(x[t], T.iv1[q], 1, p);
    exchange(x, eC, eD, padC);
    for (int t = 0; t < 64; t++) pass16_inv<false>(x[t], &L.iv2[q][0][t & 15], 16, p);
    exchange(x, eD, eA, padDA);
    for (int t = 0; t < 64; t++) {
        pass16_inv3(x[t], &L.iv3[q][0][t], 64, p);
        for (int k = 0; k < 16; k++) x[t][k] = shoup_mul_p(x[t][k], L.scale[q][k][t].w, L.scale[q][k][t].ws, p);
    }
}
static void load_A(Team x, const int32_t* a, int q)
{
    for (int t = 0; t < 64; t++)
        for (int k = 0; k < 16; k++) x[t][k] = (uint32_t)(a[eA(t, k)] + (int32_t)kP[q]);
}
static void key_to_ntt(uint32_t* dst, const uint32_t* src, int q)
{
    Team x;
    for (int t = 0; t < 64; t++)
        for (int k = 0; k < 16; k++) x[t][k] = src[eA(t, k)] % kP[q];
    fwd(x, q);
    for (int t = 0; t < 64; t++)
        for (int k = 0; k < 16; k++) {
            const int32_t m = umontmul_p(x[t][k], T.r2[q], kP[q], q ? Prime<1>::pinv : Prime<0>::pinv);
            dst[bk16_slot(t, k)] = (uint32_t)(m + ((m >> 31) & (int32_t)kP[q]));
        }
}
static int32_t mm(uint32_t a, uint32_t b, int q) { return umontmul_p(a, b, kP[q], q ? Prime<1>::pinv : Prime<0>::pinv); }

static int check_polymul(std::mt19937_64& rng, int signed_b)
{
    uint32_t a[kN], b[kN], ref[kN], out[kN];
    for (int i = 0; i < kN; i++) {
        a[i] = (uint32_t)rng();
        b[i] = signed_b ? (uint32_t)((int)(rng() & 255) - 128) : (uint32_t)(rng() & 255);
    }
    orc_polymul_schoolbook(ref, a, b);
    static uint32_t an[2][kN];
    Team x[2];
    int32_t bs[kN];
    for (int i = 0; i < kN; i++) bs[i] = (int32_t)b[i];
    for (int q = 0; q < 2; q++) {
        key_to_ntt(an[q], a, q);
        load_A(x[q], bs, q);
        fwd(x[q], q);
        for (int t = 0; t < 64; t++)
            for (int k = 0; k < 16; k++) x[q][t][k] = (uint32_t)(mm(x[q][t][k], an[q][bk16_slot(t, k)], q) + (int32_t)kP[q]);
        inv(x[q], q);
    }
    for (int t = 0; t < 64; t++)
        for (int k = 0; k < 16; k++) out[eA(t, k)] = crt_lift32(x[0][t][k], x[1][t][k], T.crt_p0inv.w, T.crt_p0inv.ws);
    return memcmp(out, ref, sizeof ref) != 0;
}

static int check_cmux(std::mt19937_64& rng)
{
    static uint32_t bk[8 * kN], bkn[2][8 * kN], acc[2 * kN], ref[2 * kN];
    for (auto& v : bk) v = (uint32_t)rng();
    for (auto& v : acc) v = (uint32_t)rng();
    uint32_t abar = (uint32_t)(rng() % 2048);
    memcpy(ref, acc, sizeof acc);
    orc_cmux_step(ref, abar, bk, 0);
    for (int pl = 0; pl < 8; pl++)
        for (int q = 0; q < 2; q++) key_to_ntt(bkn[q] + pl * kN, bk + pl * kN, q);
    static int32_t sum[2][2][64][16];   // [prime][out][T][k]
    memset(sum, 0, sizeof sum);
    for (int j = 0; j < 2; j++)
        for (int d = 0; d < 2; d++) {
            int32_t dig[kN];
            for (uint32_t c = 0; c < (uint32_t)kN; c++) {
                uint32_t rot = acc[j * kN + ((c - abar) & (kN - 1))];
                if (rot_neg(c, abar)) rot = 0u - rot;
                dig[c] = digit(rot - acc[j * kN + c] + kDecompAdd, d);
            }
            const int row = 2 * j + d;
            for (int q = 0; q < 2; q++) {
                Team x;
                load_A(x, dig, q);
                fwd(x, q);
                for (int o = 0; o < 2; o++)
                    for (int t = 0; t < 64; t++)
                        for (int k = 0; k < 16; k++)
                            sum[q][o][t][k] += mm(x[t][k], bkn[q][(2 * row + o) * kN + bk16_slot(t, k)], q);
            }
        }
    for (int o = 0; o < 2; o++) {
        Team y[2];
        for (int q = 0; q < 2; q++) {
            for (int t = 0; t < 64; t++)
                for (int k = 0; k < 16; k++) y[q][t][k] = (uint32_t)(sum[q][o][t][k] + 4 * (int32_t)kP[q]);
            inv(y[q], q);
        }
        for (int t = 0; t < 64; t++)
            for (int k = 0; k < 16; k++)
                acc[o * kN + eA(t, k)] += crt_lift32(y[0][t][k], y[1][t][k], T.crt_p0inv.w, T.crt_p0inv.ws);
    }
    return memcmp(acc, ref, sizeof ref) != 0;
}

// pass F1 from the gadget-digit look-up tables == pass F1 by butterflies, as residues, and within
// the lazy bound the later passes assume
static int check_digit_pass(std::mt19937_64& rng)
{
    int bad = 0;
    for (int q = 0; q < 2; q++) {
        const uint32_t p = kP[q];
        for (int it = 0; it < 2000; it++) {
            uint32_t d[16], a[16], b[16];
            for (int k = 0; k < 16; k++) {
                d[k] = it == 0 ? 0u : it == 1 ? 255u : (uint32_t)(rng() & 0xFF);
                a[k] = d[k] + p - 128u;
                b[k] = d[k];
            }
            pass16_fwd<4, 8>(a, T.fw1[q], 1, p);
            pass16_fwd_digits(b, L.lut1[q], L.lut24[q], L.lut35[q], T.fw1[q], p);
            for (int k = 0; k < 16; k++)
                if (a[k] % p != b[k] % p || b[k] >= 8u * p) bad++;
        }
    }
    return bad;
}

// row-pair / four-row key multiply (umontmul2_p, umontmul4_p: one Montgomery reduction for the
// 64-bit sum) == the sum of the single products, as residues, and inside the stated bounds
static int check_pair_mac(std::mt19937_64& rng)
{
    int bad = 0;
    for (int it = 0; it < 200000; it++) {
        const int q = it & 1;
        const uint32_t p = kP[q], pinv = q ? Prime<1>::pinv : Prime<0>::pinv;
        uint32_t a[4], b[4];
        for (int i = 0; i < 4; i++) {
            a[i] = it < 16 ? 0xFFFFFFFFu : (uint32_t)rng();              // any lazy 32-bit residue
            b[i] = it < 16 ? p - 1 : (uint32_t)(rng() % p);             // key word in Montgomery form
        }
        int64_t one = 0, two = 0;
        for (int i = 0; i < 4; i++) {
            one += umontmul_p(a[i], b[i], p, pinv);
            if (i < 2) two += umontmul_p(a[i], b[i], p, pinv);
        }
        const int32_t r2 = (int32_t)umontmul2_p(a[0], b[0], a[1], b[1], p, pinv);
        const int32_t r4 = (int32_t)umontmul4_p(a[0], b[0], a[1], b[1], a[2], b[2], a[3], b[3], p, pinv);
        if (((int64_t)r2 - two) % (int64_t)p != 0 || r2 <= -(int32_t)p || r2 >= 2 * (int32_t)p) bad++;
        if (((int64_t)r4 - one) % (int64_t)p != 0 || r4 <= -(int32_t)p || r4 >= 4 * (int32_t)p) bad++;
    }
    return bad;
}

int main(int argc, char** argv)
{
    make_tables(T);
    make_tables16(L);
    std::mt19937_64 rng(20261104);
    int trials = argc > 1 ? atoi(argv[1]) : 20;
    const int bad0 = 0;
    const int bad3 = check_digit_pass(rng);
    printf("digit look-up pass mismatches: %d\n", bad3);
    int bad = 0, bad2 = 0;
    for (int i = 0; i < trials; i++) bad += check_polymul(rng, i & 1);
    printf("polymul mismatches: %d / %d\n", bad, trials);
    for (int i = 0; i < trials; i++) bad2 += check_cmux(rng);
    printf("cmux mismatches: %d / %d\n", bad2, trials);
    const int bad4 = check_pair_mac(rng);
    printf("pair / four-row multiply mismatches: %d\n", bad4);
    printf("bank conflicts: %d\n", g_conflicts);
    return (bad0 || bad || bad2 || bad3 || bad4 || g_conflicts) ? 1 : 0;
}
