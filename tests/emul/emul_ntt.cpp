// CPU emulation of the warp-level NTT / CMUX dataflow used by the CUDA kernels.
// Runs the SAME host+device functions (cufhe_b200/csrc/*.cuh) lane by lane, with the
// same swizzled exchange indices, and checks them against the oracle.  Test
// infrastructure only (links oracle/liborc.so).
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <vector>
#include "../../cufhe_b200/csrc/ntt_core.cuh"
#include "../../cufhe_b200/csrc/tfhe_core.cuh"
#include "../../oracle/cufhe_oracle.h"

using namespace b2;
static NttTables T;

// one emulated warp: x[lane][k]
typedef uint32_t Warp[32][32];

static void exchange(Warp x)
{
    static uint32_t buf[1024];
    for (int lane = 0; lane < 32; lane++)
        for (int v = 0; v < 8; v++)
            for (int e = 0; e < 4; e++) buf[xchg_wr_word(lane, v) + e] = x[lane][4 * v + e];
    for (int lane = 0; lane < 32; lane++)
        for (int j = 0; j < 32; j++) x[lane][j] = buf[xchg_rd_word(lane, j)];
}
template <int Q> static void fwd(Warp x)   // layout A in, layout B out
{
    for (int lane = 0; lane < 32; lane++) pass_fwd<Q>(x[lane], T.fw1[Q], 1);
    exchange(x);
    for (int lane = 0; lane < 32; lane++) pass_fwd<Q>(x[lane], &T.fw2[Q][0][lane], 32);
}
template <int Q> static void inv(Warp x)   // layout B in, layout A out (scaled)
{
    for (int lane = 0; lane < 32; lane++) pass_inv<Q>(x[lane], T.iv1[Q], 1);
    exchange(x);
    for (int lane = 0; lane < 32; lane++) {
        pass_inv<Q>(x[lane], &T.iv2[Q][0][lane], 32);
        for (int k = 0; k < 32; k++) x[lane][k] = shoup_mul<Q>(x[lane][k], T.scale[Q][k][lane].w, T.scale[Q][k][lane].ws);
    }
}
template <int Q> static void load_A(Warp x, const int32_t* a)
{
    for (int lane = 0; lane < 32; lane++)
        for (int k = 0; k < 32; k++) x[lane][k] = (uint32_t)(a[lane + 32 * k] + (int32_t)Prime<Q>::p);
}
// NTT-domain key polynomial in Montgomery form, stored at bk_slot()
template <int Q> static void key_to_ntt(uint32_t* dst, const uint32_t* src)
{
    Warp x;
    for (int lane = 0; lane < 32; lane++)
        for (int k = 0; k < 32; k++) x[lane][k] = src[lane + 32 * k] % Prime<Q>::p;
    fwd<Q>(x);
    for (int lane = 0; lane < 32; lane++)
        for (int k = 0; k < 32; k++) dst[bk_slot(lane, k)] = canon<Q>(umontmul<Q>(x[lane][k], T.r2[Q]));
}

static int check_polymul(std::mt19937_64& rng, int signed_b)
{
    uint32_t a[kN], b[kN], ref[kN], out[kN];
    for (int i = 0; i < kN; i++) {
        a[i] = (uint32_t)rng();
        b[i] = signed_b ? (uint32_t)((int)(rng() & 255) - 128) : (uint32_t)(rng() & 255);
    }
    orc_polymul_schoolbook(ref, a, b);
    static uint32_t an[2][kN];
    key_to_ntt<0>(an[0], a);
    key_to_ntt<1>(an[1], a);
    Warp x0, x1;
    int32_t bs[kN];
    for (int i = 0; i < kN; i++) bs[i] = (int32_t)b[i];
    load_A<0>(x0, bs); fwd<0>(x0);
    load_A<1>(x1, bs); fwd<1>(x1);
    for (int lane = 0; lane < 32; lane++)
        for (int k = 0; k < 32; k++) {
            x0[lane][k] = (uint32_t)(umontmul<0>(x0[lane][k], an[0][bk_slot(lane, k)]) + (int32_t)Prime<0>::p);
            x1[lane][k] = (uint32_t)(umontmul<1>(x1[lane][k], an[1][bk_slot(lane, k)]) + (int32_t)Prime<1>::p);
        }
    inv<0>(x0); inv<1>(x1);
    for (int lane = 0; lane < 32; lane++)
        for (int k = 0; k < 32; k++) out[lane + 32 * k] = crt_lift32(x0[lane][k], x1[lane][k], T.crt_p0inv.w, T.crt_p0inv.ws);
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
    for (int pl = 0; pl < 8; pl++) {
        key_to_ntt<0>(bkn[0] + pl * kN, bk + pl * kN);
        key_to_ntt<1>(bkn[1] + pl * kN, bk + pl * kN);
    }
    static int32_t sum[2][2][32][32];   // [prime][out][lane][k]
    memset(sum, 0, sizeof sum);
    for (int j = 0; j < 2; j++)
        for (int d = 0; d < 2; d++) {
            int32_t dig[kN];
            for (uint32_t c = 0; c < (uint32_t)kN; c++) {
                uint32_t rot = acc[j * kN + ((c - abar) & (kN - 1))];
                if (rot_neg(c, abar)) rot = 0u - rot;
                dig[c] = digit(rot - acc[j * kN + c] + kDecompAdd, d);
            }
            Warp x0, x1;
            load_A<0>(x0, dig); fwd<0>(x0);
            load_A<1>(x1, dig); fwd<1>(x1);
            const int row = 2 * j + d;
            for (int o = 0; o < 2; o++)
                for (int lane = 0; lane < 32; lane++)
                    for (int k = 0; k < 32; k++) {
                        sum[0][o][lane][k] += umontmul<0>(x0[lane][k], bkn[0][(2 * row + o) * kN + bk_slot(lane, k)]);
                        sum[1][o][lane][k] += umontmul<1>(x1[lane][k], bkn[1][(2 * row + o) * kN + bk_slot(lane, k)]);
                    }
        }
    for (int o = 0; o < 2; o++) {
        Warp y0, y1;
        for (int lane = 0; lane < 32; lane++)
            for (int k = 0; k < 32; k++) {
                y0[lane][k] = (uint32_t)(sum[0][o][lane][k] + 4 * (int32_t)Prime<0>::p);
                y1[lane][k] = (uint32_t)(sum[1][o][lane][k] + 4 * (int32_t)Prime<1>::p);
            }
        inv<0>(y0); inv<1>(y1);
        for (int lane = 0; lane < 32; lane++)
            for (int k = 0; k < 32; k++)
                acc[o * kN + lane + 32 * k] += crt_lift32(y0[lane][k], y1[lane][k], T.crt_p0inv.w, T.crt_p0inv.ws);
    }
    return memcmp(acc, ref, sizeof ref) != 0;
}

int main(int argc, char** argv)
{
    make_tables(T);
    std::mt19937_64 rng(20261104);
    int trials = argc > 1 ? atoi(argv[1]) : 20, bad = 0;
    for (int i = 0; i < trials; i++) bad += check_polymul(rng, i & 1);
    printf("polymul mismatches: %d / %d\n", bad, trials);
    int bad2 = 0;
    for (int i = 0; i < trials; i++) bad2 += check_cmux(rng);
    printf("cmux mismatches: %d / %d\n", bad2, trials);
    return (bad || bad2) ? 1 : 0;
}
