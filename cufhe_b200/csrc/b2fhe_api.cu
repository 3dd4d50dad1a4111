// b2fhe_api.cu -- C ABI (include/b2fhe.h) over the sm_100a kernels.
// Host-side counterpart of the reference's launch wrappers and key upload
// (src/bootstrap_gpu.cu:97-155,782-1292, src/keyswitch_gpu.cu, src/cufhe_gates_gpu.cu).
#include <atomic>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/b2fhe.h"
#include "k_blindrotate.cuh"
#include "k_blindrotate_s.cuh"
#include "k_blindrotate_cl4s.cuh"
#include "k_cmux.cuh"
#include "k_keyswitch.cuh"
#include "k_ntt_s.cuh"

using namespace b2;

namespace {

constexpr int kMaxDev = 16;
struct DevCtx {
    bool ready = false;
    bool owns_keys = false;
    uint32_t* bk = nullptr;         // NTT-domain key [636][4][2 primes][2][1024], Montgomery residues in [0, p), bk16_slot order
    uint32_t* ksk = nullptr;        // T[1024][7][3][320]
    TeamSTables3* tabs = nullptr;   // team transform tables, three primes (the blind rotation stages the first two)
    int sm_count = 0;
};
DevCtx g_ctx[kMaxDev];
std::mutex g_mu;
int g_gates_per_cta = 0;
int g_ks_variant = 0;            // 0 = by batch size, 1 = k_keyswitch_lat, 2 = k_keyswitch
int g_ks_lat_max_per_sm = 32;    // k_keyswitch_lat up to this many gates per SM (measured faster up to there)
int g_br_cluster_max = 70;       // k_blind_rotate_cl4s up to this many gates: 33 four-CTA clusters run one CTA per SM (1.43 ms), up to 70 two per SM (1.91 ms); beyond that k_blind_rotate_s with one gate per CTA (3.3 ms up to 148 gates) wins
int g_latency_mode = 1;          // small calls: 1 = latency kernels (one gate over several SMs), 0 = always the throughput kernel (b2fhe_set_latency_mode)
bool g_br_tail_split = true;     // last partial wave of k_blind_rotate_s as its own launch (b2fhe_set_br_variant(-1) / (-2) = off / on)
bool g_ntt_prefetch = true;     // stand-alone NTT / INTT: inputs prefetched by bulk copies; products: 5 / 3 per CTA (b2fhe_set_br_variant(-3) / (-4) = off / on)
int g_br_variant = 0;            // 0 = by batch size, 5 = k_blind_rotate_cl4s, 40 = k_blind_rotate_s, 6 / 41 = their schedule-fuzzing builds, 48 = the key-row-0 timing experiment
std::atomic<uint64_t> g_launches{0};
thread_local std::string t_err;

int fail(int code, const char* what, const char* file, int line)
{
    char buf[512];
    snprintf(buf, sizeof buf, "b2fhe: %s failed at %s:%d : %s", what, file, line,
             code >= 10000 ? "invalid use" : cudaGetErrorString((cudaError_t)code));
    t_err = buf;
    return code;
}
#define CK(expr)                                                         \
    do {                                                                 \
        cudaError_t e_ = (expr);                                         \
        if (e_ != cudaSuccess) return fail((int)e_, #expr, __FILE__, __LINE__); \
    } while (0)
#define REQUIRE(cond, code)                                              \
    do {                                                                 \
        if (!(cond)) return fail(code, #cond, __FILE__, __LINE__);       \
    } while (0)
#define LAUNCHED() do { g_launches.fetch_add(1, std::memory_order_relaxed); CK(cudaGetLastError()); } while (0)

template <int G> int set_brs_attr()
{
    CK(cudaFuncSetAttribute(k_blind_rotate_s<G>, cudaFuncAttributeMaxDynamicSharedMemorySize, brs_smem_bytes(G)));
    return 0;
}

int noalloc_setup(int device);   // b2fhe_noalloc.inl: device address of the fixed scratch array

// twiddle tables + kernel attributes on `device` (InitializeNTThandlers, src/bootstrap_gpu.cu:97-109)
int setup_device(int device)
{
    int ndev = 0;
    CK(cudaGetDeviceCount(&ndev));
    REQUIRE(device >= 0 && device < ndev && device < kMaxDev, B2FHE_E_NO_DEVICE);
    CK(cudaSetDevice(device));
    DevCtx& c = g_ctx[device];
    if (c.tabs) return 0;
    static TeamSTables3 LS;
    static TeamSConst CS;
    static bool made = false;
    if (!made) { make_tables16s(LS, CS); made = true; }
    CK(cudaMemcpyToSymbol(c_s, &CS, sizeof CS));
    CK(cudaMalloc(&c.tabs, sizeof(TeamSTables3)));
    CK(cudaMemcpy(c.tabs, &LS, sizeof LS, cudaMemcpyHostToDevice));
    CK(cudaDeviceGetAttribute(&c.sm_count, cudaDevAttrMultiProcessorCount, device));
    // keep stream-ordered scratch cached between calls
    cudaMemPool_t pool;
    CK(cudaDeviceGetDefaultMemPool(&pool, device));
    uint64_t keep = ~0ull;
    CK(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep));
    if (int r = set_brs_attr<1>()) return r;
    if (int r = set_brs_attr<2>()) return r;
    if (int r = set_brs_attr<3>()) return r;
    if (int r = set_brs_attr<4>()) return r;
    CK(cudaFuncSetAttribute(k_blind_rotate_s<4, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, brs_smem_bytes(4)));
    CK(cudaFuncSetAttribute(k_blind_rotate_cl4s<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kCl4SmemBytes));
    CK(cudaFuncSetAttribute(k_blind_rotate_cl4s<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kCl4SmemBytes));
    CK(cudaFuncSetAttribute(k_blind_rotate_s<1, false, ABL_JIT>, cudaFuncAttributeMaxDynamicSharedMemorySize, brs_smem_bytes(1)));
    CK(cudaFuncSetAttribute(k_blind_rotate_s<4, false, ABL_JIT>, cudaFuncAttributeMaxDynamicSharedMemorySize, brs_smem_bytes(4)));
    CK(cudaFuncSetAttribute(k_cmux_s<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, cmux_smem_bytes(1)));
    CK(cudaFuncSetAttribute(k_cmux_s<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, cmux_smem_bytes(4)));
    CK(cudaFuncSetAttribute(k_keyswitch_step<32, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 32 * kN * 2));
    CK(cudaFuncSetAttribute(k_keyswitch_step<32, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 32 * kN * 2));
    CK(cudaFuncSetAttribute(k_keyswitch_step<32, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 32 * kN * 2));
    CK(cudaFuncSetAttribute(k_keyswitch_step<32, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 32 * kN * 2));
    CK(cudaFuncSetAttribute(k_keyswitch_step<32, 16>, cudaFuncAttributeMaxDynamicSharedMemorySize, 32 * kN * 2));
    CK(cudaFuncSetAttribute(k_ntt_fwd_s<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, ntt_fwd_smem_bytes<kNttTeams>()));
    CK(cudaFuncSetAttribute(k_ntt_fwd_s<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, ntt_fwd_smem_bytes<kNttTeams>()));
    CK(cudaFuncSetAttribute(k_ntt_fwd_s<0, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, ntt_fwd_pf_smem_bytes<kNttTeams>()));
    CK(cudaFuncSetAttribute(k_ntt_inv_s<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, ntt_inv_smem_bytes<kNttTeams>()));
    CK(cudaFuncSetAttribute(k_ntt_inv_s<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, ntt_inv_pf_smem_bytes<kNttTeams>()));
    CK(cudaFuncSetAttribute(k_polymul_s<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, ntt_s_smem_bytes<2, 4>()));
    CK(cudaFuncSetAttribute(k_polymul_s<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, ntt_s_smem_bytes<3, 6>()));
    CK(cudaFuncSetAttribute(k_polymul_s<2, 5>, cudaFuncAttributeMaxDynamicSharedMemorySize, ntt_s_smem_bytes<2, 10>()));
    CK(cudaFuncSetAttribute(k_polymul_s<3, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, ntt_s_smem_bytes<3, 9>()));
    CK(cudaFuncSetAttribute(k_polymul_s<2, 5, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, polymul_pf_smem_bytes<2, 5>()));
    CK(cudaFuncSetAttribute(k_polymul_s<3, 3, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, polymul_pf_smem_bytes<3, 3>()));
    return noalloc_setup(device);
}

// persistent CTAs of the stand-alone transform kernels: `per_sm` per SM, fewer for small batches
int ntt_grid(const DevCtx& c, size_t npoly, int polys_per_cta, int per_sm)
{
    size_t blocks = (npoly + polys_per_cta - 1) / polys_per_cta;
    size_t cap = (size_t)c.sm_count * per_sm;
    return (int)(blocks < cap ? (blocks ? blocks : 1) : cap);
}

// BootstrappingKeyToNTT + KeySwitchingKeyToDevice on one device (src/bootstrap_gpu.cu:111-138,
// src/keyswitch_gpu.cu:6-16)
int convert_keys(int device, const uint32_t* bk_raw, const uint16_t* ksk_raw)
{
    DevCtx& c = g_ctx[device];
    CK(cudaSetDevice(device));
    uint32_t* d_raw = nullptr;
    uint16_t* d_kraw = nullptr;
    CK(cudaMalloc(&c.bk, B2FHE_BK_NTT_BYTES));
    CK(cudaMalloc(&c.ksk, B2FHE_KSK_DEV_BYTES));
    c.owns_keys = true;
    CK(cudaMalloc(&d_raw, B2FHE_BK_RAW_WORDS * 4));
    CK(cudaMalloc(&d_kraw, B2FHE_KSK_RAW_WORDS * 2));
    CK(cudaMemcpy(d_raw, bk_raw, B2FHE_BK_RAW_WORDS * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d_kraw, ksk_raw, B2FHE_KSK_RAW_WORDS * 2, cudaMemcpyHostToDevice));
    const int npoly = kLwe * kRows * 2;
    k_ntt_fwd_s<1><<<ntt_grid(c, npoly, kNttTeams / 2, 2), kNttThreadsS, ntt_fwd_smem_bytes<kNttTeams>()>>>(d_raw, c.bk, npoly, c.tabs->pr);
    LAUNCHED();
    k_ksk_layout<<<c.sm_count * 8, 256>>>(d_kraw, c.ksk);
    LAUNCHED();
    CK(cudaDeviceSynchronize());
    CK(cudaFree(d_raw));
    CK(cudaFree(d_kraw));
    c.ready = true;
    return 0;
}

int defer_flush_all();      // b2fhe_defer.inl
int defer_destroy_all();
void defer_mark_idle_all();

// replicate GPU 0's converted keys device-to-device (NVLink peer copies) instead of
// re-uploading and re-converting per GPU as the reference does
int replicate_keys(int ngpu)
{
    for (int d = 1; d < ngpu; d++) {
        DevCtx& c = g_ctx[d];
        CK(cudaSetDevice(d));
        CK(cudaMalloc(&c.bk, B2FHE_BK_NTT_BYTES));
        CK(cudaMalloc(&c.ksk, B2FHE_KSK_DEV_BYTES));
        c.owns_keys = true;
        CK(cudaMemcpyPeer(c.bk, d, g_ctx[0].bk, 0, B2FHE_BK_NTT_BYTES));
        CK(cudaMemcpyPeer(c.ksk, d, g_ctx[0].ksk, 0, B2FHE_KSK_DEV_BYTES));
        c.ready = true;
    }
    return 0;
}

int get_ctx(int device, DevCtx*& c)
{
    REQUIRE(device >= 0 && device < kMaxDev, B2FHE_E_NO_DEVICE);
    c = &g_ctx[device];
    REQUIRE(c->ready, B2FHE_E_NOT_INITIALIZED);
    CK(cudaSetDevice(device));
    return 0;
}

}  // namespace

extern "C" {

const char* b2fhe_last_error_string(void) { return t_err.c_str(); }
uint64_t b2fhe_kernel_launches(void) { return g_launches.load(); }
int b2fhe_set_gates_per_cta(int gates)
{
    REQUIRE(gates >= 0 && gates <= 8, B2FHE_E_BAD_ARGUMENT);
    g_gates_per_cta = gates;
    return 0;
}
int b2fhe_set_ks_variant(int variant)
{
    REQUIRE((variant >= 0 && variant <= 3) || (variant >= 10 && variant <= 21), B2FHE_E_BAD_ARGUMENT);
    g_ks_variant = variant;
    return 0;
}
int b2fhe_set_br_variant(int variant)
{
    if (variant == -1 || variant == -2) { g_br_tail_split = variant == -2; return 0; }
    if (variant == -3 || variant == -4) { g_ntt_prefetch = variant == -4; return 0; }
    REQUIRE(variant == 0 || variant == 5 || variant == 6 || variant == 40 || variant == 41 || variant == 48, B2FHE_E_BAD_ARGUMENT);
    g_br_variant = variant;
    return 0;
}
int b2fhe_set_latency_mode(int on)
{
    REQUIRE(on == 0 || on == 1, B2FHE_E_BAD_ARGUMENT);
    g_latency_mode = on;
    return 0;
}
int b2fhe_num_devices(void)
{
    int n = 0;
    for (int i = 0; i < kMaxDev; i++) n += g_ctx[i].ready;
    return n;
}

int b2fhe_setup_devices(int ngpu)
{
    std::lock_guard<std::mutex> lk(g_mu);
    REQUIRE(ngpu >= 1 && ngpu <= kMaxDev, B2FHE_E_BAD_ARGUMENT);
    for (int d = 0; d < ngpu; d++)
        if (int r = setup_device(d)) return r;
    return 0;
}

int b2fhe_init_device(int device, const uint32_t* bk_raw, const uint16_t* ksk_raw)
{
    std::lock_guard<std::mutex> lk(g_mu);
    REQUIRE(bk_raw && ksk_raw, B2FHE_E_BAD_ARGUMENT);
    if (int r = setup_device(device)) return r;
    REQUIRE(!g_ctx[device].ready, B2FHE_E_BAD_ARGUMENT);
    return convert_keys(device, bk_raw, ksk_raw);
}

int b2fhe_init(int ngpu, const uint32_t* bk_raw, const uint16_t* ksk_raw)
{
    std::lock_guard<std::mutex> lk(g_mu);
    REQUIRE(ngpu >= 1 && ngpu <= kMaxDev && bk_raw && ksk_raw, B2FHE_E_BAD_ARGUMENT);
    for (int d = 0; d < ngpu; d++) {
        if (int r = setup_device(d)) return r;
        REQUIRE(!g_ctx[d].ready, B2FHE_E_BAD_ARGUMENT);
    }
    if (int r = convert_keys(0, bk_raw, ksk_raw)) return r;
    return replicate_keys(ngpu);
}

int b2fhe_adopt_keys(int device, const void* bk_ntt_dev, const void* ksk_dev)
{
    std::lock_guard<std::mutex> lk(g_mu);
    REQUIRE(bk_ntt_dev && ksk_dev, B2FHE_E_BAD_ARGUMENT);
    if (int r = setup_device(device)) return r;
    DevCtx& c = g_ctx[device];
    REQUIRE(!c.ready, B2FHE_E_BAD_ARGUMENT);
    c.bk = (uint32_t*)bk_ntt_dev;
    c.ksk = (uint32_t*)ksk_dev;
    c.owns_keys = false;
    c.ready = true;
    return 0;
}

int b2fhe_key_pointers(int device, void** bk_ntt_dev, void** ksk_dev)
{
    DevCtx* c;
    if (int r = get_ctx(device, c)) return r;
    if (bk_ntt_dev) *bk_ntt_dev = c->bk;
    if (ksk_dev) *ksk_dev = c->ksk;
    return 0;
}

int b2fhe_cleanup(void)
{
    if (int r = defer_flush_all()) return r;
    std::lock_guard<std::mutex> lk(g_mu);
    for (int d = 0; d < kMaxDev; d++) {
        DevCtx& c = g_ctx[d];
        if (!c.tabs && !c.ready) continue;
        CK(cudaSetDevice(d));
        CK(cudaDeviceSynchronize());
        if (c.owns_keys) { CK(cudaFree(c.bk)); CK(cudaFree(c.ksk)); }
        if (c.tabs) CK(cudaFree(c.tabs));
        c = DevCtx();
    }
    defer_destroy_all();
    return 0;
}

int b2fhe_synchronize(void)
{
    if (int r = defer_flush_all()) return r;      // queued per-gate calls run now
    for (int d = 0; d < kMaxDev; d++)
        if (g_ctx[d].ready) { CK(cudaSetDevice(d)); CK(cudaDeviceSynchronize()); }
    defer_mark_idle_all();
    return 0;
}

}  // extern "C"

#include "b2fhe_gates.inl"
#include "b2fhe_defer.inl"
#include "b2fhe_wire.inl"
#include "b2fhe_noalloc.inl"
