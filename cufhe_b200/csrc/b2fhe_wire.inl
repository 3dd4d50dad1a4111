// b2fhe_wire.inl -- key / ciphertext wire formats, host code, included
// by b2fhe_api.cu.
//
// 1. TFHEpp's own serialisation: cereal `portable_binary` archives written by
//    `EvalKey::serialize` (TFHEpp/include/cloudkey.hpp:103-107) and of
//    `std::vector<TLWE<P>>` (TFHEpp/tutorial/client.cpp, cloud.cpp), default 128-bit
//    parameter set, little-endian archives.  Layout (restated from cereal's published
//    format -- archives/portable_binary.hpp, types/{array,memory,tuple,unordered_map,vector}.hpp
//    -- and pinned byte-for-byte against the real library in tests/test_cpu_tfhepp.py):
//      u8   1                      archive byte order flag (1 = little endian)
//      lweParams                   24 fields, u32 / f64 as declared (lweParams.hpp:202-209)
//      28 tuple members in order:  shared_ptr<T>  -> u32 id: 0 = null, 0x80000000|k = k-th
//                                                    new object, followed by its bytes;
//                                  unordered_map  -> u64 size (+ entries)
//      std::array<arithmetic, N>   its bytes, no length prefix (nested arrays concatenate)
//      std::vector<T>              u64 size, then the elements
// 2. A cache of THIS engine's converted keys (NTT-domain bootstrapping key in the interchange
//    order + device key-switching key), so that a later process can skip the conversion.
#include <cstring>

namespace {

constexpr size_t kWireBkBytes = B2FHE_BK_RAW_WORDS * 4;
constexpr size_t kWireKskBytes = B2FHE_KSK_RAW_WORDS * 2;
constexpr size_t kWireParamBytes = 112;
constexpr int kWireTupleMembers = 28;
// member index -> 0: shared_ptr, 1: unordered_map   (cloudkey.hpp:20-83, default build flags)
constexpr uint8_t kWireMemberIsMap[kWireTupleMembers] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0,
                                                         0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 0, 0, 0, 0};
// payload bytes of the members this engine knows how to skip (0 = unknown)
constexpr size_t wire_member_bytes(int idx)
{
    return idx == 0 ? kWireBkBytes                              // BootstrappingKey<lvl01param>
         : idx == 4 ? (size_t)636 * 4 * 2 * 1024 * 8            // BootstrappingKeyFFT<lvl01param> (double)
         : idx == 8 ? (size_t)636 * 4 * 2 * 1024 * 8            // BootstrappingKeyNTT<lvl01param> (u64)
         : idx == 12 ? kWireKskBytes                            // KeySwitchingKey<lvl10param>
                     : 0;
}

struct WireOut {
    uint8_t* p; size_t cap, n = 0;
    WireOut(void* b, size_t c) : p(static_cast<uint8_t*>(b)), cap(c) {}
    void raw(const void* s, size_t len) { if (p && n + len <= cap) memcpy(p + n, s, len); n += len; }
    void u8(uint8_t v) { raw(&v, 1); }
    void u32(uint32_t v) { raw(&v, 4); }      // the engine runs on little-endian hosts only
    void u64(uint64_t v) { raw(&v, 8); }
    void f64(double v) { raw(&v, 8); }
};

// lweParams of the default 128-bit set (lweParams.hpp:190-209, params/128bit.hpp)
void wire_params(WireOut& w)
{
    const double a0 = 0.0000925119974676756;            // lvl0param::alpha
    w.u32(636); w.f64(a0); w.u32(16);                   // lvl0: n, alpha, approx_bit
    w.u32(10); w.u32(1024); w.u32(2); w.u32(8);         // lvl1: nbit, n, l, Bgbit
    w.f64(0.0);                                         //   alpha: never initialised by TFHEpp (lweParams.hpp:43-50); readers must ignore it
    w.u32(32);                                          //   approx_bit
    w.u32(11); w.u32(2048); w.u32(4); w.u32(10); w.u32(64);   // lvl2
    w.u32(7); w.u32(2); w.f64(a0);                      // lvl10: t, basebit, alpha
    w.u32(7); w.u32(2); w.f64(a0);                      // lvl20
    w.u32(8); w.u32(3);                                 // lvl21
    w.u32(38); w.u32(1);                                // lvl22
}

size_t wire_evalkey_bytes() { return 1 + kWireParamBytes + 24 * 4 + 4 * 8 + kWireBkBytes + kWireKskBytes; }

int wire_fail(const char* what)
{
    t_err = std::string("b2fhe wire format: ") + what;
    return B2FHE_E_BAD_FORMAT;
}

}  // namespace

extern "C" {

size_t b2fhe_wire_evalkey_size(void) { return wire_evalkey_bytes(); }

int b2fhe_wire_evalkey_write(void* blob, size_t cap, const uint32_t* bk_raw, const uint16_t* ksk_raw, size_t* written)
{
    REQUIRE(blob && bk_raw && ksk_raw && cap >= wire_evalkey_bytes(), B2FHE_E_BAD_ARGUMENT);
    WireOut w(blob, cap);
    w.u8(1);
    wire_params(w);
    uint32_t next_id = 1;
    for (int m = 0; m < kWireTupleMembers; m++) {
        if (kWireMemberIsMap[m]) { w.u64(0); continue; }
        if (m == 0) { w.u32(0x80000000u | next_id++); w.raw(bk_raw, kWireBkBytes); }
        else if (m == 12) { w.u32(0x80000000u | next_id++); w.raw(ksk_raw, kWireKskBytes); }
        else w.u32(0);
    }
    if (written) *written = w.n;
    return 0;
}

/* Finds the raw bootstrapping key and key-switching key inside an EvalKey archive.  Offsets are
 * byte offsets into `blob` (NOT aligned).  Members this engine does not use are skipped when
 * their size is known (the FFT / NTT forms of the same key), refused otherwise. */
int b2fhe_wire_evalkey_parse(const void* blob, size_t len, size_t* bk_offset, size_t* ksk_offset)
{
    REQUIRE(blob && bk_offset && ksk_offset, B2FHE_E_BAD_ARGUMENT);
    const uint8_t* p = static_cast<const uint8_t*>(blob);
    if (len < 1 + kWireParamBytes) return wire_fail("truncated header");
    if (p[0] != 1) return wire_fail("big-endian archive");
    uint8_t want[1 + kWireParamBytes];
    WireOut w(want, sizeof want);
    w.u8(1);
    wire_params(w);
    for (size_t i = 1; i < sizeof want; i++) {
        if (i >= 33 && i < 41) continue;                        // lvl1 alpha: indeterminate in TFHEpp's writer
        if (p[i] != want[i]) return wire_fail("parameter set is not TFHEpp's default 128-bit set");
    }
    size_t pos = sizeof want;
    size_t found[kWireTupleMembers] = {};
    bool have[kWireTupleMembers] = {};
    for (int m = 0; m < kWireTupleMembers; m++) {
        if (kWireMemberIsMap[m]) {
            uint64_t cnt;
            if (pos + 8 > len) return wire_fail("truncated");
            memcpy(&cnt, p + pos, 8);
            pos += 8;
            if (cnt != 0) return wire_fail("private key-switching keys present: unsupported");
            continue;
        }
        uint32_t id;
        if (pos + 4 > len) return wire_fail("truncated");
        memcpy(&id, p + pos, 4);
        pos += 4;
        if (id == 0) continue;
        if (!(id & 0x80000000u)) return wire_fail("shared pointer back-reference: unsupported");
        const size_t bytes = wire_member_bytes(m);
        if (bytes == 0) return wire_fail("archive holds a key type this engine cannot skip");
        if (pos + bytes > len) return wire_fail("truncated key");
        found[m] = pos; have[m] = true;
        pos += bytes;
    }
    if (pos != len) return wire_fail("trailing bytes");
    if (!have[0]) return wire_fail("no BootstrappingKey<lvl01param> (emplacebk<lvl01param>)");
    if (!have[12]) return wire_fail("no KeySwitchingKey<lvl10param> (emplaceiksk<lvl10param>)");
    *bk_offset = found[0];
    *ksk_offset = found[12];
    return 0;
}

/* Initialize(EvalKey) straight from the serialised key (e.g. TFHEpp tutorial's cloud.key) */
int b2fhe_init_from_evalkey(int ngpu, const void* blob, size_t len)
{
    size_t ob, ok;
    if (int r = b2fhe_wire_evalkey_parse(blob, len, &ob, &ok)) return r;
    std::vector<uint32_t> bk(B2FHE_BK_RAW_WORDS);          // the archive does not align its arrays
    std::vector<uint16_t> ksk(B2FHE_KSK_RAW_WORDS);
    memcpy(bk.data(), static_cast<const uint8_t*>(blob) + ob, kWireBkBytes);
    memcpy(ksk.data(), static_cast<const uint8_t*>(blob) + ok, kWireKskBytes);
    return b2fhe_init(ngpu, bk.data(), ksk.data());
}

/* std::vector<TLWE<P>>: u8 flag, u64 count, count x (637 u16 | 1025 u32) */
size_t b2fhe_wire_tlwes_size(int level, size_t count)
{
    return 9 + count * (level == 0 ? (size_t)kLvl0Words * 2 : (size_t)kLvl1Words * 4);
}

int b2fhe_wire_tlwes_write(void* blob, size_t cap, int level, const void* cts, size_t count, size_t* written)
{
    REQUIRE(blob && (cts || count == 0) && (level == 0 || level == 1), B2FHE_E_BAD_ARGUMENT);
    REQUIRE(cap >= b2fhe_wire_tlwes_size(level, count), B2FHE_E_BAD_ARGUMENT);
    WireOut w(blob, cap);
    w.u8(1);
    w.u64(count);
    w.raw(cts, b2fhe_wire_tlwes_size(level, count) - 9);
    if (written) *written = w.n;
    return 0;
}

int b2fhe_wire_tlwes_parse(const void* blob, size_t len, int level, size_t* count, size_t* offset)
{
    REQUIRE(blob && count && offset && (level == 0 || level == 1), B2FHE_E_BAD_ARGUMENT);
    const uint8_t* p = static_cast<const uint8_t*>(blob);
    if (len < 9) return wire_fail("truncated header");
    if (p[0] != 1) return wire_fail("big-endian archive");
    uint64_t cnt;
    memcpy(&cnt, p + 1, 8);
    if (cnt > (uint64_t)1 << 40 || b2fhe_wire_tlwes_size(level, (size_t)cnt) != len)
        return wire_fail("length does not match the element count at this level");
    *count = (size_t)cnt;
    *offset = 9;
    return 0;
}

/* ---- cache of the converted keys ---------------------------------------------------------
 * "B2FHEKC1" | u32 layout version | u32 reserved | u64 bk bytes | u64 ksk bytes |
 * u64 FNV-1a of the payload | bk (bk16_slot order) | ksk (device layout) */
namespace {
constexpr size_t kCacheHeader = 8 + 4 + 4 + 8 + 8 + 8;
constexpr uint32_t kCacheLayoutVersion = 2;      // 2: bk16_slot order, primes 2^27 - 2^11 + 1 / 2^27 - 2^21 + 1
uint64_t fnv1a64(const uint8_t* p, size_t n)
{
    // 8 interleaved lanes so that ~70 MB hash in tens of milliseconds
    uint64_t h[8];
    for (int k = 0; k < 8; k++) h[k] = 0xcbf29ce484222325ull + k;
    size_t i = 0;
    for (; i + 8 <= n; i += 8)
        for (int k = 0; k < 8; k++) h[k] = (h[k] ^ p[i + k]) * 0x100000001b3ull;
    uint64_t r = 0xcbf29ce484222325ull;
    for (; i < n; i++) r = (r ^ p[i]) * 0x100000001b3ull;
    for (int k = 0; k < 8; k++) r = (r ^ h[k]) * 0x100000001b3ull;
    return r;
}
}  // namespace

size_t b2fhe_key_cache_size(void) { return kCacheHeader + B2FHE_BK_NTT_BYTES + B2FHE_KSK_DEV_BYTES; }

int b2fhe_key_cache_export(int device, void* blob, size_t cap, size_t* written)
{
    DevCtx* c;
    if (int r = get_ctx(device, c)) return r;
    REQUIRE(blob && cap >= b2fhe_key_cache_size(), B2FHE_E_BAD_ARGUMENT);
    uint8_t* p = static_cast<uint8_t*>(blob);
    uint8_t* payload = p + kCacheHeader;
    CK(cudaMemcpy(payload, c->bk, B2FHE_BK_NTT_BYTES, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(payload + B2FHE_BK_NTT_BYTES, c->ksk, B2FHE_KSK_DEV_BYTES, cudaMemcpyDeviceToHost));
    WireOut w(p, kCacheHeader);
    w.raw("B2FHEKC1", 8);
    w.u32(kCacheLayoutVersion); w.u32(0);
    w.u64(B2FHE_BK_NTT_BYTES); w.u64(B2FHE_KSK_DEV_BYTES);
    w.u64(fnv1a64(payload, B2FHE_BK_NTT_BYTES + B2FHE_KSK_DEV_BYTES));
    if (written) *written = b2fhe_key_cache_size();
    return 0;
}

int b2fhe_init_from_key_cache(int ngpu, const void* blob, size_t len)
{
    REQUIRE(blob && ngpu >= 1 && ngpu <= kMaxDev, B2FHE_E_BAD_ARGUMENT);
    const uint8_t* p = static_cast<const uint8_t*>(blob);
    if (len != b2fhe_key_cache_size() || memcmp(p, "B2FHEKC1", 8) != 0) return wire_fail("not a key cache");
    uint32_t ver; uint64_t nb, nk, sum;
    memcpy(&ver, p + 8, 4); memcpy(&nb, p + 16, 8); memcpy(&nk, p + 24, 8); memcpy(&sum, p + 32, 8);
    if (ver != kCacheLayoutVersion || nb != B2FHE_BK_NTT_BYTES || nk != B2FHE_KSK_DEV_BYTES)
        return wire_fail("key cache written by an incompatible layout version");
    const uint8_t* payload = p + kCacheHeader;
    if (fnv1a64(payload, nb + nk) != sum) return wire_fail("key cache checksum mismatch");
    std::lock_guard<std::mutex> lk(g_mu);
    for (int d = 0; d < ngpu; d++) {
        if (int r = setup_device(d)) return r;
        REQUIRE(!g_ctx[d].ready, B2FHE_E_BAD_ARGUMENT);
    }
    DevCtx& c = g_ctx[0];
    CK(cudaSetDevice(0));
    CK(cudaMalloc(&c.bk, B2FHE_BK_NTT_BYTES));
    CK(cudaMalloc(&c.ksk, B2FHE_KSK_DEV_BYTES));
    c.owns_keys = true;
    CK(cudaMemcpy(c.bk, payload, B2FHE_BK_NTT_BYTES, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(c.ksk, payload + B2FHE_BK_NTT_BYTES, B2FHE_KSK_DEV_BYTES, cudaMemcpyHostToDevice));
    c.ready = true;
    return replicate_keys(ngpu);
}

}  // extern "C"
