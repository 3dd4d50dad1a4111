// tfhepp_ref.cpp -- thin C wrapper over the REFERENCE's own CPU library (TFHEpp, vendored
// under /root/reference/thirdparties/TFHEpp), compiled from the sources where they lie
// by oracle/Makefile into oracle/_ref/libtfhepp_ref.so (git-ignored build output).
//
// TEST INFRASTRUCTURE ONLY.  It provides
//   * key / ciphertext generation and decryption by TFHEpp itself, so that the oracle's
//     and the CUDA engine's reading of TFHEpp's raw key layouts is pinned at decrypt
//     level against the real library (TFHEpp's ciphertext bits differ from
//     cuFHE's by design, decryption must agree);
//   * the timed CPU baseline: TFHEpp HomNAND<lvl01param, mu, lvl10param> (lvl0 in/out,
//     spqlios FFT), the reference's CPU implementation of the same gate
//     (TFHEpp/include/gate.hpp:60-67, TFHEpp/test/nand.cpp:42-58);
//   * the wire format: EvalKey and ciphertext vectors written / read by TFHEpp's own cereal
//     portable_binary serialisation (TFHEpp/include/cloudkey.hpp:103-107,
//     TFHEpp/tutorial/client.cpp, cloud.cpp), the pin for cufhe_b200's b2fhe_wire_* reader
//     and writer.
// None of this file is copied from the reference; it only calls its public API.
#include <omp.h>
#include <sstream>
#include <cereal/archives/portable_binary.hpp>
#include <cereal/types/vector.hpp>
#include <chrono>
#include <cstring>
#include <memory>
#include <vector>
#include <tfhe++.hpp>

using namespace TFHEpp;
static std::unique_ptr<SecretKey> g_sk;
static std::unique_ptr<EvalKey> g_ek;

extern "C" {

int ref_keygen(void)
{
    g_sk = std::make_unique<SecretKey>();
    g_ek = std::make_unique<EvalKey>();
    g_ek->emplacebk<lvl01param>(*g_sk);
    g_ek->emplacebkfft<lvl01param>(*g_sk);
    g_ek->emplaceiksk<lvl10param>(*g_sk);
    return 0;
}
// raw arrays in TFHEpp's own layouts (what cufhe::Initialize(EvalKey) uploads:
// reference src/bootstrap_gpu.cu:111-138, src/keyswitch_gpu.cu:6-16)
size_t ref_bk_bytes(void) { return sizeof(BootstrappingKey<lvl01param>); }
size_t ref_ksk_bytes(void) { return sizeof(KeySwitchingKey<lvl10param>); }
int ref_export_keys(int32_t* sk0, int32_t* sk1, void* bk, void* ksk)
{
    if (!g_sk) ref_keygen();
    const auto k0 = g_sk->key.get<lvl0param>();
    const auto k1 = g_sk->key.get<lvl1param>();
    for (size_t i = 0; i < k0.size(); i++) sk0[i] = (int32_t)(std::make_signed_t<lvl0param::T>)k0[i];
    for (size_t i = 0; i < k1.size(); i++) sk1[i] = (int32_t)(std::make_signed_t<lvl1param::T>)k1[i];
    std::memcpy(bk, &g_ek->getbk<lvl01param>(), sizeof(BootstrappingKey<lvl01param>));
    std::memcpy(ksk, &g_ek->getiksk<lvl10param>(), sizeof(KeySwitchingKey<lvl10param>));
    return 0;
}
// message = bit ? mu : -mu (reference test/test_util.h:23-27)
void ref_encrypt(int level, const uint8_t* bits, int n, void* out)
{
    if (!g_sk) ref_keygen();
    for (int i = 0; i < n; i++) {
        if (level == 0) {
            TLWE<lvl0param> c;
            tlweSymEncrypt<lvl0param>(c, bits[i] ? lvl0param::μ : -lvl0param::μ, *g_sk);
            std::memcpy((char*)out + (size_t)i * sizeof c, c.data(), sizeof c);
        } else {
            TLWE<lvl1param> c;
            tlweSymEncrypt<lvl1param>(c, bits[i] ? lvl1param::μ : -lvl1param::μ, *g_sk);
            std::memcpy((char*)out + (size_t)i * sizeof c, c.data(), sizeof c);
        }
    }
}
void ref_decrypt(int level, const void* in, int n, uint8_t* bits)
{
    for (int i = 0; i < n; i++) {
        if (level == 0) {
            TLWE<lvl0param> c;
            std::memcpy(c.data(), (const char*)in + (size_t)i * sizeof c, sizeof c);
            bits[i] = tlweSymDecrypt<lvl0param>(c, g_sk->key.get<lvl0param>());
        } else {
            TLWE<lvl1param> c;
            std::memcpy(c.data(), (const char*)in + (size_t)i * sizeof c, sizeof c);
            bits[i] = tlweSymDecrypt<lvl1param>(c, g_sk->key.get<lvl1param>());
        }
    }
}
// TFHEpp's own gate (lvl0 in/out), OpenMP over independent gates
void ref_homnand_lvl0(void* out, const void* in0, const void* in1, int n, int threads)
{
    const size_t sz = sizeof(TLWE<lvl0param>);
#pragma omp parallel for num_threads(threads) schedule(dynamic, 1)
    for (int i = 0; i < n; i++) {
        TLWE<lvl0param> a, b, r;
        std::memcpy(a.data(), (const char*)in0 + i * sz, sz);
        std::memcpy(b.data(), (const char*)in1 + i * sz, sz);
        HomNAND<lvl01param, lvl1param::μ, lvl10param>(r, a, b, *g_ek);
        std::memcpy((char*)out + i * sz, r.data(), sz);
    }
}
// seconds to evaluate n lvl0 NANDs over `threads` host threads (keys made on first call)
double ref_bench_nand(int n, int threads)
{
    if (!g_sk) ref_keygen();
    std::vector<uint8_t> bits(n);
    for (int i = 0; i < n; i++) bits[i] = (i * 7 + 3) & 1;
    std::vector<TLWE<lvl0param>> a(n), b(n), r(n);
    ref_encrypt(0, bits.data(), n, a.data());
    ref_encrypt(0, bits.data(), n, b.data());
    ref_homnand_lvl0(r.data(), a.data(), b.data(), threads < n ? threads : n, threads);   // warm-up
    auto t0 = std::chrono::steady_clock::now();
    ref_homnand_lvl0(r.data(), a.data(), b.data(), n, threads);
    auto t1 = std::chrono::steady_clock::now();
    return std::chrono::duration<double>(t1 - t0).count();
}
// ---- wire format pins -------------------------------------------------------------------
// EvalKey as TFHEpp writes it (tutorial/client.cpp: `ek.serialize(ar)` on a
// PortableBinaryOutputArchive).  which: 0 = bk<lvl01> + iksk<lvl10> (what cuFHE's tests
// create), 1 = also bkfft<lvl01> (what the TFHEpp tutorial's cloud.key holds)
size_t ref_serialize_evalkey(int which, void* buf, size_t cap)
{
    if (!g_sk) ref_keygen();
    EvalKey ek(*g_sk);
    ek.get<BootstrappingKey<lvl01param>>() = g_ek->get<BootstrappingKey<lvl01param>>();
    ek.get<KeySwitchingKey<lvl10param>>() = g_ek->get<KeySwitchingKey<lvl10param>>();
    if (which == 1) ek.get<BootstrappingKeyFFT<lvl01param>>() = g_ek->get<BootstrappingKeyFFT<lvl01param>>();
    std::ostringstream os(std::ios::binary);
    {
        cereal::PortableBinaryOutputArchive ar(os);
        ek.serialize(ar);
    }
    const std::string s = os.str();
    if (buf && s.size() <= cap) std::memcpy(buf, s.data(), s.size());
    return s.size();
}
// reads an EvalKey blob with TFHEpp's own loader; copies the raw bk / iksk out. 0 = ok
int ref_deserialize_evalkey(const void* blob, size_t len, void* bk, void* ksk)
{
    try {
        std::istringstream is(std::string((const char*)blob, len), std::ios::binary);
        cereal::PortableBinaryInputArchive ar(is);
        EvalKey ek;
        ek.serialize(ar);
        if (!ek.get<BootstrappingKey<lvl01param>>() || !ek.get<KeySwitchingKey<lvl10param>>()) return 2;
        if (!(ek.params == lweParams())) return 3;
        std::memcpy(bk, ek.get<BootstrappingKey<lvl01param>>().get(), sizeof(BootstrappingKey<lvl01param>));
        std::memcpy(ksk, ek.get<KeySwitchingKey<lvl10param>>().get(), sizeof(KeySwitchingKey<lvl10param>));
        return 0;
    } catch (...) {
        return 1;
    }
}
// std::vector<TLWE<P>> as tutorial/client.cpp writes cloud.data
size_t ref_serialize_tlwes(int level, const void* cts, int n, void* buf, size_t cap)
{
    std::ostringstream os(std::ios::binary);
    {
        cereal::PortableBinaryOutputArchive ar(os);
        if (level == 0) {
            std::vector<TLWE<lvl0param>> v(n);
            std::memcpy((void*)v.data(), cts, (size_t)n * sizeof(TLWE<lvl0param>));
            ar(v);
        } else {
            std::vector<TLWE<lvl1param>> v(n);
            std::memcpy((void*)v.data(), cts, (size_t)n * sizeof(TLWE<lvl1param>));
            ar(v);
        }
    }
    const std::string s = os.str();
    if (buf && s.size() <= cap) std::memcpy(buf, s.data(), s.size());
    return s.size();
}
// returns the element count (or -1); copies at most cap_n ciphertexts out
int ref_deserialize_tlwes(int level, const void* blob, size_t len, void* out, int cap_n)
{
    try {
        std::istringstream is(std::string((const char*)blob, len), std::ios::binary);
        cereal::PortableBinaryInputArchive ar(is);
        if (level == 0) {
            std::vector<TLWE<lvl0param>> v;
            ar(v);
            std::memcpy(out, v.data(), std::min<size_t>(v.size(), cap_n) * sizeof(TLWE<lvl0param>));
            return (int)v.size();
        }
        std::vector<TLWE<lvl1param>> v;
        ar(v);
        std::memcpy(out, v.data(), std::min<size_t>(v.size(), cap_n) * sizeof(TLWE<lvl1param>));
        return (int)v.size();
    } catch (...) {
        return -1;
    }
}
}
