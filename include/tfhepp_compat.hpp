// tfhepp_compat.hpp -- the handful of TFHEpp names the cufhe:: API surface needs, for
// builds where the real TFHEpp headers are not on the include path (e.g. the GPU test
// box).  With TFHEpp available, compile with -DB2FHE_WITH_TFHEPP and this file is unused.
// Values: TFHEpp/include/params/128bit.hpp:7-72,226-235 (default 128-bit set).
#pragma once
#include <array>
#include <cstddef>
#include <cstdint>
#include <vector>

namespace TFHEpp {
struct lvl0param {
    using T = uint16_t;
    static constexpr int32_t k = 1;
    static constexpr uint32_t n = 636;
    static constexpr T μ = 1U << 13;
};
struct lvl1param {
    using T = uint32_t;
    static constexpr int32_t k = 1;
    static constexpr uint32_t n = 1024;
    static constexpr uint32_t l = 2;
    static constexpr uint32_t Bgbit = 8;
    static constexpr T μ = 1U << 29;
};
struct lvl01param { using domainP = lvl0param; using targetP = lvl1param; };
struct lvl10param {
    using domainP = lvl1param; using targetP = lvl0param;
    static constexpr uint32_t t = 7;
    static constexpr uint32_t basebit = 2;
};
template <class P> using TLWE = std::array<typename P::T, P::k * P::n + 1>;
template <class P> using TRLWE = std::array<std::array<typename P::T, P::n>, P::k + 1>;
template <class P> using TRGSW = std::array<TRLWE<P>, (P::k + 1) * P::l>;
template <class P> using BootstrappingKey = std::array<TRGSW<typename P::targetP>, P::domainP::k * P::domainP::n>;
template <class P>
using KeySwitchingKey = std::array<std::array<std::array<TLWE<typename P::targetP>, (1 << (P::basebit - 1))>, P::t>,
                                   P::domainP::k * P::domainP::n>;
// raw-array stand-in for TFHEpp::EvalKey: same accessors the engine uses
struct EvalKey {
    std::vector<uint32_t> bk;     // BootstrappingKey<lvl01param> bytes
    std::vector<uint16_t> iksk;   // KeySwitchingKey<lvl10param> bytes
    template <class P> const BootstrappingKey<P>& getbk() const { return *reinterpret_cast<const BootstrappingKey<P>*>(bk.data()); }
    template <class P> const KeySwitchingKey<P>& getiksk() const { return *reinterpret_cast<const KeySwitchingKey<P>*>(iksk.data()); }
};
}  // namespace TFHEpp
