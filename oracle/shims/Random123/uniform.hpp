// ORACLE / TEST INFRASTRUCTURE ONLY. Restatement of the public definitions in
// Random123 uniform.hpp for the instantiations the reference uses:
//   u01<float>(uint32)    = float(u) * 2^-32 + 2^-33
//   uneg11<float>(uint32) = float(int32(u)) * 2^-31 + 2^-32
//   uneg11all<T>(array)   = lane-wise uneg11
// PARITY UNPINNED against a real Random123 checkout (none is available offline; the
// reference's histograms for these are Threefry-only, test_r123.cc:612-629).
// Consumers: /root/reference/RandBLAS/random_gen.hh:171, util.hh:291-325.
#pragma once
#include "features/compilerfeatures.h"
#include "array.h"
#include <array>
#include <limits>
#include <type_traits>

namespace r123 {

template <typename T> constexpr T maxTvalue() { return std::numeric_limits<T>::max(); }

template <typename Ftype, typename Itype>
R123_STATIC_INLINE Ftype u01(Itype in) {
    typedef typename std::make_unsigned<Itype>::type Utype;
    constexpr Ftype factor = Ftype(1.) / (Ftype(maxTvalue<Utype>()) + Ftype(1.));
    constexpr Ftype halffactor = Ftype(0.5) * factor;
    return Utype(in) * factor + halffactor;
}

template <typename Ftype, typename Itype>
R123_STATIC_INLINE Ftype uneg11(Itype in) {
    typedef typename std::make_signed<Itype>::type Stype;
    constexpr Ftype factor = Ftype(1.) / (Ftype(maxTvalue<Stype>()) + Ftype(1.));
    constexpr Ftype halffactor = Ftype(0.5) * factor;
    return Stype(in) * factor + halffactor;
}

template <typename Ftype, typename CollType>
static inline std::array<Ftype, CollType::static_size> u01all(CollType in) {
    std::array<Ftype, CollType::static_size> ret;
    for (size_t i = 0; i < (size_t) CollType::static_size; ++i) ret[i] = u01<Ftype>(in[i]);
    return ret;
}

template <typename Ftype, typename CollType>
static inline std::array<Ftype, CollType::static_size> uneg11all(CollType in) {
    std::array<Ftype, CollType::static_size> ret;
    for (size_t i = 0; i < (size_t) CollType::static_size; ++i) ret[i] = uneg11<Ftype>(in[i]);
    return ret;
}

}  // namespace r123
