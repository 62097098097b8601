// ORACLE / TEST INFRASTRUCTURE ONLY. Restatement of the public interface of
// Random123's r123array4x32 / r123array2x32 (Random123 array.h): a POD array
// `v[N]` with operator[], static_size, value_type and a little-endian
// multi-word `incr(n)` that carries across words.
//
// Pinned by: /root/reference/test/test_basic_rng/test_r123.cc:679-766 (carry
// semantics of incr) -- re-expressed in tests/test_oracle_pins.py.
// Consumers: /root/reference/RandBLAS/base.hh:166-232 (ctr_type/key_type,
// .v, static_size, value_type), dense_skops.hh:128-168 (.incr), sparse_skops.hh:71-91.
#pragma once
#include "features/compilerfeatures.h"
#include <cstddef>
#include <cstdint>
#include <iterator>

template <typename T, size_t N>
struct r123array_shim {
    typedef T value_type;
    typedef T* iterator;
    typedef const T* const_iterator;
    enum { static_size = N };
    T v[N];

    T& operator[](size_t i) { return v[i]; }
    const T& operator[](size_t i) const { return v[i]; }
    size_t size() const { return N; }
    T* data() { return v; }
    const T* data() const { return v; }
    iterator begin() { return v; }
    iterator end() { return v + N; }
    const_iterator begin() const { return v; }
    const_iterator end() const { return v + N; }
    bool operator==(const r123array_shim& o) const {
        for (size_t i = 0; i < N; ++i) if (v[i] != o.v[i]) return false;
        return true;
    }
    bool operator!=(const r123array_shim& o) const { return !(*this == o); }

    // Add the 64-bit quantity n to the N*W-bit little-endian integer held in v.
    r123array_shim& incr(R123_ULONG_LONG n = 1) {
        const unsigned W = 8 * sizeof(T);
        unsigned __int128 carry = n;
        for (size_t i = 0; i < N && carry != 0; ++i) {
            unsigned __int128 s = (unsigned __int128) v[i] + (carry & (((unsigned __int128)1 << W) - 1));
            v[i] = (T) s;
            carry = (carry >> W) + (s >> W);
        }
        return *this;
    }
};

typedef r123array_shim<uint32_t, 4> r123array4x32;
typedef r123array_shim<uint32_t, 2> r123array2x32;
typedef r123array_shim<uint32_t, 1> r123array1x32;
typedef r123array_shim<uint64_t, 4> r123array4x64;
typedef r123array_shim<uint64_t, 2> r123array2x64;
