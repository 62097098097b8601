// ORACLE / TEST INFRASTRUCTURE ONLY. Restatement of Philox4x32-10 (Salmon et al.,
// "Parallel random numbers: as easy as 1, 2, 3", SC'11; Random123 philox.h).
// Pinned by the known-answer vectors the reference carries:
// /root/reference/test/test_basic_rng/r123_kat_vectors.txt:16-21 (philox4x32 7 and 10 rounds).
// Consumers: /root/reference/RandBLAS/random_gen.hh:143,171; sparse_skops.hh:78; base.hh:161.
#pragma once
#include "array.h"

namespace r123 {

template <unsigned ROUNDS>
struct Philox4x32_R {
    typedef r123array4x32 ctr_type;
    typedef r123array2x32 key_type;
    typedef r123array2x32 ukey_type;
    static const unsigned rounds = ROUNDS;

    static inline void mulhilo(uint32_t a, uint32_t b, uint32_t& hi, uint32_t& lo) {
        uint64_t p = (uint64_t) a * (uint64_t) b;
        hi = (uint32_t)(p >> 32);
        lo = (uint32_t) p;
    }

    ctr_type operator()(ctr_type c, key_type k) const {
        const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
        const uint32_t W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
        for (unsigned r = 0; r < ROUNDS; ++r) {
            if (r > 0) { k.v[0] += W0; k.v[1] += W1; }
            uint32_t hi0, lo0, hi1, lo1;
            mulhilo(M0, c.v[0], hi0, lo0);
            mulhilo(M1, c.v[2], hi1, lo1);
            ctr_type o = {{ hi1 ^ c.v[1] ^ k.v[0], lo1, hi0 ^ c.v[3] ^ k.v[1], lo0 }};
            c = o;
        }
        return c;
    }
    key_type keyinit(ukey_type uk) const { return uk; }
};

typedef Philox4x32_R<10> Philox4x32;

}  // namespace r123
