// ORACLE / TEST INFRASTRUCTURE ONLY. Restatement of Random123 boxmuller.hpp,
// float overload: boxmuller(u0,u1) = (sin, cos)(pi*uneg11<float>(u0)) * sqrtf(-2*logf(u01<float>(u1))).
// `sincospif` is NOT defined here: on the host it is supplied by the reference itself
// (/root/reference/RandBLAS/random_gen.hh:61-65, = sincosf(3.1415926535897932f*x)), which is
// why that header is included before this one (random_gen.hh:36-77).
// PARITY UNPINNED against a real Random123 checkout (see uniform.hpp).
// Consumer: /root/reference/RandBLAS/random_gen.hh:105.
#pragma once
#include "features/compilerfeatures.h"
#include "uniform.hpp"
#include <math.h>

namespace r123 {

typedef struct { float x, y; } float2;
typedef struct { double x, y; } double2;

R123_STATIC_INLINE float2 boxmuller(uint32_t u0, uint32_t u1) {
    float r;
    float2 f;
    sincospif(uneg11<float>(u0), &f.x, &f.y);
    r = sqrtf(-2.f * logf(u01<float>(u1)));  // u01 never returns 0
    f.x *= r;
    f.y *= r;
    return f;
}

R123_STATIC_INLINE double2 boxmuller(uint64_t u0, uint64_t u1) {
    double r;
    double2 f;
    sincospi(uneg11<double>(u0), &f.x, &f.y);
    r = sqrt(-2. * log(u01<double>(u1)));
    f.x *= r;
    f.y *= r;
    return f;
}

}  // namespace r123
