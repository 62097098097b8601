// ORACLE / TEST INFRASTRUCTURE ONLY -- never part of the shipped product path.
//
// Minimal stand-in for Random123's <Random123/features/compilerfeatures.h>.
// Random123 (DEShawResearch/random123, release 1.14.0; un-pinned by the reference,
// see /root/reference/.github/workflows/core-linux.yaml:35-40) is NOT vendored in
// /root/reference, so the reference headers cannot compile without it. This
// directory restates the few public pieces the reference uses so that the
// reference's own headers can be compiled, untouched, into oracle/_ref/.
//
// Consumers: /root/reference/RandBLAS/random_gen.hh:34,49,54 (R123_CUDA_DEVICE,
// R123_STATIC_INLINE).
#pragma once
#include <stdint.h>
#include <stddef.h>

#ifndef R123_CUDA_DEVICE
#define R123_CUDA_DEVICE
#endif
#ifndef R123_STATIC_INLINE
#define R123_STATIC_INLINE static inline
#endif
#ifndef R123_FORCE_INLINE
#define R123_FORCE_INLINE(decl) decl __attribute__((always_inline))
#endif
#ifndef R123_CONSTEXPR
#define R123_CONSTEXPR constexpr
#endif
#ifndef R123_ULONG_LONG
#define R123_ULONG_LONG unsigned long long
#endif
#ifndef R123_NO_SINCOS
#define R123_NO_SINCOS 0
#endif
