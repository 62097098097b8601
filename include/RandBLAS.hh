// RandBLAS.hh -- drop-in for the reference's single include (RandBLAS.hh:33-41): put this repository's include/ directory
// ahead of the reference's on the include path and `#include <RandBLAS.hh>` resolves to the B200 implementation
// (RandBLAS_b200.hh over the rbx C ABI, link librbx.so). The per-component includes the reference offers
// (<RandBLAS/base.hh>, <RandBLAS/skge.hh>, ...) forward here as well (include/RandBLAS/*.hh).
#ifndef RandBLAS_HH
#define RandBLAS_HH
#include "RandBLAS_b200.hh"
#endif
