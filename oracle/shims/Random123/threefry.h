// ORACLE / TEST INFRASTRUCTURE ONLY. /root/reference/RandBLAS/random_gen.hh:75 includes
// <Random123/threefry.h> but nothing on the north-star path instantiates Threefry
// (RNG = r123::Philox4x32 only). Intentionally empty.
#pragma once
#include "array.h"
