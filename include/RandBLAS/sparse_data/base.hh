#include "../../RandBLAS_b200.hh"   // forwards to the single B200 header (reference: RandBLAS/sparse_data/base.hh)
