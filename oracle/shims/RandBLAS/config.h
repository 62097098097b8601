// ORACLE / TEST INFRASTRUCTURE ONLY. Hand-written stand-in for the CMake-generated
// RandBLAS/config.h (/root/reference/RandBLAS/config.h.in:44): OpenMP on.
#pragma once
#define RandBLAS_FULL_VERSION "oracle-shim"
#define RandBLAS_VERSION_MAJOR 0
#define RandBLAS_VERSION_MINOR 2
#define RandBLAS_VERSION_PATCH 0
#define RandBLAS_HAS_OpenMP
