// gq_platform.h - compilation-mode glue.
//
// The product is compiled by nvcc for sm_100a (GQ_EMULATE undefined).  The same
// sources can also be compiled by g++ with -DGQ_EMULATE against tests/emul/, a
// fibre-based single-process stand-in for the CUDA execution model that exists
// ONLY so the `-m "not gpu"` test tier can exercise host logic and kernel index
// arithmetic in a container without a GPU.  The emulated library is never loaded
// by gpqhe_b200/capi.py unless a test passes its path explicitly.
#pragma once

#include <stddef.h>
#include <stdint.h>

#ifdef GQ_EMULATE
#include "cuda_emul.h"   // tests/emul/cuda_emul.h (test infrastructure)
#define GQ_LAUNCH(kernel, grid, block, stream, ...) \
  ::gq_emul::launch((grid), (block), [&]() { kernel(__VA_ARGS__); })
#else
#include <cuda_runtime.h>
#define GQ_LAUNCH(kernel, grid, block, stream, ...) \
  kernel<<<(grid), (block), 0, (stream)>>>(__VA_ARGS__)
#endif

// dynamic shared memory of a kernel launched with a byte count
#ifdef GQ_EMULATE
#define GQ_DYN_SMEM(type, name) type *name = reinterpret_cast<type *>(::gq_emul::dyn_smem())
#else
#define GQ_DYN_SMEM(type, name) extern __shared__ __align__(16) unsigned char name##_raw_[]; \
  type *name = reinterpret_cast<type *>(name##_raw_)
#endif

#define GQ_HD __host__ __device__ __forceinline__
#define GQ_DEV __device__ __forceinline__

typedef unsigned __int128 gq_u128;

// B200: 148 SMs (2 dies x 74).  Grid sizes for grid-stride kernels are multiples of it.
#define GQ_NUM_SMS 148
