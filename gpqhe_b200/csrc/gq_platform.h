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
#define GQ_LAUNCH_DYN(kernel, grid, block, smem, stream, ...) \
  ::gq_emul::launch_dyn((grid), (block), (smem), [&]() { kernel(__VA_ARGS__); })
#else
#include <cuda_runtime.h>
// Programmatic dependent launch (sm_90+): every kernel of the library starts with gq_pdl_trigger() - its successor in
// the stream may be scheduled as soon as all of this grid's blocks are resident or done - and calls gq_pdl_wait()
// before it touches anything a predecessor wrote (or writes anything a predecessor reads): launch latency and the
// constant-table prologues overlap the tail of the previous kernel instead of following it (one he_mul + he_rs:
// 296 -> 267 us when replayed from a graph, profiles/tools/graph_op.py; this gets most of that without a graph).
// gq_pdl_enabled: 1 unless GPQHE_B200_NO_PDL is set (read once, gq_api.cu); GQ_NO_PDL at compile time for libraries
// whose kernels do not carry the two calls (profiles/probe).
extern int gq_pdl_enabled;
template <typename... KA, typename... A>
static inline void gq_launch_ex(void (*kernel)(KA...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream, A &&...a) {
#ifndef GQ_NO_PDL
  if (gq_pdl_enabled) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    cudaLaunchKernelEx(&cfg, kernel, KA(a)...);
    return;
  }
#endif
  kernel<<<grid, block, smem, stream>>>(KA(a)...);
}
#define GQ_LAUNCH(kernel, grid, block, stream, ...) gq_launch_ex(kernel, (grid), (block), 0, (stream), __VA_ARGS__)
#define GQ_LAUNCH_DYN(kernel, grid, block, smem, stream, ...) gq_launch_ex(kernel, (grid), (block), (smem), (stream), __VA_ARGS__)
#endif

// device side of the programmatic dependent launch (no-ops under emulation)
#if defined(__CUDA_ARCH__) && !defined(GQ_NO_PDL)
#define GQ_PDL_TRIGGER() asm volatile("griddepcontrol.launch_dependents;" ::: "memory")
#define GQ_PDL_WAIT() asm volatile("griddepcontrol.wait;" ::: "memory")
#else
#define GQ_PDL_TRIGGER() ((void)0)
#define GQ_PDL_WAIT() ((void)0)
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
