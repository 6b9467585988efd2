// Common macros for code that is compiled both by nvcc (device, the product) and by g++ (host-side planner
// code and the test-only lockstep emulation of the kernel, tests/emul).
#pragma once

#if defined(__CUDACC__)
#define MPCB_HD __host__ __device__ __forceinline__
#define MPCB_HD_NOINLINE __host__ __device__ __noinline__
#else
#define MPCB_HD inline
#define MPCB_HD_NOINLINE inline
#endif

#if defined(__CUDA_ARCH__)
#define MPCB_UNROLL _Pragma("unroll")
#else
#define MPCB_UNROLL
#endif

// Panda model constants: one copy in __constant__ memory for device code, one plain copy for host code.
namespace mpcb {
namespace panda_host {
#define PANDA_CONST static const
#include "panda_model_constants.h"
#undef PANDA_CONST
}  // namespace panda_host
#if defined(__CUDACC__)
namespace panda_dev {
#define PANDA_CONST static __constant__ const
#include "panda_model_constants.h"
#undef PANDA_CONST
}  // namespace panda_dev
#endif
}  // namespace mpcb

#if defined(__CUDA_ARCH__)
#define PANDA(x) (::mpcb::panda_dev::x)
#else
#define PANDA(x) (::mpcb::panda_host::x)
#endif
