// Programmatic dependent launch (PDL): a kernel launched with launch_pdl may start while the previous
// kernel of the stream is still draining; it must call pdl_wait() before it touches anything the
// previous kernels wrote (or overwrites anything they read).  pdl_trigger() in the earlier kernel lets
// the next one be scheduled as soon as every CTA of this one has started, so launch latency and the
// next kernel's independent prologue overlap this kernel's tail.  Both are no-ops in a kernel that
// was launched the ordinary way.
#pragma once
#include <cuda_runtime.h>

namespace sfmba {

__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                              Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

}  // namespace sfmba
