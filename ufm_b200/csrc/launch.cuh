// Kernel launch helper: every kernel of the engine is launched with programmatic stream serialisation (PDL), so the
// launch latency and the prologue of kernel N+1 overlap the tail of kernel N.  Contract: every kernel executes
// pdl_wait() (griddepcontrol.wait: returns when the preceding kernel in the stream has completed and its writes are
// visible) in EVERY CTA before it touches global memory and before any early return -- a CTA that skipped the wait
// would let its grid "complete" while the grid before it still runs, and break the chain for the grid after it.
#pragma once
#include <cuda_runtime.h>

#include <utility>

namespace ufm {

int pdl_mode();            // 1 = launch with the programmatic-serialisation attribute (default), 0 = plain launches
void set_pdl_mode(int v);  // tuning hook (ufm_debug_flag 4)

__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

template <typename... KArgs, typename... Args>
inline cudaError_t launch_kc(void (*kernel)(KArgs...), int cluster, dim3 grid, dim3 block, size_t smem, cudaStream_t s,
                             Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute at[2];
  unsigned n = 0;
  if (cluster > 1) {
    at[n].id = cudaLaunchAttributeClusterDimension;
    at[n].val.clusterDim.x = (unsigned)cluster;
    at[n].val.clusterDim.y = 1;
    at[n].val.clusterDim.z = 1;
    ++n;
  }
  if (pdl_mode()) {
    at[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[n].val.programmaticStreamSerializationAllowed = 1;
    ++n;
  }
  cfg.attrs = at;
  cfg.numAttrs = n;
  return cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}

template <typename... KArgs, typename... Args>
inline cudaError_t launch_k(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args&&... args) {
  return launch_kc(kernel, 1, grid, block, smem, s, std::forward<Args>(args)...);
}

}  // namespace ufm
