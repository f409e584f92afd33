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

// Early trigger (small batches only): after the wait a CTA may at once allow the NEXT grid to be scheduled
// (griddepcontrol.launch_dependents).  The dependent grid is launched when ALL CTAs of this grid have got there (or exited), i.e.
// when every CTA of this grid is resident and holds its shared / tensor memory, so early dependents can never starve a CTA of
// this grid; they run their prologue (barrier init, TMEM allocation as soon as columns are free, descriptor prefetch) on whatever
// SM resources are free and then block in their own griddepcontrol.wait until this grid has completed and flushed.  Measured
// (profiles/r02_sweep_batch_pdl_early.csv): B = 1 6.64 -> 6.29 ms, B = 2 9.39 -> 9.19 ms, but B = 4 15.3 -> 15.6 ms and B = 16
// 54.6 -> 57.2 ms (early-resident dependents take issue slots and TMEM / smem headroom from a full-chip grid), so the engine switches
// it on per pass for <= 2 pairs only.  The flag lives in constant memory, one copy per translation unit (no -rdc), set through
// the ufm_set_pdl_early_<tu>() functions the macro below defines.
static __constant__ int c_pdl_early;
__device__ __forceinline__ void pdl_wait() {
  asm volatile("griddepcontrol.wait;" ::: "memory");
  if (c_pdl_early) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}
#define UFM_DEFINE_PDL_EARLY_SETTER(tu)                                                                   \
  cudaError_t ufm_set_pdl_early_##tu(int v, cudaStream_t s) {                                             \
    return cudaMemcpyToSymbolAsync(c_pdl_early, &v, sizeof(int), 0, cudaMemcpyHostToDevice, s);           \
  }

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
