// launch.cuh — kernel launches with programmatic dependent launch (PDL).
//
// The hot path is chains of short dependent kernels (a key switch is ~25 launches of one to three waves).  Launched with
// cudaLaunchAttributeProgrammaticStreamSerialization, a kernel's thread blocks may be placed on the SMs while the previous
// kernel of the stream is still draining; they run their prologue (index arithmetic, per-prime constants) and stop at
// pdl_wait() — griddepcontrol.wait — until the previous grid has completed and its writes are visible.  Every kernel launched
// through launch_k() calls pdl_wait() before its first access to data another kernel may have written or may still be
// reading, so the ordering the stream guarantees is unchanged; only the launch latency and the prologue move under the
// predecessor's tail.  pdl_trigger() lets the NEXT kernel start placing blocks once all blocks of this one have started.
// Launched without the attribute the two instructions are no-ops.
#pragma once
#include <cuda_runtime.h>

#include <utility>

namespace b200fhe {

// capi_host.cu: grids of at least this many thread blocks are launched with the attribute (0 = never; default 1 = all;
// B200FHE_PDL_MIN sets the threshold, B200FHE_PDL=0 switches the attribute off).  Measured on the B200 (scratch/pdl_ab.sh):
// bootstrap 42.2 -> 41.3 ms, sparse bootstrap 22.2 -> 21.4 ms, headline transform +1 %; with a threshold of 1000 blocks 41.7 / 21.7.
unsigned pdl_min_blocks();
// launches made between pdl_suppress(true) and pdl_suppress(false) on this host thread go without the attribute (the digit
// lanes of a key switch: their kernels run beside each other on priority streams and early placement starves the neighbours)
void pdl_suppress(bool on);
bool pdl_suppressed();

template <class... KArgs, class... Args>
inline cudaError_t launch_k(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args &&...args) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at{};
    at.id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at.val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = &at;
    const unsigned min_blocks = pdl_min_blocks();
    cfg.numAttrs = min_blocks && !pdl_suppressed() && size_t(grid.x) * grid.y * grid.z >= min_blocks ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(std::forward<Args>(args))...);
}

__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

}  // namespace b200fhe
