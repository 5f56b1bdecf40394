// Programmatic dependent launch (PDL): every libb200seg kernel is launched with the programmatic-stream-serialization
// attribute and begins with pdl_prologue().  A training step is ~600 launches of 5-300 us kernels on one stream; with
// PDL the next kernel's CTAs are scheduled (and run their prologue: shared-memory carve-up, mbarrier init, TMEM
// allocation, tensor-map prefetch) while the previous kernel drains, instead of paying a full launch latency after it.
//
// Rules every kernel here follows:
//   * griddepcontrol.wait comes BEFORE the first global-memory access of any kind (reads of a predecessor's output,
//     and writes — the caching allocator may hand us memory a predecessor is still reading);
//   * griddepcontrol.launch_dependents is issued at the top: the dependent grid still waits (at its own
//     griddepcontrol.wait) for this grid to complete and flush, so early triggering is only a scheduling hint.
// Kernels of other libraries interleaved on the stream are launched normally and serialise as usual.
// B200SEG_PDL=0 in the environment turns the attribute off (the device-side instructions are then no-ops).
#pragma once
#include <cuda_runtime.h>
#include <stdlib.h>

#include <utility>

namespace b200seg {

__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_prologue() {
    pdl_launch_dependents();
    pdl_wait();
}

inline bool pdl_enabled() {
    static const bool on = [] {
        const char* e = getenv("B200SEG_PDL");
        return !(e && e[0] == '0');
    }();
    return on;
}

template <typename... KArgs, typename... Args>
cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}

// the same launch with a thread-block cluster of `cluster` CTAs along x (grid.x must be a multiple of it)
template <typename... KArgs, typename... Args>
cudaError_t launch_pdl_cluster(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, int cluster,
                               Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[2];
    int n = 0;
    if (cluster > 1) {
        attr[n].id = cudaLaunchAttributeClusterDimension;
        attr[n].val.clusterDim.x = cluster;
        attr[n].val.clusterDim.y = 1;
        attr[n].val.clusterDim.z = 1;
        ++n;
    }
    if (pdl_enabled()) {
        attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[n].val.programmaticStreamSerializationAllowed = 1;
        ++n;
    }
    cfg.attrs = attr;
    cfg.numAttrs = n;
    return cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}

}  // namespace b200seg
