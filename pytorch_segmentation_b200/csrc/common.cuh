// Host-side helpers shared by the C-ABI translation units: error reporting and TMA descriptor encoding.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/b200seg.h"

namespace b200seg {

int set_error(const char* fmt, ...);   // formats into the thread-local message, returns 1
void clear_error();

#define B200SEG_CHECK(cond, ...)                      \
    do {                                              \
        if (!(cond)) return ::b200seg::set_error(__VA_ARGS__); \
    } while (0)

#define B200SEG_CUDA(expr)                                                                         \
    do {                                                                                           \
        cudaError_t _e = (expr);                                                                   \
        if (_e != cudaSuccess)                                                                     \
            return ::b200seg::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
    } while (0)

// after a kernel launch
#define B200SEG_LAUNCH_CHECK() B200SEG_CUDA(cudaGetLastError())

int sm_count();   // SMs of the current device (cached), 0 on failure

// bf16 tensor map over up to 4 dims (dim 0 innermost, contiguous).  strides_bytes[i] is the stride of dim i+1.
// SWIZZLE_128B, zero OOB fill, box[0] must be 64 (128 bytes).  Returns 0 on success.
int encode_tmap_bf16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                     const uint32_t* box);

// validates a bf16 NHWC channel slice (multiples of 8 channels, 16-byte aligned base); 0 = OK
int check_slice(const char* what, const void* p, int c_total, int c_off, int c);
// grid size for a 256-thread grid-stride elementwise kernel over `total` items (capped at 16 CTAs per SM)
int elementwise_grid(int64_t total);
int gemm_sm_count();

inline int ceil_div(int a, int b) { return (a + b - 1) / b; }
inline int64_t ceil_div64(int64_t a, int64_t b) { return (a + b - 1) / b; }
inline int round_up(int a, int b) { return ceil_div(a, b) * b; }

}  // namespace b200seg
