// Layout conversion at the drop-in boundary (fp32 NCHW <-> bf16 NHWC slices) and conv weight packing.
#include "common.cuh"
#include "pdl.cuh"
#include "ptx.cuh"

namespace b200seg {
namespace {

// 32x32 (channel x pixel) transpose through shared memory: reads coalesced along pixels (NCHW), writes
// coalesced along channels (NHWC).
__global__ void nchw_to_nhwc_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, int c, int hw,
                                    int dst_c_total, int dst_c_offset) {
    pdl_prologue();
    __shared__ float tile[32][33];
    const int n = blockIdx.z;
    const int p0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
    const float* s = src + static_cast<size_t>(n) * c * hw;
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const int cc = c0 + i, pp = p0 + threadIdx.x;
        tile[i][threadIdx.x] = (cc < c && pp < hw) ? s[static_cast<size_t>(cc) * hw + pp] : 0.f;
    }
    __syncthreads();
    __nv_bfloat16* d = dst + static_cast<size_t>(n) * hw * dst_c_total + dst_c_offset;
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const int pp = p0 + i, cc = c0 + threadIdx.x;
        if (pp < hw && cc < c) d[static_cast<size_t>(pp) * dst_c_total + cc] = __float2bfloat16(tile[threadIdx.x][i]);
    }
}

__global__ void nhwc_to_nchw_kernel(const __nv_bfloat16* __restrict__ src, float* __restrict__ dst, int c, int hw,
                                    int src_c_total, int src_c_offset) {
    pdl_prologue();
    __shared__ float tile[32][33];
    const int n = blockIdx.z;
    const int p0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
    const __nv_bfloat16* s = src + static_cast<size_t>(n) * hw * src_c_total + src_c_offset;
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const int pp = p0 + i, cc = c0 + threadIdx.x;
        tile[i][threadIdx.x] = (pp < hw && cc < c) ? __bfloat162float(s[static_cast<size_t>(pp) * src_c_total + cc]) : 0.f;
    }
    __syncthreads();
    float* d = dst + static_cast<size_t>(n) * c * hw;
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const int cc = c0 + i, pp = p0 + threadIdx.x;
        if (cc < c && pp < hw) d[static_cast<size_t>(cc) * hw + pp] = tile[threadIdx.x][i];
    }
}

// One thread per packed element; both packings in one launch (grid.y selects which).
__global__ void pack_weights_kernel(const float* __restrict__ w, int k, int c, int r, int s, __nv_bfloat16* wf,
                                    int k_rows, __nv_bfloat16* wd, int c_rows, int k_cols) {
    pdl_prologue();
    const int taps = r * s;
    const long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
    if (blockIdx.y == 0) {
        if (!wf) return;
        const long long total = static_cast<long long>(taps) * k_rows * c;
        if (i >= total) return;
        const int cc = static_cast<int>(i % c);
        const int kk = static_cast<int>((i / c) % k_rows);
        const int tap = static_cast<int>(i / (static_cast<long long>(c) * k_rows));
        const float v = kk < k ? w[(static_cast<size_t>(kk) * c + cc) * taps + tap] : 0.f;
        wf[i] = __float2bfloat16(v);
    } else {
        if (!wd) return;
        const long long total = static_cast<long long>(taps) * c_rows * k_cols;
        if (i >= total) return;
        const int kk = static_cast<int>(i % k_cols);
        const int cc = static_cast<int>((i / k_cols) % c_rows);
        const int tap = static_cast<int>(i / (static_cast<long long>(k_cols) * c_rows));
        // dgrad reads dy at +offset(tap') where tap' mirrors the forward tap: (r', s') = (R-1-r, S-1-s)
        const int src_tap = taps - 1 - tap;
        const float v = (cc < c && kk < k) ? w[(static_cast<size_t>(kk) * c + cc) * taps + src_tap] : 0.f;
        wd[i] = __float2bfloat16(v);
    }
}

}  // namespace
}  // namespace b200seg

extern "C" {

int b200seg_nchw_to_nhwc(const float* src, void* dst, int n, int c, int h, int w, int dst_c_total, int dst_c_offset,
                         b200seg_stream_t st) {
    using namespace b200seg;
    clear_error();
    B200SEG_CHECK(src && dst && n > 0 && c > 0 && h > 0 && w > 0, "nchw_to_nhwc: bad arguments");
    B200SEG_CHECK(dst_c_offset + c <= dst_c_total, "nchw_to_nhwc: slice exceeds buffer");
    const int hw = h * w;
    dim3 grid(ceil_div(hw, 32), ceil_div(c, 32), n), block(32, 8);
    B200SEG_CUDA(launch_pdl(nchw_to_nhwc_kernel, grid, block, 0, static_cast<cudaStream_t>(st), src, static_cast<__nv_bfloat16*>(dst), c, hw,
                                                                             dst_c_total, dst_c_offset));
    return 0;
}

int b200seg_nhwc_to_nchw(const void* src, float* dst, int n, int c, int h, int w, int src_c_total, int src_c_offset,
                         b200seg_stream_t st) {
    using namespace b200seg;
    clear_error();
    B200SEG_CHECK(src && dst && n > 0 && c > 0 && h > 0 && w > 0, "nhwc_to_nchw: bad arguments");
    B200SEG_CHECK(src_c_offset + c <= src_c_total, "nhwc_to_nchw: slice exceeds buffer");
    const int hw = h * w;
    dim3 grid(ceil_div(hw, 32), ceil_div(c, 32), n), block(32, 8);
    B200SEG_CUDA(launch_pdl(nhwc_to_nchw_kernel, grid, block, 0, static_cast<cudaStream_t>(st), static_cast<const __nv_bfloat16*>(src), dst, c,
                                                                             hw, src_c_total, src_c_offset));
    return 0;
}

int b200seg_pack_weights(const float* w_oihw, int k, int c, int r, int s, void* w_fprop, int k_rows, void* w_dgrad,
                         int c_rows, int k_cols, b200seg_stream_t st) {
    using namespace b200seg;
    clear_error();
    B200SEG_CHECK(w_oihw && k > 0 && c > 0 && r > 0 && s > 0, "pack_weights: bad arguments");
    B200SEG_CHECK(!w_fprop || k_rows >= k, "pack_weights: k_rows %d < k %d", k_rows, k);
    B200SEG_CHECK(!w_dgrad || (c_rows >= c && k_cols >= k), "pack_weights: c_rows %d < c %d or k_cols %d < k %d", c_rows,
                  c, k_cols, k);
    const long long tf = w_fprop ? static_cast<long long>(r) * s * k_rows * c : 0;
    const long long td = w_dgrad ? static_cast<long long>(r) * s * c_rows * k_cols : 0;
    const long long tmax = tf > td ? tf : td;
    if (tmax == 0) return 0;
    dim3 grid(static_cast<unsigned>((tmax + 255) / 256), 2);
    B200SEG_CUDA(launch_pdl(pack_weights_kernel, grid, 256, 0, static_cast<cudaStream_t>(st), 
        w_oihw, k, c, r, s, static_cast<__nv_bfloat16*>(w_fprop), k_rows, static_cast<__nv_bfloat16*>(w_dgrad), c_rows,
        k_cols));
    return 0;
}

}  // extern "C"
