// 3x3 stride-2 pad-1 max pooling of the ResNet stem (backbones.resnet50, used at models/deeplabv3plus.py:17-19,29) on
// NHWC bf16.  HBM-bound: one thread per (pixel, 8-channel group), 128-bit accesses.  The forward also stores the
// window position (0..8) of each maximum — first maximum in row-major window order, strict '>' like ATen — so the
// backward is a gather over the <= 4 windows that contain an input pixel, with no atomics.
#include "common.cuh"
#include "pdl.cuh"
#include "vec.cuh"

namespace b200seg {
namespace {

__global__ void __launch_bounds__(256) maxpool_fwd_kernel(const __nv_bfloat16* __restrict__ x, int x_ct, int x_off, int h,
                                                           int w, __nv_bfloat16* __restrict__ y, int y_ct, int y_off,
                                                           uint8_t* __restrict__ idx, int oh, int ow, int groups,
                                                           int64_t total) {
    pdl_prologue();
    for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
         i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        const Idx4 id = decode4(i, groups, ow, oh, total);
        const int g = id.g, ox = id.x, oy = id.y;
        const int64_t n = id.n;
        float best[8];
        int bi[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            best[j] = -INFINITY;
            bi[j] = 0;
        }
#pragma unroll
        for (int r = 0; r < 3; ++r) {
            const int iy = 2 * oy - 1 + r;
            if (iy < 0 || iy >= h) continue;
#pragma unroll
            for (int s = 0; s < 3; ++s) {
                const int ix = 2 * ox - 1 + s;
                if (ix < 0 || ix >= w) continue;
                float f[8];
                unpack8(ldg128(x + ((n * h + iy) * w + ix) * x_ct + x_off + g * 8), f);
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    if (f[j] > best[j] || f[j] != f[j]) {
                        best[j] = f[j];
                        bi[j] = r * 3 + s;
                    }
                }
            }
        }
        const int64_t op = (n * oh + oy) * ow + ox;
        st128(y + op * y_ct + y_off + g * 8, pack8(best));
        uint2 packed;
        packed.x = bi[0] | (bi[1] << 8) | (bi[2] << 16) | (bi[3] << 24);
        packed.y = bi[4] | (bi[5] << 8) | (bi[6] << 16) | (bi[7] << 24);
        *reinterpret_cast<uint2*>(idx + (op * groups + g) * 8) = packed;
    }
}

__global__ void __launch_bounds__(256) maxpool_bwd_kernel(const __nv_bfloat16* __restrict__ dy, int dy_ct, int dy_off,
                                                           const uint8_t* __restrict__ idx, int oh, int ow,
                                                           __nv_bfloat16* __restrict__ dx, int dx_ct, int dx_off, int h,
                                                           int w, int groups, int64_t total) {
    pdl_prologue();
    for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
         i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        const Idx4 id = decode4(i, groups, w, h, total);
        const int g = id.g, ix = id.x, iy = id.y;
        const int64_t n = id.n;
        float acc[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = 0.f;
        // windows o with 2o-1 <= i <= 2o+1: o = i/2 (tap 1) for even i; o = (i-1)/2 (tap 2) and (i+1)/2 (tap 0) for odd i.
        // All (up to four) window reads are issued before any is consumed: the dependent-latency chain of the looped
        // form kept this kernel at a quarter of the HBM rate.
        const int oy0 = iy >> 1, oy1 = (iy + 1) >> 1;
        const int ox0 = ix >> 1, ox1 = (ix + 1) >> 1;
        uint2 packed[4];
        uint4 dv[4];
        int tapv[4];
#pragma unroll
        for (int a = 0; a < 2; ++a)
#pragma unroll
            for (int b = 0; b < 2; ++b) {
                const int oy = a ? oy1 : oy0, ox = b ? ox1 : ox0;
                const bool ok = (a == 0 || oy1 != oy0) && (b == 0 || ox1 != ox0) && oy < oh && ox < ow;
                tapv[a * 2 + b] = ok ? (iy - (2 * oy - 1)) * 3 + (ix - (2 * ox - 1)) : -1;
                packed[a * 2 + b] = make_uint2(0u, 0u);
                dv[a * 2 + b] = make_uint4(0u, 0u, 0u, 0u);
                if (ok) {
                    const int64_t op = (n * oh + oy) * ow + ox;
                    packed[a * 2 + b] = __ldg(reinterpret_cast<const uint2*>(idx + (op * groups + g) * 8));
                    dv[a * 2 + b] = ldg128(dy + op * dy_ct + dy_off + g * 8);
                }
            }
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            float d[8];
            unpack8(dv[q], d);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const int sel = ((j < 4 ? packed[q].x : packed[q].y) >> ((j & 3) * 8)) & 0xff;
                if (sel == tapv[q]) acc[j] += d[j];
            }
        }
        st128(dx + ((n * h + iy) * w + ix) * dx_ct + dx_off + g * 8, pack8(acc));
    }
}

}  // namespace
}  // namespace b200seg

extern "C" {

int b200seg_maxpool3x3s2_fwd(const void* x, int x_c_total, int x_c_offset, int n, int h, int w, void* y, int y_c_total,
                             int y_c_offset, uint8_t* idx, int c, b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    if (check_slice("maxpool_fwd(x)", x, x_c_total, x_c_offset, c)) return 1;
    if (check_slice("maxpool_fwd(y)", y, y_c_total, y_c_offset, c)) return 1;
    B200SEG_CHECK(idx && n > 0 && h > 0 && w > 0, "maxpool_fwd: bad arguments");
    B200SEG_CHECK((reinterpret_cast<uintptr_t>(idx) & 7) == 0, "maxpool_fwd: index buffer not 8-byte aligned");
    const int oh = (h + 1) / 2, ow = (w + 1) / 2;
    const int64_t total = static_cast<int64_t>(n) * oh * ow * (c / 8);
    B200SEG_CUDA(launch_pdl(maxpool_fwd_kernel, elementwise_grid(total), 256, 0, static_cast<cudaStream_t>(stream), 
        static_cast<const __nv_bfloat16*>(x), x_c_total, x_c_offset, h, w, static_cast<__nv_bfloat16*>(y), y_c_total,
        y_c_offset, idx, oh, ow, c / 8, total));
    return 0;
}

int b200seg_maxpool3x3s2_bwd(const void* dy, int dy_c_total, int dy_c_offset, const uint8_t* idx, int n, int h, int w,
                             void* dx, int dx_c_total, int dx_c_offset, int c, b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    if (check_slice("maxpool_bwd(dy)", dy, dy_c_total, dy_c_offset, c)) return 1;
    if (check_slice("maxpool_bwd(dx)", dx, dx_c_total, dx_c_offset, c)) return 1;
    B200SEG_CHECK(idx && n > 0 && h > 0 && w > 0, "maxpool_bwd: bad arguments");
    const int oh = (h + 1) / 2, ow = (w + 1) / 2;
    const int64_t total = static_cast<int64_t>(n) * h * w * (c / 8);
    B200SEG_CUDA(launch_pdl(maxpool_bwd_kernel, elementwise_grid(total), 256, 0, static_cast<cudaStream_t>(stream), 
        static_cast<const __nv_bfloat16*>(dy), dy_c_total, dy_c_offset, idx, oh, ow, static_cast<__nv_bfloat16*>(dx),
        dx_c_total, dx_c_offset, h, w, c / 8, total));
    return 0;
}

}  // extern "C"
