// 3x3 stride-2 pad-1 max pooling of the ResNet stem (backbones.resnet50, used at models/deeplabv3plus.py:17-19,29) on
// NHWC bf16.  HBM-bound: one thread per (pixel, 8-channel group), 128-bit accesses.  The forward also stores the
// window position (0..8) of each maximum — first maximum in row-major window order, strict '>' like ATen — so the
// backward is a gather over the <= 4 windows that contain an input pixel, with no atomics (one thread per 2 x 2
// input block and channel group).
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

// dx for the 2 x 2 input block (2py..2py+1, 2px..2px+1) of one channel group.  Input rows / columns of even index lie
// in one window (centre tap), odd ones in two (last tap of window p, first tap of window p + 1), so the block reads the
// four windows (py..py+1, px..px+1) once — 4 gradient + 4 position loads per 4 outputs instead of 2.25 of each per
// output — and every (input, window) pair is one byte-wise compare of the stored positions against its tap number
// (__vcmpeq4), widened to a bf16 lane mask; control flow is uniform away from the borders.  The one-thread-per-input
// form spent ~250 instructions per 16 stored bytes and ran at a third of the HBM rate.
__device__ __forceinline__ void add_selected(const uint4& d, const uint2& pos, unsigned tap, float (&acc)[8]) {
    const unsigned pat = tap * 0x01010101u;
    const unsigned m0 = __vcmpeq4(pos.x, pat), m1 = __vcmpeq4(pos.y, pat);      // 0xff per matching channel
    const unsigned w[4] = {d.x & __byte_perm(m0, 0, 0x1100), d.y & __byte_perm(m0, 0, 0x3322),
                           d.z & __byte_perm(m1, 0, 0x1100), d.w & __byte_perm(m1, 0, 0x3322)};
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        acc[2 * q] += __uint_as_float(w[q] << 16);
        acc[2 * q + 1] += __uint_as_float(w[q] & 0xFFFF0000u);
    }
}

__global__ void __launch_bounds__(256) maxpool_bwd_kernel(const __nv_bfloat16* __restrict__ dy, int dy_ct, int dy_off,
                                                           const uint8_t* __restrict__ idx, int oh, int ow,
                                                           __nv_bfloat16* __restrict__ dx, int dx_ct, int dx_off, int h,
                                                           int w, int groups, int64_t total) {
    pdl_prologue();
    for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
         i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        const Idx4 id = decode4(i, groups, ow, oh, total);
        const int g = id.g, px = id.x, py = id.y;
        const int64_t n = id.n;
        const bool row1 = 2 * py + 1 < h, col1 = 2 * px + 1 < w;          // the odd row / column of the block exists
        const bool wy1 = py + 1 < oh, wx1 = px + 1 < ow;                  // the windows below / to the right exist
        // windows: q = a * 2 + b -> (py + a, px + b)
        uint2 pos[4];
        uint4 dv[4];
#pragma unroll
        for (int a = 0; a < 2; ++a)
#pragma unroll
            for (int b = 0; b < 2; ++b) {
                const int q = a * 2 + b;
                pos[q] = make_uint2(0xffffffffu, 0xffffffffu);            // matches no tap
                dv[q] = make_uint4(0u, 0u, 0u, 0u);
                if ((a == 0 || (wy1 && row1)) && (b == 0 || (wx1 && col1))) {
                    const int64_t op = (n * oh + py + a) * ow + px + b;
                    pos[q] = __ldg(reinterpret_cast<const uint2*>(idx + (op * groups + g) * 8));
                    dv[q] = ldg128(dy + op * dy_ct + dy_off + g * 8);
                }
            }
        float e00[8], e01[8], e10[8], e11[8];                             // (row parity, column parity) of the input
#pragma unroll
        for (int j = 0; j < 8; ++j) e00[j] = e01[j] = e10[j] = e11[j] = 0.f;
        // tap = r * 3 + s with r = iy - (2 oy - 1), s = ix - (2 ox - 1)
        add_selected(dv[0], pos[0], 4, e00);
        add_selected(dv[0], pos[0], 5, e01);
        add_selected(dv[1], pos[1], 3, e01);
        add_selected(dv[0], pos[0], 7, e10);
        add_selected(dv[2], pos[2], 1, e10);
        add_selected(dv[0], pos[0], 8, e11);
        add_selected(dv[1], pos[1], 6, e11);
        add_selected(dv[2], pos[2], 2, e11);
        add_selected(dv[3], pos[3], 0, e11);
        __nv_bfloat16* o = dx + ((n * h + 2 * py) * w + 2 * px) * dx_ct + dx_off + g * 8;
        st128(o, pack8(e00));
        if (col1) st128(o + dx_ct, pack8(e01));
        if (row1) {
            o += static_cast<int64_t>(w) * dx_ct;
            st128(o, pack8(e10));
            if (col1) st128(o + dx_ct, pack8(e11));
        }
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
    const int64_t total = static_cast<int64_t>(n) * oh * ow * (c / 8);        // one thread per 2 x 2 input block
    B200SEG_CUDA(launch_pdl(maxpool_bwd_kernel, elementwise_grid(total), 256, 0, static_cast<cudaStream_t>(stream), 
        static_cast<const __nv_bfloat16*>(dy), dy_c_total, dy_c_offset, idx, oh, ow, static_cast<__nv_bfloat16*>(dx),
        dx_c_total, dx_c_offset, h, w, c / 8, total));
    return 0;
}

}  // extern "C"
