// Input side of the hot path: the reference normalises on the host / with three elementwise torch ops
// (utils/datasets.py:199-213: imgs.float(); imgs -= mean; imgs /= std) and ships 4-byte pixels to the stem.  Here the
// 1-byte pixels the dataset yields (utils/datasets.py:161-194) go to the device as they are and ONE kernel produces what
// the stem's tcgen05 GEMM reads: normalised, rounded to bf16, NHWC, 2x2 space-to-depth (csrc/im2col.cu has the fp32 form).
// Bit-exactness: (float(x) - mean) / std with __fsub_rn / __fdiv_rn is the reference's fp32 arithmetic, operation for
// operation, so the fp32 result equals post_fetch_fn's and the bf16 one its round-to-nearest-even.
#include "common.cuh"
#include "pdl.cuh"
#include "vec.cuh"

namespace b200seg {
namespace {

struct Norm4 {
    float mean[4], std[4];
};

__device__ __forceinline__ float norm1(unsigned int v, float mean, float std) {
    return __fdiv_rn(__fsub_rn(static_cast<float>(v), mean), std);
}

// NCHW: consecutive threads on consecutive groups of 4 pixels of a row (32-bit loads, 128-bit stores)
__global__ void __launch_bounds__(256) input_u8_normalize_nchw_kernel(const uint8_t* __restrict__ x, int C, int64_t hw,
                                                                       int64_t total4, const Norm4 nm,
                                                                       float* __restrict__ y) {
    pdl_prologue();
    const int64_t hw4 = hw >> 2;
    for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total4;
         i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        const int ch = static_cast<int>((i / hw4) % C);
        const unsigned int v = __ldcs(reinterpret_cast<const unsigned int*>(x) + i);
        const float m = nm.mean[ch], s = nm.std[ch];
        __stcs(reinterpret_cast<float4*>(y) + i, make_float4(norm1(v & 255u, m, s), norm1((v >> 8) & 255u, m, s),
                                                             norm1((v >> 16) & 255u, m, s), norm1(v >> 24, m, s)));
    }
}

// any layout / size: one element per thread
__global__ void __launch_bounds__(256) input_u8_normalize_kernel(const uint8_t* __restrict__ x, int nhwc, int C, int H, int W,
                                                                  int64_t total, const Norm4 nm, float* __restrict__ y) {
    pdl_prologue();
    for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
         i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        // i enumerates the NCHW output
        const int xw = static_cast<int>(i % W);
        const int yh = static_cast<int>((i / W) % H);
        const int ch = static_cast<int>((i / (static_cast<int64_t>(W) * H)) % C);
        const int64_t n = i / (static_cast<int64_t>(W) * H * C);
        const int64_t src = nhwc ? ((n * H + yh) * W + xw) * C + ch : i;
        y[i] = norm1(x[src], nm.mean[ch], nm.std[ch]);
    }
}

// z[n][Y][X][(2b + b')*C + ch] = bf16(norm(x[n][ch][2Y + b - pad][2X + b' - pad]))   (0 outside the image / padding lanes)
__global__ void __launch_bounds__(256) input_u8_s2d_kernel(const uint8_t* __restrict__ x, int nhwc, int C, int H, int W,
                                                            int pad, int ZH, int ZW, int cs, const Norm4 nm,
                                                            __nv_bfloat16* __restrict__ z, long long total) {
    pdl_prologue();
    const long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
    if (i >= total) return;
    const int X = static_cast<int>(i % ZW);
    const int Y = static_cast<int>((i / ZW) % ZH);
    const int n = static_cast<int>(i / (static_cast<long long>(ZW) * ZH));
    const uint8_t* xn = x + static_cast<size_t>(n) * C * H * W;
    __align__(16) __nv_bfloat16 v[16];
#pragma unroll
    for (int q = 0; q < 16; ++q) v[q] = __float2bfloat16(0.f);
#pragma unroll
    for (int b = 0; b < 2; ++b)
#pragma unroll
        for (int bb = 0; bb < 2; ++bb) {
            const int iy = 2 * Y + b - pad, ix = 2 * X + bb - pad;
            if (iy >= 0 && iy < H && ix >= 0 && ix < W)
                for (int ch = 0; ch < C; ++ch) {
                    const size_t src = nhwc ? (static_cast<size_t>(iy) * W + ix) * C + ch
                                            : (static_cast<size_t>(ch) * H + iy) * W + ix;
                    v[(2 * b + bb) * C + ch] = __float2bfloat16(norm1(__ldg(xn + src), nm.mean[ch], nm.std[ch]));
                }
        }
    uint4* out = reinterpret_cast<uint4*>(z + static_cast<size_t>(i) * cs);
    out[0] = *reinterpret_cast<const uint4*>(v);
    if (cs == 16) out[1] = *reinterpret_cast<const uint4*>(v + 8);
}

// multi_scale (utils/datasets.py:206-212): F.interpolate(imgs, (h, w)), mode 'nearest', on the 1-byte planes — nearest
// sampling commutes with the per-pixel normalisation, so resizing the uint8 image first gives the reference's float
// result bit for bit.  Source index as ATen computes it (UpSample.h nearest_neighbor_compute_source_index):
// min(int(floorf(dst * float(in) / out)), in - 1) in fp32.  One thread per 4 output bytes of a row.
__global__ void __launch_bounds__(256) resize_nearest_u8_kernel(const uint8_t* __restrict__ x, int H, int W,
                                                                 uint8_t* __restrict__ y, int OH, int OW, int ow4,
                                                                 float sh, float sw, int64_t total) {
    pdl_prologue();
    for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
         i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        const int q = static_cast<int>(i % ow4);
        const int64_t row = i / ow4;                     // plane * OH + oy
        const int oy = static_cast<int>(row % OH);
        const int64_t plane = row / OH;
        const int iy = min(static_cast<int>(floorf(oy * sh)), H - 1);
        const uint8_t* src = x + (plane * H + iy) * W;
        uint8_t* dst = y + row * OW + q * 4;
        uint8_t v[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int ox = q * 4 + k;
            v[k] = ox < OW ? __ldg(src + min(static_cast<int>(floorf(ox * sw)), W - 1)) : 0;
        }
        if (q * 4 + 4 <= OW && (reinterpret_cast<uintptr_t>(dst) & 3) == 0) {
            *reinterpret_cast<uchar4*>(dst) = make_uchar4(v[0], v[1], v[2], v[3]);
        } else {
            for (int k = 0; k < 4 && q * 4 + k < OW; ++k) dst[k] = v[k];
        }
    }
}

int fill_norm(Norm4* nm, int c, const float* mean, const float* std, const char* what) {
    B200SEG_CHECK(mean && std, "%s: null mean / std", what);
    for (int i = 0; i < 4; ++i) {
        nm->mean[i] = i < c ? mean[i] : 0.f;
        nm->std[i] = i < c ? std[i] : 1.f;
        B200SEG_CHECK(i >= c || std[i] != 0.f, "%s: std[%d] is zero", what, i);
    }
    return 0;
}

}  // namespace
}  // namespace b200seg

extern "C" {

int b200seg_input_u8_normalize(const uint8_t* x, int layout, int n, int c, int h, int w, const float* mean, const float* std,
                               float* y, b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    B200SEG_CHECK(x && y, "input_u8_normalize: null pointer");
    B200SEG_CHECK(n > 0 && c > 0 && c <= 4 && h > 0 && w > 0 && (layout == 0 || layout == 1),
                  "input_u8_normalize: bad geometry (n %d c %d h %d w %d layout %d; c <= 4)", n, c, h, w, layout);
    Norm4 nm;
    if (fill_norm(&nm, c, mean, std, "input_u8_normalize")) return 1;
    const int64_t hw = static_cast<int64_t>(h) * w, total = hw * c * n;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (layout == 0 && hw % 4 == 0 && (reinterpret_cast<uintptr_t>(x) & 3) == 0 && (reinterpret_cast<uintptr_t>(y) & 15) == 0) {
        B200SEG_CUDA(launch_pdl(input_u8_normalize_nchw_kernel, elementwise_grid(total / 4), 256, 0, st, x, c, hw, total / 4, nm, y));
    } else {
        B200SEG_CUDA(launch_pdl(input_u8_normalize_kernel, elementwise_grid(total), 256, 0, st, x, layout, c, h, w, total, nm, y));
    }
    return 0;
}

int b200seg_input_u8_s2d(const uint8_t* x, int layout, int n, int c, int h, int w, int r, int s, int pad, const float* mean,
                         const float* std, void* z, int cs, b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    B200SEG_CHECK(x && z, "input_u8_s2d: null pointer");
    B200SEG_CHECK(n > 0 && c > 0 && c <= 4 && h > 0 && w > 0 && r > 0 && s > 0 && pad >= 0 && (layout == 0 || layout == 1),
                  "input_u8_s2d: bad geometry");
    B200SEG_CHECK(cs == (4 * c + 7) / 8 * 8, "input_u8_s2d: cs %d must be 4*c rounded up to a multiple of 8", cs);
    B200SEG_CHECK(h + 2 * pad >= r && w + 2 * pad >= s, "input_u8_s2d: empty output");
    B200SEG_CHECK((reinterpret_cast<uintptr_t>(z) & 15) == 0, "input_u8_s2d: output not 16-byte aligned");
    Norm4 nm;
    if (fill_norm(&nm, c, mean, std, "input_u8_s2d")) return 1;
    const int oh = (h + 2 * pad - r) / 2 + 1, ow = (w + 2 * pad - s) / 2 + 1;
    const int zh = oh + (r + 1) / 2 - 1, zw = ow + (s + 1) / 2 - 1;
    const long long total = static_cast<long long>(n) * zh * zw;
    B200SEG_CUDA(launch_pdl(input_u8_s2d_kernel, static_cast<unsigned>((total + 255) / 256), 256, 0,
                            static_cast<cudaStream_t>(stream), x, layout, c, h, w, pad, zh, zw, cs, nm,
                            static_cast<__nv_bfloat16*>(z), total));
    return 0;
}

int b200seg_resize_nearest_u8(const uint8_t* x, int64_t planes, int h, int w, uint8_t* y, int oh, int ow,
                              b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    B200SEG_CHECK(x && y, "resize_nearest_u8: null pointer");
    B200SEG_CHECK(planes > 0 && h > 0 && w > 0 && oh > 0 && ow > 0, "resize_nearest_u8: bad geometry");
    const int ow4 = (ow + 3) / 4;
    const int64_t total = planes * oh * ow4;
    // ATen: scale = float(in) / out, computed in fp32
    const float sh = static_cast<float>(h) / static_cast<float>(oh), sw = static_cast<float>(w) / static_cast<float>(ow);
    B200SEG_CUDA(launch_pdl(resize_nearest_u8_kernel, elementwise_grid(total), 256, 0, static_cast<cudaStream_t>(stream),
                            x, h, w, y, oh, ow, ow4, sh, sw, total));
    return 0;
}

}  // extern "C"
