// Patch gather ("im2col") for convolutions whose input has too few channels for a 64-wide K block — the 3-channel
// stems (ResNet-50's 7x7 stride-2 conv1 behind models/deeplabv3plus.py:17; HRNet's ConvNormAct(3, 64, 3, 2),
// models/hrnet.py:260): the fp32 NCHW image is gathered ONCE into a bf16 matrix
//     col[n][oy][ox][(r*S + s)*C + c] = x[n][c][oy*stride + r - pad][ox*stride + s - pad]      (0 outside the image)
// of row length c_pad (R*S*C rounded up to a multiple of 8, zero padded), i.e. an NHWC tensor with c_pad channels
// that the tcgen05 1x1 implicit GEMM and the tcgen05 wgrad consume like any other activation.  The same pass does the
// fp32 -> bf16 conversion and the NCHW -> NHWC layout change of the input.
//
// One CTA = 64 consecutive output pixels of one output row: the R x ((64-1)*stride + S) x C input window is staged in
// shared memory with loads coalesced along W, then written out as 16-byte chunks (8 col channels) by consecutive
// threads — the col rows of neighbouring pixels are contiguous, so the stores are fully coalesced.
#include "common.cuh"
#include "pdl.cuh"
#include "vec.cuh"

namespace b200seg {
namespace {

constexpr int kTW = 64;          // output pixels per CTA
constexpr int kMaxCol = 512;     // LUT capacity (R*S*C rounded up)

__global__ void __launch_bounds__(256) im2col_kernel(const float* __restrict__ x, int C, int H, int W, int R, int S,
                                                      int stride, int pad, int OH, int OW, int c_pad,
                                                      __nv_bfloat16* __restrict__ col) {
    pdl_prologue();
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int cols = (kTW - 1) * stride + S;
    int* lut = reinterpret_cast<int*>(smem_raw);                                   // [c_pad]
    __nv_bfloat16* tile = reinterpret_cast<__nv_bfloat16*>(smem_raw + kMaxCol * sizeof(int));   // [C][R][cols]
    const int n = blockIdx.z, oy = blockIdx.y, ox0 = blockIdx.x * kTW;
    const int iy0 = oy * stride - pad, ix0 = ox0 * stride - pad;
    for (int j = threadIdx.x; j < c_pad; j += blockDim.x) {
        int off = -1;
        if (j < R * S * C) {
            const int c = j % C, tap = j / C;
            off = (c * R + tap / S) * cols + tap % S;
        }
        lut[j] = off;
    }
    const float* xn = x + static_cast<size_t>(n) * C * H * W;
    for (int i = threadIdx.x; i < C * R * cols; i += blockDim.x) {
        const int xx = i % cols, rr = (i / cols) % R, c = i / (cols * R);
        const int iy = iy0 + rr, ix = ix0 + xx;
        float v = 0.f;
        if (iy >= 0 && iy < H && ix >= 0 && ix < W) v = __ldg(xn + (static_cast<size_t>(c) * H + iy) * W + ix);
        tile[i] = __float2bfloat16(v);
    }
    __syncthreads();
    const int chunks = c_pad >> 3;
    const int npix = min(kTW, OW - ox0);
    __nv_bfloat16* out = col + ((static_cast<size_t>(n) * OH + oy) * OW + ox0) * c_pad;
    const unsigned short* t16 = reinterpret_cast<const unsigned short*>(tile);
    for (int i = threadIdx.x; i < npix * chunks; i += blockDim.x) {
        const int px = i / chunks, q = i - px * chunks;
        const int base = px * stride;
        unsigned int w[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const int o0 = lut[q * 8 + 2 * e], o1 = lut[q * 8 + 2 * e + 1];
            const unsigned int lo = o0 >= 0 ? t16[o0 + base] : 0u;
            const unsigned int hi = o1 >= 0 ? t16[o1 + base] : 0u;
            w[e] = lo | (hi << 16);
        }
        *reinterpret_cast<uint4*>(out + static_cast<size_t>(i) * 8) = make_uint4(w[0], w[1], w[2], w[3]);
    }
}

// wf[kk][(r*S+s)*C + c] = w[kk][c][r][s]  (bf16, rows padded to c_pad, k padded to k_rows)
__global__ void pack_weights_im2col_kernel(const float* __restrict__ w, int k, int C, int R, int S, int k_rows, int c_pad,
                                           __nv_bfloat16* __restrict__ wf) {
    pdl_prologue();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= k_rows * c_pad) return;
    const int j = i % c_pad, kk = i / c_pad;
    float v = 0.f;
    if (kk < k && j < R * S * C) {
        const int c = j % C, tap = j / C;
        v = w[(static_cast<size_t>(kk) * C + c) * R * S + tap];
    }
    wf[i] = __float2bfloat16(v);
}

// dw[kk][c][r][s] (+)= dwcol[kk][(r*S+s)*C + c]
__global__ void unpack_wgrad_im2col_kernel(const float* __restrict__ dwcol, int k, int C, int R, int S, int c_pad,
                                           float* __restrict__ dw, int accumulate) {
    pdl_prologue();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= k * C * R * S) return;
    const int tap = i % (R * S), c = (i / (R * S)) % C, kk = i / (R * S * C);
    const float v = dwcol[static_cast<size_t>(kk) * c_pad + tap * C + c];
    dw[i] = accumulate ? dw[i] + v : v;
}

}  // namespace
}  // namespace b200seg

extern "C" {

int b200seg_im2col_nchw(const float* x, int n, int c, int h, int w, int r, int s, int stride, int pad, void* col,
                        int c_pad, b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    B200SEG_CHECK(x && col, "im2col: null pointer");
    B200SEG_CHECK(n > 0 && c > 0 && h > 0 && w > 0 && r > 0 && s > 0 && stride > 0 && pad >= 0, "im2col: bad geometry");
    B200SEG_CHECK(c_pad % 8 == 0 && c_pad >= r * s * c && c_pad <= kMaxCol,
                  "im2col: c_pad %d must be a multiple of 8 in [%d, %d]", c_pad, r * s * c, kMaxCol);
    B200SEG_CHECK((reinterpret_cast<uintptr_t>(col) & 15) == 0, "im2col: output not 16-byte aligned");
    const int oh = (h + 2 * pad - r) / stride + 1, ow = (w + 2 * pad - s) / stride + 1;
    B200SEG_CHECK(oh > 0 && ow > 0, "im2col: empty output");
    const int cols = (kTW - 1) * stride + s;
    const size_t smem = kMaxCol * sizeof(int) + static_cast<size_t>(c) * r * cols * 2;
    B200SEG_CHECK(smem <= 48 * 1024, "im2col: window of %zu bytes exceeds shared memory", smem);
    B200SEG_CHECK(oh <= 65535 && n <= 65535, "im2col: grid too large");
    dim3 grid(ceil_div(ow, kTW), oh, n);
    B200SEG_CUDA(launch_pdl(im2col_kernel, grid, 256, smem, static_cast<cudaStream_t>(stream), x, c, h, w, r, s, stride, pad, oh, ow, c_pad,
                                                                         static_cast<__nv_bfloat16*>(col)));
    return 0;
}

int b200seg_pack_weights_im2col(const float* w, int k, int c, int r, int s, void* wf, int k_rows, int c_pad,
                                b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    B200SEG_CHECK(w && wf, "pack_weights_im2col: null pointer");
    B200SEG_CHECK(k > 0 && k_rows >= k && c_pad % 8 == 0 && c_pad >= r * s * c, "pack_weights_im2col: bad sizes");
    const int total = k_rows * c_pad;
    B200SEG_CUDA(launch_pdl(pack_weights_im2col_kernel, ceil_div(total, 256), 256, 0, static_cast<cudaStream_t>(stream), 
        w, k, c, r, s, k_rows, c_pad, static_cast<__nv_bfloat16*>(wf)));
    return 0;
}

int b200seg_unpack_wgrad_im2col(const float* dwcol, int k, int c, int r, int s, int c_pad, float* dw, int accumulate,
                                b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    B200SEG_CHECK(dwcol && dw, "unpack_wgrad_im2col: null pointer");
    B200SEG_CHECK(k > 0 && c_pad >= r * s * c, "unpack_wgrad_im2col: bad sizes");
    const int total = k * c * r * s;
    B200SEG_CUDA(launch_pdl(unpack_wgrad_im2col_kernel, ceil_div(total, 256), 256, 0, static_cast<cudaStream_t>(stream), dwcol, k, c, r, s, c_pad,
                                                                                                   dw, accumulate));
    return 0;
}

}  // extern "C"
