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


// ---- space-to-depth form of a stride-2 few-channel conv -------------------------------------------------------
// A stride-2 R x S conv over C (<= 4) channels is a stride-1 ceil(R/2) x ceil(S/2) conv over the 2x2 space-to-depth
// transform of the padded image:   r = 2a + b,  s = 2a' + b'
//     y[oy][ox][k] = sum_{a, a'} sum_{b, b', ch} Z[oy + a][ox + a'][(2b + b')*C + ch] * w[k][ch][2a + b][2a' + b']
//     Z[Y][X][(2b + b')*C + ch] = x[ch][2Y + b - pad][2X + b' - pad]                (0 outside the image)
// Z is an NHWC bf16 tensor with cs = round_up(4C, 8) channels, 1/4 of the pixels of the image (for ResNet's 7x7 stem
// 34 MB at 16 x 512 x 512, against 319 MB of gathered patches).  The a' positions of one filter row are CONTIGUOUS in
// memory — ceil(S/2) * cs elements starting at pixel (oy + a, ox) — so the tcgen05 implicit GEMM reads them as one
// K block per row tap through an anchored window source (include/b200seg.h): nothing is gathered.
__global__ void __launch_bounds__(256) space_to_depth_kernel(const float* __restrict__ x, int C, int H, int W, int pad,
                                                              int ZH, int ZW, int cs, __nv_bfloat16* __restrict__ z,
                                                              long long total) {
    pdl_prologue();
    const long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
    if (i >= total) return;
    const int X = static_cast<int>(i % ZW);
    const int Y = static_cast<int>((i / ZW) % ZH);
    const int n = static_cast<int>(i / (static_cast<long long>(ZW) * ZH));
    const float* xn = x + static_cast<size_t>(n) * C * H * W;
    __align__(16) __nv_bfloat16 v[16];
#pragma unroll
    for (int q = 0; q < 16; ++q) v[q] = __float2bfloat16(0.f);
#pragma unroll
    for (int b = 0; b < 2; ++b)
#pragma unroll
        for (int bb = 0; bb < 2; ++bb) {
            const int iy = 2 * Y + b - pad, ix = 2 * X + bb - pad;
            if (iy >= 0 && iy < H && ix >= 0 && ix < W)
                for (int ch = 0; ch < C; ++ch)
                    v[(2 * b + bb) * C + ch] = __float2bfloat16(__ldg(xn + (static_cast<size_t>(ch) * H + iy) * W + ix));
        }
    uint4* out = reinterpret_cast<uint4*>(z + static_cast<size_t>(i) * cs);
    out[0] = *reinterpret_cast<const uint4*>(v);
    if (cs == 16) out[1] = *reinterpret_cast<const uint4*>(v + 8);
}

// wp[a][kk][a'*cs + (2b + b')*C + ch] = w[kk][ch][2a + b][2a' + b']     (bf16; 0 for taps past R / S and padding lanes)
__global__ void pack_weights_s2d_kernel(const float* __restrict__ w, int k, int C, int R, int S, int k_rows, int cs,
                                        __nv_bfloat16* __restrict__ wp) {
    pdl_prologue();
    const int A = (R + 1) / 2, A2 = (S + 1) / 2, cw = A2 * cs;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= A * k_rows * cw) return;
    const int e = i % cw, kk = (i / cw) % k_rows, a = i / (cw * k_rows);
    const int a2 = e / cs, q = e % cs;
    float v = 0.f;
    if (kk < k && q < 4 * C) {
        const int ch = q % C, bb = (q / C) & 1, b = q / (2 * C);
        const int r = 2 * a + b, sx = 2 * a2 + bb;
        if (r < R && sx < S) v = w[((static_cast<size_t>(kk) * C + ch) * R + r) * S + sx];
    }
    wp[i] = __float2bfloat16(v);
}

// dw[kk][ch][r][s] (+)= dwin[kk][a'*cs + (2b + b')*C + ch][a]       (dwin = the window-weight gradient [k][cw][A][1])
__global__ void unpack_wgrad_s2d_kernel(const float* __restrict__ dwin, int k, int C, int R, int S, int cs,
                                        float* __restrict__ dw, int accumulate) {
    pdl_prologue();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= k * C * R * S) return;
    const int sx = i % S, r = (i / S) % R, ch = (i / (R * S)) % C, kk = i / (R * S * C);
    const int A = (R + 1) / 2, cw = ((S + 1) / 2) * cs;
    const int e = (sx >> 1) * cs + (2 * (r & 1) + (sx & 1)) * C + ch;
    const float v = dwin[(static_cast<size_t>(kk) * cw + e) * A + (r >> 1)];
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

int b200seg_space_to_depth_nchw(const float* x, int n, int c, int h, int w, int r, int s, int pad, void* z, int cs,
                                b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    B200SEG_CHECK(x && z, "space_to_depth: null pointer");
    B200SEG_CHECK(n > 0 && c > 0 && c <= 4 && h > 0 && w > 0 && r > 0 && s > 0 && pad >= 0, "space_to_depth: bad geometry");
    B200SEG_CHECK(cs == (4 * c + 7) / 8 * 8, "space_to_depth: cs %d must be 4*c rounded up to a multiple of 8", cs);
    B200SEG_CHECK(h + 2 * pad >= r && w + 2 * pad >= s, "space_to_depth: empty output");
    B200SEG_CHECK((reinterpret_cast<uintptr_t>(z) & 15) == 0, "space_to_depth: output not 16-byte aligned");
    const int oh = (h + 2 * pad - r) / 2 + 1, ow = (w + 2 * pad - s) / 2 + 1;
    const int zh = oh + (r + 1) / 2 - 1, zw = ow + (s + 1) / 2 - 1;
    const long long total = static_cast<long long>(n) * zh * zw;
    B200SEG_CUDA(launch_pdl(space_to_depth_kernel, static_cast<unsigned>((total + 255) / 256), 256, 0,
                            static_cast<cudaStream_t>(stream), x, c, h, w, pad, zh, zw, cs, static_cast<__nv_bfloat16*>(z), total));
    return 0;
}

int b200seg_pack_weights_s2d(const float* w, int k, int c, int r, int s, void* wp, int k_rows, int cs,
                             b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    B200SEG_CHECK(w && wp, "pack_weights_s2d: null pointer");
    B200SEG_CHECK(k > 0 && c > 0 && c <= 4 && r > 0 && s > 0 && k_rows >= k && cs == (4 * c + 7) / 8 * 8,
                  "pack_weights_s2d: bad geometry");
    const int total = ((r + 1) / 2) * k_rows * ((s + 1) / 2) * cs;
    B200SEG_CUDA(launch_pdl(pack_weights_s2d_kernel, ceil_div(total, 256), 256, 0, static_cast<cudaStream_t>(stream), w, k, c, r,
                            s, k_rows, cs, static_cast<__nv_bfloat16*>(wp)));
    return 0;
}

int b200seg_unpack_wgrad_s2d(const float* dwin, int k, int c, int r, int s, int cs, float* dw, int accumulate,
                             b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    B200SEG_CHECK(dwin && dw, "unpack_wgrad_s2d: null pointer");
    B200SEG_CHECK(k > 0 && c > 0 && c <= 4 && r > 0 && s > 0 && cs == (4 * c + 7) / 8 * 8, "unpack_wgrad_s2d: bad geometry");
    const int total = k * c * r * s;
    B200SEG_CUDA(launch_pdl(unpack_wgrad_s2d_kernel, ceil_div(total, 256), 256, 0, static_cast<cudaStream_t>(stream), dwin, k,
                            c, r, s, cs, dw, accumulate));
    return 0;
}

}  // extern "C"
