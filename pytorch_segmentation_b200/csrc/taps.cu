// Narrow-output 3x3 convolutions (cls_conv: models/deeplabv3plus.py:22,39 384 -> 21 classes; models/unet.py:23,51
// 88 -> nc) restated so the wide activation tensor is read ONCE instead of once per filter tap.
//
// With N = 21 (padded 24) output channels the per-tap implicit GEMM is bound by re-reading the 201 MB input nine times
// through L2 (fprop 230 us, wgrad 331 us measured) and its data gradient by a 2.7x zero-padded reduction dimension.
// Instead:
//   fprop   Z[p][tap*ncp + k] = sum_c X[p][c] * W[k][c][tap]        one 1x1 tcgen05 GEMM, N = 9*ncp = 216
//           y[p][k] = bias[k] + sum_tap Z[p + off(tap)][tap*ncp + k]  tap_scatter_add (this file), fp32 out
//   dgrad   col[p][tap*ncp + k] = dY[p + off(tap)][k]                 tap_gather (this file), 216 channels
//           dX[p][c] = sum_j col[p][j] * Wcol[c][j]                   one 1x1 tcgen05 GEMM, K = 216 dense
//   wgrad   dWcol[j][c] = sum_p col[p][j] * X[p][c]                   one 1x1 tcgen05 wgrad over the same col
//           dW[k][c][tap] = dWcol[mirror(tap)*ncp + k][c]             unpack (this file)
// off(tap) = ((r - R/2) * dil, (s - S/2) * dil); pixels outside the image contribute zero (the padding).
#include "common.cuh"
#include "pdl.cuh"
#include "vec.cuh"

namespace b200seg {
namespace {

// col[p][tap*c + ch] = x[p + off(tap)][ch]: one thread per (pixel, tap, 8-channel group), 16-byte copies
__global__ void __launch_bounds__(256) tap_gather_kernel(const __nv_bfloat16* __restrict__ x, int x_ct, int x_off, int c,
                                                          int H, int W, int R, int S, int dil,
                                                          __nv_bfloat16* __restrict__ col, int col_ct, int64_t total) {
    pdl_prologue();
    const int groups = c >> 3;
    const int per_pixel = R * S * groups;
    for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
         i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        int q, px, py;
        const int64_t p = divmod_small(i, per_pixel, total, q);
        const int tap = q / groups, g = q - tap * groups;
        divmod_small(divmod_small(p, W, total, px), H, total, py);
        const int sy = py + (tap / S - R / 2) * dil, sx = px + (tap % S - S / 2) * dil;
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (sy >= 0 && sy < H && sx >= 0 && sx < W) {
            const int64_t sp = p + static_cast<int64_t>(sy - py) * W + (sx - px);
            v = ldg128(x + sp * x_ct + x_off + g * 8);
        }
        st128(col + p * col_ct + tap * c + g * 8, v);
    }
}

// y[p][k] = bias[k] + sum_tap z[p + off(tap)][tap*ncp + k]: one thread per (pixel, 8-channel group), fp32 accumulate
__global__ void __launch_bounds__(256) tap_scatter_add_kernel(const __nv_bfloat16* __restrict__ z, int z_ct, int ncp, int H,
                                                               int W, int R, int S, int dil, const float* __restrict__ bias,
                                                               float* __restrict__ y, int64_t total) {
    pdl_prologue();
    const int groups = ncp >> 3;
    for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
         i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        int g, px, py;
        const int64_t p = divmod_small(i, groups, total, g);
        divmod_small(divmod_small(p, W, total, px), H, total, py);
        float acc[8];
        if (bias) {
            load8f(bias + g * 8, acc);
        } else {
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[j] = 0.f;
        }
        for (int tap = 0; tap < R * S; ++tap) {
            const int sy = py + (tap / S - R / 2) * dil, sx = px + (tap % S - S / 2) * dil;
            if (sy < 0 || sy >= H || sx < 0 || sx >= W) continue;
            const int64_t sp = p + static_cast<int64_t>(sy - py) * W + (sx - px);
            float f[8];
            unpack8(ldg128(z + sp * z_ct + tap * ncp + g * 8), f);
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[j] += f[j];
        }
        float4* dst = reinterpret_cast<float4*>(y + p * ncp + g * 8);
        dst[0] = make_float4(acc[0], acc[1], acc[2], acc[3]);
        dst[1] = make_float4(acc[4], acc[5], acc[6], acc[7]);
    }
}

// wz[tap*ncp + k][c] = w[k][c][tap];  wcol[c][tap*ncp + k] = w[k][c][mirror(tap)]   (bf16, zero padded)
__global__ void pack_weights_taps_kernel(const float* __restrict__ w, int k, int c, int taps, int ncp, int z_rows,
                                         __nv_bfloat16* __restrict__ wz, int c_rows, __nv_bfloat16* __restrict__ wcol) {
    pdl_prologue();
    const int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
    const int cols = taps * ncp;
    if (blockIdx.y == 0) {
        if (!wz || i >= static_cast<int64_t>(z_rows) * c) return;
        const int cc = static_cast<int>(i % c), j = static_cast<int>(i / c);
        const int tap = j / ncp, kk = j - tap * ncp;
        const float v = (j < cols && kk < k) ? w[(static_cast<size_t>(kk) * c + cc) * taps + tap] : 0.f;
        wz[i] = __float2bfloat16(v);
    } else {
        if (!wcol || i >= static_cast<int64_t>(c_rows) * cols) return;
        const int j = static_cast<int>(i % cols), cc = static_cast<int>(i / cols);
        const int tap = j / ncp, kk = j - tap * ncp;
        const float v = (cc < c && kk < k) ? w[(static_cast<size_t>(kk) * c + cc) * taps + (taps - 1 - tap)] : 0.f;
        wcol[i] = __float2bfloat16(v);
    }
}

// dw[k][c][tap] (+)= dwcol[mirror(tap)*ncp + k][c]
__global__ void unpack_wgrad_taps_kernel(const float* __restrict__ dwcol, int k, int c, int taps, int ncp,
                                         float* __restrict__ dw, int accumulate) {
    pdl_prologue();
    const int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
    if (i >= static_cast<int64_t>(k) * c * taps) return;
    const int tap = static_cast<int>(i % taps), cc = static_cast<int>((i / taps) % c), kk = static_cast<int>(i / (taps * c));
    const float v = dwcol[(static_cast<size_t>(taps - 1 - tap) * ncp + kk) * c + cc];
    dw[i] = accumulate ? dw[i] + v : v;
}

}  // namespace
}  // namespace b200seg

extern "C" {

int b200seg_tap_gather(const void* x, int x_c_total, int x_c_offset, int c, int n, int h, int w, int r, int s, int dil,
                       void* col, int col_c_total, b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    if (check_slice("tap_gather(x)", x, x_c_total, x_c_offset, c)) return 1;
    B200SEG_CHECK(col && n > 0 && h > 0 && w > 0 && r > 0 && s > 0 && dil > 0, "tap_gather: bad arguments");
    B200SEG_CHECK(col_c_total % 8 == 0 && col_c_total >= r * s * c, "tap_gather: col has %d channels, needs %d",
                  col_c_total, r * s * c);
    B200SEG_CHECK((reinterpret_cast<uintptr_t>(col) & 15) == 0, "tap_gather: output not 16-byte aligned");
    const int64_t total = static_cast<int64_t>(n) * h * w * r * s * (c / 8);
    B200SEG_CUDA(launch_pdl(tap_gather_kernel, elementwise_grid(total), 256, 0, static_cast<cudaStream_t>(stream),
                            static_cast<const __nv_bfloat16*>(x), x_c_total, x_c_offset, c, h, w, r, s, dil,
                            static_cast<__nv_bfloat16*>(col), col_c_total, total));
    return 0;
}

int b200seg_tap_scatter_add(const void* z, int z_c_total, int ncp, int n, int h, int w, int r, int s, int dil,
                            const float* bias, float* y, b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    B200SEG_CHECK(z && y && n > 0 && h > 0 && w > 0 && r > 0 && s > 0 && dil > 0, "tap_scatter_add: bad arguments");
    B200SEG_CHECK(ncp > 0 && ncp % 8 == 0 && z_c_total % 8 == 0 && z_c_total >= r * s * ncp,
                  "tap_scatter_add: z has %d channels, needs %d (ncp %d)", z_c_total, r * s * ncp, ncp);
    B200SEG_CHECK(((reinterpret_cast<uintptr_t>(z) | reinterpret_cast<uintptr_t>(y)) & 15) == 0,
                  "tap_scatter_add: pointers not 16-byte aligned");
    const int64_t total = static_cast<int64_t>(n) * h * w * (ncp / 8);
    B200SEG_CUDA(launch_pdl(tap_scatter_add_kernel, elementwise_grid(total), 256, 0, static_cast<cudaStream_t>(stream),
                            static_cast<const __nv_bfloat16*>(z), z_c_total, ncp, h, w, r, s, dil, bias, y, total));
    return 0;
}

int b200seg_pack_weights_taps(const float* w, int k, int c, int r, int s, int ncp, void* wz, int z_rows, void* wcol,
                              int c_rows, b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    B200SEG_CHECK(w && (wz || wcol) && k > 0 && c > 0 && r > 0 && s > 0, "pack_weights_taps: bad arguments");
    B200SEG_CHECK(ncp >= k && ncp % 8 == 0 && z_rows >= r * s * ncp && c_rows >= c, "pack_weights_taps: bad sizes");
    const int taps = r * s;
    const int64_t tz = static_cast<int64_t>(z_rows) * c, tc = static_cast<int64_t>(c_rows) * taps * ncp;
    const int64_t total = tz > tc ? tz : tc;
    B200SEG_CUDA(launch_pdl(pack_weights_taps_kernel, dim3(static_cast<unsigned>(ceil_div64(total, 256)), 2), 256, 0,
                            static_cast<cudaStream_t>(stream), w, k, c, taps, ncp, z_rows, static_cast<__nv_bfloat16*>(wz),
                            c_rows, static_cast<__nv_bfloat16*>(wcol)));
    return 0;
}

int b200seg_unpack_wgrad_taps(const float* dwcol, int k, int c, int r, int s, int ncp, float* dw, int accumulate,
                              b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    B200SEG_CHECK(dwcol && dw && k > 0 && c > 0 && r > 0 && s > 0 && ncp >= k, "unpack_wgrad_taps: bad arguments");
    const int64_t total = static_cast<int64_t>(k) * c * r * s;
    B200SEG_CUDA(launch_pdl(unpack_wgrad_taps_kernel, static_cast<unsigned>(ceil_div64(total, 256)), 256, 0,
                            static_cast<cudaStream_t>(stream), dwcol, k, c, r * s, ncp, dw, accumulate));
    return 0;
}

}  // extern "C"
