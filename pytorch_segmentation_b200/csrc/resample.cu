// Bilinear resampling between NHWC bf16 channel slices (decoder upsample + concat placement), its adjoint in
// gather form, and the H*W reductions / broadcasts of the ASPP image-pooling branch.  HBM-bound: one thread per
// (pixel, 8-channel group), 128-bit accesses, consecutive threads on consecutive channel groups.
#include "common.cuh"
#include "pdl.cuh"
#include "vec.cuh"

namespace b200seg {
namespace {

// A thread owns one output column (ox, 8-channel group) over a strip of kUpRows output rows: the horizontal
// interpolation of a source row is done once and reused by every output row that reads it (x4 upsampling: 2-3 source
// rows per 8 output rows), so an output pixel costs ~16 FMAs and a fraction of a 2-load row fetch instead of four
// loads and 24 FMAs — the one-thread-per-output form was issue-bound at a quarter of the HBM write rate.
constexpr int kUpRows = 8;

__global__ void __launch_bounds__(256) upsample_fwd_kernel(const __nv_bfloat16* __restrict__ x, int x_ct, int x_off, int h,
                                                            int w, __nv_bfloat16* __restrict__ y, int y_ct, int y_off,
                                                            int oh, int ow, int groups, int strips, int64_t total,
                                                            float sh, float sw, int ac) {
    pdl_prologue();
    for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
         i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        const Idx4 id = decode4(i, groups, ow, strips, total);
        const int g = id.g, ox = id.x, n = id.n;
        const int oy0 = id.y * kUpRows, oy1 = min(oy0 + kUpRows, oh);
        const Lerp lx = bilinear_src(sw, ox, w, ac);
        const __nv_bfloat16* b = x + (static_cast<int64_t>(n) * h * w) * x_ct + x_off + g * 8;
        const __nv_bfloat16* c0 = b + static_cast<int64_t>(lx.i0) * x_ct;
        const __nv_bfloat16* c1 = b + static_cast<int64_t>(lx.i1) * x_ct;
        const int64_t row_pitch = static_cast<int64_t>(w) * x_ct;
        auto lerp_row = [&](int r, float (&o)[8]) {
            float a[8], d[8];
            unpack8(ldg128(c0 + r * row_pitch), a);
            unpack8(ldg128(c1 + r * row_pitch), d);
#pragma unroll
            for (int j = 0; j < 8; ++j) o[j] = lx.w0 * a[j] + lx.w1 * d[j];
        };
        float top[8], bot[8];
        int r0 = -1, r1 = -1;                 // source rows held in top / bot
        __nv_bfloat16* yp = y + ((static_cast<int64_t>(n) * oh + oy0) * ow + ox) * y_ct + y_off + g * 8;
        const int64_t y_pitch = static_cast<int64_t>(ow) * y_ct;
        for (int oy = oy0; oy < oy1; ++oy, yp += y_pitch) {
            const Lerp ly = bilinear_src(sh, oy, h, ac);
            if (ly.i0 != r0) {
                if (ly.i0 == r1) {
#pragma unroll
                    for (int j = 0; j < 8; ++j) top[j] = bot[j];
                } else {
                    lerp_row(ly.i0, top);
                }
                r0 = ly.i0;
            }
            if (ly.i1 != r1) {
                if (ly.i1 == r0) {
#pragma unroll
                    for (int j = 0; j < 8; ++j) bot[j] = top[j];
                } else {
                    lerp_row(ly.i1, bot);
                }
                r1 = ly.i1;
            }
            float o[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) o[j] = ly.w0 * top[j] + ly.w1 * bot[j];
            st128(yp, pack8(o));
        }
    }
}

// Weight with which destination index `dst` reads source index `src_i`: the tent 1 - |src - src_i| around PyTorch's
// source coordinate, clamped to [0, in - 1] (the clamp reproduces bilinear_src's edge rule, where both taps fall on the
// last pixel and their weights add up to 1).
__device__ __forceinline__ float bilinear_weight(float scale, int dst, int src_i, int in, int align_corners) {
    float src = align_corners ? scale * dst : fmaxf(scale * (dst + 0.5f) - 0.5f, 0.f);
    src = fminf(src, static_cast<float>(in - 1));
    return fmaxf(1.f - fabsf(src - static_cast<float>(src_i)), 0.f);
}

// dx[n, y, x, :] = sum over destination pixels (oy, ox) that read (y, x) of weight * dy[n, oy, ox, :] — the general
// form (any ratio, any width): one thread per (source pixel, channel group) gathers its whole footprint.  Used when the
// separable kernel below does not apply (downsampling ratios, rows too wide for its shared-memory line).
__global__ void __launch_bounds__(256) upsample_bwd_gather_kernel(const __nv_bfloat16* __restrict__ dy, int dy_ct, int dy_off,
                                                            int oh, int ow, __nv_bfloat16* __restrict__ dx, int dx_ct,
                                                            int dx_off, int h, int w, int groups, int64_t total, float sh,
                                                            float sw, int ac) {
    pdl_prologue();
    for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
         i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        const Idx4 id = decode4(i, groups, w, h, total);
        const int g = id.g, sx = id.x, sy = id.y, n = id.n;
        int ylo, yhi, xlo, xhi;
        bilinear_dst_range(sh, sy, h, oh, ac, ylo, yhi);
        bilinear_dst_range(sw, sx, w, ow, ac, xlo, xhi);
        float acc[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = 0.f;
        const __nv_bfloat16* b = dy + (static_cast<int64_t>(n) * oh * ow) * dy_ct + dy_off + g * 8;
        // the range is a superset: trim the columns that do not read sx, then up to x4.5 a source column feeds <= 10
        auto col_weight = [&](int ox) {
            const Lerp lx = bilinear_src(sw, ox, w, ac);
            return (lx.i0 == sx ? lx.w0 : 0.f) + (lx.i1 == sx ? lx.w1 : 0.f);
        };
        while (xlo < xhi && col_weight(xlo) == 0.f) ++xlo;
        while (xhi > xlo && col_weight(xhi) == 0.f) --xhi;
        constexpr int kSpan = 10;
        if (xhi - xlo < kSpan) {
            // column weights once per element; per destination row all loads are issued before they are consumed
            float wxs[kSpan];
#pragma unroll
            for (int k = 0; k < kSpan; ++k) {
                const int ox = xlo + k;
                const Lerp lx = bilinear_src(sw, min(ox, ow - 1), w, ac);
                wxs[k] = ox <= xhi ? (lx.i0 == sx ? lx.w0 : 0.f) + (lx.i1 == sx ? lx.w1 : 0.f) : 0.f;
            }
            for (int oy = ylo; oy <= yhi; ++oy) {
                const Lerp ly = bilinear_src(sh, oy, h, ac);
                const float wy = (ly.i0 == sy ? ly.w0 : 0.f) + (ly.i1 == sy ? ly.w1 : 0.f);
                if (wy == 0.f) continue;
                const __nv_bfloat16* rowp = b + (static_cast<int64_t>(oy) * ow + xlo) * dy_ct;
                uint4 v[kSpan];
#pragma unroll
                for (int k = 0; k < kSpan; ++k)
                    v[k] = wxs[k] != 0.f ? ldg128(rowp + static_cast<int64_t>(k) * dy_ct) : make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
                for (int k = 0; k < kSpan; ++k) {
                    float f[8];
                    unpack8(v[k], f);
                    const float wt = wy * wxs[k];
#pragma unroll
                    for (int j = 0; j < 8; ++j) acc[j] = fmaf(wt, f[j], acc[j]);
                }
            }
        } else {
            for (int oy = ylo; oy <= yhi; ++oy) {
                const Lerp ly = bilinear_src(sh, oy, h, ac);
                const float wy = (ly.i0 == sy ? ly.w0 : 0.f) + (ly.i1 == sy ? ly.w1 : 0.f);
                if (wy == 0.f) continue;
                for (int ox = xlo; ox <= xhi; ++ox) {
                    const Lerp lx = bilinear_src(sw, ox, w, ac);
                    const float wx = (lx.i0 == sx ? lx.w0 : 0.f) + (lx.i1 == sx ? lx.w1 : 0.f);
                    if (wx == 0.f) continue;
                    float f[8];
                    unpack8(ldg128(b + (static_cast<int64_t>(oy) * ow + ox) * dy_ct), f);
                    const float wt = wy * wx;
#pragma unroll
                    for (int j = 0; j < 8; ++j) acc[j] = fmaf(wt, f[j], acc[j]);
                }
            }
        }
        st128(dx + ((static_cast<int64_t>(n) * h + sy) * w + sx) * dx_ct + dx_off + g * 8, pack8(acc));
    }
}

// The adjoint of an UPsampling, separably: a block owns one source row (n, sy) and a chunk of channels.  Phase 1 sums
// the destination rows that read sy, weighted, into a shared-memory line tmp[ox][channel] (fp32) — coalesced 128-bit
// loads, every dy element read by the two source rows it feeds instead of the four source pixels; phase 2 contracts the
// line along x into the w source pixels.  The gather form above re-reads every dy element four times through L2 and
// keeps only n*h*w*c/8 threads busy (88 us for the decoder's 134 MB gradient); this one reads it twice.
constexpr int kUpBwdRowBatch = 5;
constexpr int kUpBwdSmemBytes = 32 * 1024;

__global__ void __launch_bounds__(256) upsample_bwd_rows_kernel(const __nv_bfloat16* __restrict__ dy, int dy_ct, int dy_off,
                                                                 int oh, int ow, __nv_bfloat16* __restrict__ dx, int dx_ct,
                                                                 int dx_off, int h, int w, int c, int cc, float sh,
                                                                 float sw, int ac) {
    pdl_prologue();
    extern __shared__ __align__(16) float tmp[];                        // [ow][gcur][8]
    const int sy = blockIdx.y, n = blockIdx.z, ch0 = blockIdx.x * cc;
    const int gcur = min(cc, c - ch0) / 8;
    int ylo, yhi;
    bilinear_dst_range(sh, sy, h, oh, ac, ylo, yhi);
    while (ylo < yhi && bilinear_weight(sh, ylo, sy, h, ac) == 0.f) ++ylo;
    while (yhi > ylo && bilinear_weight(sh, yhi, sy, h, ac) == 0.f) --yhi;
    const __nv_bfloat16* b = dy + (static_cast<int64_t>(n) * oh * ow) * dy_ct + dy_off + ch0;
    const int64_t row_pitch = static_cast<int64_t>(ow) * dy_ct;
    const int items = ow * gcur;
    for (int it = threadIdx.x; it < items; it += 256) {
        const int ox = it / gcur, g = it - ox * gcur;
        const __nv_bfloat16* col = b + static_cast<int64_t>(ox) * dy_ct + g * 8;
        float acc[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = 0.f;
        for (int oy = ylo; oy <= yhi; oy += kUpBwdRowBatch) {
            uint4 v[kUpBwdRowBatch];
            float wy[kUpBwdRowBatch];
#pragma unroll
            for (int k = 0; k < kUpBwdRowBatch; ++k) {
                wy[k] = oy + k <= yhi ? bilinear_weight(sh, oy + k, sy, h, ac) : 0.f;
                v[k] = wy[k] != 0.f ? ldg128(col + (oy + k) * row_pitch) : make_uint4(0u, 0u, 0u, 0u);
            }
#pragma unroll
            for (int k = 0; k < kUpBwdRowBatch; ++k) {
                float f[8];
                unpack8(v[k], f);
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[j] = fmaf(wy[k], f[j], acc[j]);
            }
        }
        float4* d = reinterpret_cast<float4*>(tmp + static_cast<size_t>(it) * 8);
        d[0] = make_float4(acc[0], acc[1], acc[2], acc[3]);
        d[1] = make_float4(acc[4], acc[5], acc[6], acc[7]);
    }
    __syncthreads();
    const int outs = w * gcur;
    for (int it = threadIdx.x; it < outs; it += 256) {
        const int sx = it / gcur, g = it - sx * gcur;
        int xlo, xhi;
        bilinear_dst_range(sw, sx, w, ow, ac, xlo, xhi);
        float acc[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = 0.f;
        for (int ox = xlo; ox <= xhi; ++ox) {
            const float wx = bilinear_weight(sw, ox, sx, w, ac);
            if (wx == 0.f) continue;
            const float4* p = reinterpret_cast<const float4*>(tmp + (static_cast<size_t>(ox) * gcur + g) * 8);
            const float4 a = p[0], q = p[1];
            acc[0] = fmaf(wx, a.x, acc[0]); acc[1] = fmaf(wx, a.y, acc[1]);
            acc[2] = fmaf(wx, a.z, acc[2]); acc[3] = fmaf(wx, a.w, acc[3]);
            acc[4] = fmaf(wx, q.x, acc[4]); acc[5] = fmaf(wx, q.y, acc[5]);
            acc[6] = fmaf(wx, q.z, acc[6]); acc[7] = fmaf(wx, q.w, acc[7]);
        }
        st128(dx + ((static_cast<int64_t>(n) * h + sy) * w + sx) * dx_ct + dx_off + ch0 + g * 8, pack8(acc));
    }
}

// acc[n][c] += sum over a chunk of H*W (fp32 atomics), then a tiny kernel scales and rounds to bf16.
__global__ void __launch_bounds__(256) reduce_hw_kernel(const __nv_bfloat16* __restrict__ x, int x_ct, int x_off, int hw,
                                                         int c, int gpb, int rows, int chunk, float* __restrict__ acc) {
    pdl_prologue();
    __shared__ float red[256 * 8];
    const int groups = c >> 3;
    const int g_local = threadIdx.x % gpb, row = threadIdx.x / gpb;
    const int group = blockIdx.y * gpb + g_local;
    const int n = blockIdx.z;
    float a[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) a[j] = 0.f;
    if (row < rows && group < groups) {
        const int p0 = blockIdx.x * chunk, p1 = min(p0 + chunk, hw);
        const __nv_bfloat16* b = x + static_cast<int64_t>(n) * hw * x_ct + x_off + group * 8;
        for (int p = p0 + row; p < p1; p += rows) {
            float f[8];
            unpack8(ldg128(b + static_cast<int64_t>(p) * x_ct), f);
#pragma unroll
            for (int j = 0; j < 8; ++j) a[j] += f[j];
        }
    }
    if (row < rows) {
#pragma unroll
        for (int j = 0; j < 8; ++j) red[(row * gpb + g_local) * 8 + j] = a[j];
    }
    __syncthreads();
    for (int t = threadIdx.x; t < gpb * 8; t += blockDim.x) {
        const int gl = t >> 3, j = t & 7;
        const int gg = blockIdx.y * gpb + gl;
        if (gg < groups) {
            float s = 0.f;
            for (int r = 0; r < rows; ++r) s += red[(r * gpb + gl) * 8 + j];
            atomicAdd(acc + static_cast<int64_t>(n) * c + gg * 8 + j, s);
        }
    }
}

__global__ void scale_inplace_kernel(float* __restrict__ y, int64_t total, float scale) {
    pdl_prologue();
    const int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
    if (i < total) y[i] *= scale;
}

// y[n, p, slice] (+)= scale * v[n, :]   (v fp32 [n][c])
__global__ void __launch_bounds__(256) broadcast_hw_kernel(const float* __restrict__ v, int c, float scale,
                                                            __nv_bfloat16* __restrict__ y, int y_ct, int y_off, int hw,
                                                            int groups, int64_t total, int accumulate) {
    pdl_prologue();
    for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
         i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        const int g = static_cast<int>(i % groups);
        const int64_t p = i / groups;          // n * hw + pixel
        const int n = static_cast<int>(p / hw);
        float f[8];
        load8f(v + static_cast<int64_t>(n) * c + g * 8, f);
        __nv_bfloat16* dst = y + p * y_ct + y_off + g * 8;
        if (accumulate) {
            float o[8];
            unpack8(*reinterpret_cast<const uint4*>(dst), o);
#pragma unroll
            for (int j = 0; j < 8; ++j) f[j] = fmaf(scale, f[j], o[j]);
        } else if (scale != 1.f) {
#pragma unroll
            for (int j = 0; j < 8; ++j) f[j] *= scale;
        }
        st128(dst, pack8(f));
    }
}

int reduce_hw(const char* what, const void* x, int x_c_total, int x_c_offset, int n, int hw, int c, float scale,
              float* y, cudaStream_t st) {
    clear_error();
    if (check_slice(what, x, x_c_total, x_c_offset, c)) return 1;
    B200SEG_CHECK(n > 0 && hw > 0 && y, "%s: bad arguments", what);
    const int64_t total = static_cast<int64_t>(n) * c;
    B200SEG_CUDA(cudaMemsetAsync(y, 0, sizeof(float) * total, st));
    const int groups = c / 8;
    const int gpb = groups < 256 ? groups : 256;
    const int rows = 256 / gpb;
    const int gy = ceil_div(groups, gpb);
    int sms = sm_count();
    if (sms <= 0) sms = 148;
    int splits = ceil_div(sms * 8, n * gy);
    const int max_splits = ceil_div(hw, rows * 4);
    if (splits > max_splits) splits = max_splits;
    if (splits < 1) splits = 1;
    const int chunk = round_up(ceil_div(hw, splits), rows);
    B200SEG_CUDA(launch_pdl(reduce_hw_kernel, dim3(ceil_div(hw, chunk), gy, n), 256, 0, st, static_cast<const __nv_bfloat16*>(x), x_c_total,
                                                                      x_c_offset, hw, c, gpb, rows, chunk, y));
    if (scale != 1.f) {
        B200SEG_CUDA(launch_pdl(scale_inplace_kernel, static_cast<unsigned>(ceil_div64(total, 256)), 256, 0, st, y, total, scale));
    }
    return 0;
}

}  // namespace
}  // namespace b200seg

extern "C" {

int b200seg_upsample_fwd(const void* x, int x_c_total, int x_c_offset, int n, int h, int w, void* y, int y_c_total,
                         int y_c_offset, int oh, int ow, int c, int align_corners, b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    if (check_slice("upsample_fwd(x)", x, x_c_total, x_c_offset, c)) return 1;
    if (check_slice("upsample_fwd(y)", y, y_c_total, y_c_offset, c)) return 1;
    B200SEG_CHECK(n > 0 && h > 0 && w > 0 && oh > 0 && ow > 0, "upsample_fwd: bad shape");
    const int strips = ceil_div(oh, kUpRows);
    const int64_t total = static_cast<int64_t>(n) * strips * ow * (c / 8);
    B200SEG_CUDA(launch_pdl(upsample_fwd_kernel, elementwise_grid(total), 256, 0, static_cast<cudaStream_t>(stream),
        static_cast<const __nv_bfloat16*>(x), x_c_total, x_c_offset, h, w, static_cast<__nv_bfloat16*>(y), y_c_total,
        y_c_offset, oh, ow, c / 8, strips, total, bilinear_scale(h, oh, align_corners),
        bilinear_scale(w, ow, align_corners), align_corners));
    return 0;
}

int b200seg_upsample_bwd(const void* dy, int dy_c_total, int dy_c_offset, int n, int oh, int ow, void* dx, int dx_c_total,
                         int dx_c_offset, int h, int w, int c, int align_corners, b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    if (check_slice("upsample_bwd(dy)", dy, dy_c_total, dy_c_offset, c)) return 1;
    if (check_slice("upsample_bwd(dx)", dx, dx_c_total, dx_c_offset, c)) return 1;
    B200SEG_CHECK(n > 0 && h > 0 && w > 0 && oh > 0 && ow > 0, "upsample_bwd: bad shape");
    const cudaStream_t st = static_cast<cudaStream_t>(stream);
    const float sh = bilinear_scale(h, oh, align_corners), sw = bilinear_scale(w, ow, align_corners);
    // channels per block of the separable kernel: what fits its fp32 line of ow pixels
    const int cc = min(c, kUpBwdSmemBytes / 4 / ow / 8 * 8);
    if (oh >= h && ow >= w && cc >= 8 && n <= 65535 && h <= 65535) {
        const size_t smem = static_cast<size_t>(ow) * cc * sizeof(float);
        B200SEG_CUDA(launch_pdl(upsample_bwd_rows_kernel, dim3(ceil_div(c, cc), h, n), 256, smem, st,
            static_cast<const __nv_bfloat16*>(dy), dy_c_total, dy_c_offset, oh, ow, static_cast<__nv_bfloat16*>(dx),
            dx_c_total, dx_c_offset, h, w, c, cc, sh, sw, align_corners));
        return 0;
    }
    const int64_t total = static_cast<int64_t>(n) * h * w * (c / 8);
    B200SEG_CUDA(launch_pdl(upsample_bwd_gather_kernel, elementwise_grid(total), 256, 0, st,
        static_cast<const __nv_bfloat16*>(dy), dy_c_total, dy_c_offset, oh, ow, static_cast<__nv_bfloat16*>(dx), dx_c_total,
        dx_c_offset, h, w, c / 8, total, sh, sw, align_corners));
    return 0;
}

int b200seg_gap_fwd(const void* x, int x_c_total, int x_c_offset, int n, int hw, int c, float* y,
                    b200seg_stream_t stream) {
    return b200seg::reduce_hw("gap_fwd", x, x_c_total, x_c_offset, n, hw, c, 1.f / static_cast<float>(hw), y,
                              static_cast<cudaStream_t>(stream));
}

int b200seg_sum_hw(const void* x, int x_c_total, int x_c_offset, int n, int hw, int c, float* y,
                   b200seg_stream_t stream) {
    return b200seg::reduce_hw("sum_hw", x, x_c_total, x_c_offset, n, hw, c, 1.f, y, static_cast<cudaStream_t>(stream));
}

int b200seg_broadcast_hw(const float* v, int n, int c, float scale, void* y, int y_c_total, int y_c_offset, int hw,
                         int accumulate, b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    if (check_slice("broadcast_hw(y)", y, y_c_total, y_c_offset, c)) return 1;
    B200SEG_CHECK(v && n > 0 && hw > 0 && (reinterpret_cast<uintptr_t>(v) & 15) == 0, "broadcast_hw: bad arguments");
    const int64_t total = static_cast<int64_t>(n) * hw * (c / 8);
    B200SEG_CUDA(launch_pdl(broadcast_hw_kernel, elementwise_grid(total), 256, 0, static_cast<cudaStream_t>(stream), 
        v, c, scale, static_cast<__nv_bfloat16*>(y), y_c_total, y_c_offset, hw, c / 8, total, accumulate));
    return 0;
}

}  // extern "C"
