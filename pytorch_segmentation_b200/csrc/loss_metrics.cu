// Class-map resampling, pixel-wise softmax cross-entropy and the mIoU confusion histogram.
// All HBM-bound.  Logits at the drop-in boundary are NCHW fp32 (what the reference's forward returns), so the
// per-pixel kernels put consecutive threads on consecutive pixels of one class plane (coalesced 128-byte rows)
// and walk the class dimension in registers.
#include "common.cuh"
#include "pdl.cuh"
#include "vec.cuh"

namespace b200seg {
namespace {

// ------------------------------------------------------------------------------------------------------------
// final bilinear of the class map: NHWC fp32 [n][h][w][ct] -> NCHW fp32 [n][nc][oh][ow]
__global__ void __launch_bounds__(256) logits_up_fwd_kernel(const float* __restrict__ x, int ct, int h, int w, int nc,
                                                             float* __restrict__ y, int oh, int ow, int64_t total,
                                                             float sh, float sw, int ac) {
    pdl_prologue();
    for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
         i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        const int ox = static_cast<int>(i % ow);
        int64_t p = i / ow;
        const int oy = static_cast<int>(p % oh);
        const int n = static_cast<int>(p / oh);
        const Lerp ly = bilinear_src(sh, oy, h, ac), lx = bilinear_src(sw, ox, w, ac);
        const float* b = x + static_cast<int64_t>(n) * h * w * ct;
        const float4* p00 = reinterpret_cast<const float4*>(b + (static_cast<int64_t>(ly.i0) * w + lx.i0) * ct);
        const float4* p01 = reinterpret_cast<const float4*>(b + (static_cast<int64_t>(ly.i0) * w + lx.i1) * ct);
        const float4* p10 = reinterpret_cast<const float4*>(b + (static_cast<int64_t>(ly.i1) * w + lx.i0) * ct);
        const float4* p11 = reinterpret_cast<const float4*>(b + (static_cast<int64_t>(ly.i1) * w + lx.i1) * ct);
        float* out = y + (static_cast<int64_t>(n) * nc * oh + oy) * ow + ox;
        const int64_t plane = static_cast<int64_t>(oh) * ow;
        for (int c4 = 0; c4 * 4 < nc; ++c4) {
            const float4 a = __ldg(p00 + c4), bb = __ldg(p01 + c4), cc = __ldg(p10 + c4), d = __ldg(p11 + c4);
            const float v[4] = {ly.w0 * (lx.w0 * a.x + lx.w1 * bb.x) + ly.w1 * (lx.w0 * cc.x + lx.w1 * d.x),
                                ly.w0 * (lx.w0 * a.y + lx.w1 * bb.y) + ly.w1 * (lx.w0 * cc.y + lx.w1 * d.y),
                                ly.w0 * (lx.w0 * a.z + lx.w1 * bb.z) + ly.w1 * (lx.w0 * cc.z + lx.w1 * d.z),
                                ly.w0 * (lx.w0 * a.w + lx.w1 * bb.w) + ly.w1 * (lx.w0 * cc.w + lx.w1 * d.w)};
#pragma unroll
            for (int j = 0; j < 4; ++j)
                if (c4 * 4 + j < nc) __stcs(out + (c4 * 4 + j) * plane, v[j]);
        }
    }
}

// adjoint: NCHW fp32 gradient -> NHWC bf16 [n][h][w][ct]; one thread per source pixel gathers its window
template <int NCMAX>
__global__ void __launch_bounds__(128) logits_up_bwd_kernel(const float* __restrict__ dy, int nc, int oh, int ow,
                                                             __nv_bfloat16* __restrict__ dx, int ct, int h, int w,
                                                             int64_t total, float sh, float sw, int ac) {
    pdl_prologue();
    for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
         i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        const int sx = static_cast<int>(i % w);
        int64_t p = i / w;
        const int sy = static_cast<int>(p % h);
        const int n = static_cast<int>(p / h);
        int ylo, yhi, xlo, xhi;
        bilinear_dst_range(sh, sy, h, oh, ac, ylo, yhi);
        bilinear_dst_range(sw, sx, w, ow, ac, xlo, xhi);
        float acc[NCMAX];
#pragma unroll
        for (int c = 0; c < NCMAX; ++c) acc[c] = 0.f;
        const int64_t plane = static_cast<int64_t>(oh) * ow;
        const float* b = dy + static_cast<int64_t>(n) * nc * plane;
        for (int oy = ylo; oy <= yhi; ++oy) {
            const Lerp ly = bilinear_src(sh, oy, h, ac);
            const float wy = (ly.i0 == sy ? ly.w0 : 0.f) + (ly.i1 == sy ? ly.w1 : 0.f);
            if (wy == 0.f) continue;
            for (int ox = xlo; ox <= xhi; ++ox) {
                const Lerp lx = bilinear_src(sw, ox, w, ac);
                const float wx = (lx.i0 == sx ? lx.w0 : 0.f) + (lx.i1 == sx ? lx.w1 : 0.f);
                if (wx == 0.f) continue;
                const float wt = wy * wx;
                const float* q = b + static_cast<int64_t>(oy) * ow + ox;
#pragma unroll
                for (int c = 0; c < NCMAX; ++c)
                    if (c < nc) acc[c] = fmaf(wt, __ldcs(q + c * plane), acc[c]);
            }
        }
        __nv_bfloat16* o = dx + ((static_cast<int64_t>(n) * h + sy) * w + sx) * ct;
#pragma unroll
        for (int c8 = 0; c8 < NCMAX / 8; ++c8) {
            if (c8 * 8 < ct) {
                float f[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) f[j] = (c8 * 8 + j < nc) ? acc[c8 * 8 + j] : 0.f;
                st128(o + c8 * 8, pack8(f));
            }
        }
    }
}

// Same adjoint, tiled: a block owns 128 consecutive source pixels of one source row and four class planes at a time;
// thread (x, j) gathers the window of source pixel x in class plane j + 4k.  Lanes walk neighbouring, overlapping
// windows of the same dy rows (L1 serves the re-reads), the per-pixel x weights are computed once per thread, and
// the bf16 NHWC result leaves through shared memory as one contiguous run.  Needs an x window of at most kWinMax
// taps (up-scaling factors up to ~6); the host falls back to the kernel above otherwise.
constexpr int kWinMax = 16;
__global__ void __launch_bounds__(512) logits_up_bwd_tiled_kernel(const float* __restrict__ dy, int nc, int oh, int ow,
                                                                   __nv_bfloat16* __restrict__ dx, int ct, int h, int w,
                                                                   float sh, float sw, int ac) {
    pdl_prologue();
    extern __shared__ __align__(16) unsigned char tile_raw[];
    __nv_bfloat16* tile = reinterpret_cast<__nv_bfloat16*>(tile_raw);     // [128][ct]
    const int tx = threadIdx.x, ty = threadIdx.y;
    const int sx = blockIdx.x * 128 + tx, sy = blockIdx.y, n = blockIdx.z;
    int ylo, yhi, xlo = 0, xhi = -1;
    bilinear_dst_range(sh, sy, h, oh, ac, ylo, yhi);
    float wx[kWinMax];
#pragma unroll
    for (int k = 0; k < kWinMax; ++k) wx[k] = 0.f;
    if (sx < w) {
        bilinear_dst_range(sw, sx, w, ow, ac, xlo, xhi);
#pragma unroll
        for (int k = 0; k < kWinMax; ++k) {
            const int ox = xlo + k;
            if (ox <= xhi) {
                const Lerp lx = bilinear_src(sw, ox, w, ac);
                wx[k] = (lx.i0 == sx ? lx.w0 : 0.f) + (lx.i1 == sx ? lx.w1 : 0.f);
            }
        }
    }
    const int64_t plane = static_cast<int64_t>(oh) * ow;
    for (int c = ty; c < ct; c += blockDim.y) {
        float acc = 0.f;
        if (c < nc && sx < w) {
            const float* b = dy + (static_cast<int64_t>(n) * nc + c) * plane;
            for (int oy = ylo; oy <= yhi; ++oy) {
                const Lerp ly = bilinear_src(sh, oy, h, ac);
                const float wy = (ly.i0 == sy ? ly.w0 : 0.f) + (ly.i1 == sy ? ly.w1 : 0.f);
                if (wy == 0.f) continue;
                const float* q = b + static_cast<int64_t>(oy) * ow + xlo;
                float r = 0.f;
#pragma unroll
                for (int k = 0; k < kWinMax; ++k)
                    if (xlo + k <= xhi) r = fmaf(wx[k], __ldg(q + k), r);
                acc = fmaf(wy, r, acc);
            }
        }
        tile[tx * ct + c] = __float2bfloat16(acc);
    }
    __syncthreads();
    // the block's pixels are consecutive in NHWC: one contiguous run of npix * ct bf16
    const int npix = min(128, w - blockIdx.x * 128);
    const int vecs = npix * ct / 8;
    uint4* dst = reinterpret_cast<uint4*>(dx + ((static_cast<int64_t>(n) * h + sy) * w + blockIdx.x * 128) * ct);
    const uint4* src = reinterpret_cast<const uint4*>(tile);
    for (int i = ty * 128 + tx; i < vecs; i += 128 * blockDim.y) dst[i] = src[i];
}

// generic NCHW fp32 bilinear resize (compute_loss when the logits and the targets differ in size)
__global__ void __launch_bounds__(256) resize_nchw_fwd_kernel(const float* __restrict__ x, int h, int w,
                                                               float* __restrict__ y, int oh, int ow, int64_t total,
                                                               float sh, float sw, int ac) {
    pdl_prologue();
    for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
         i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        const int ox = static_cast<int>(i % ow);
        int64_t p = i / ow;
        const int oy = static_cast<int>(p % oh);
        const int64_t nc = p / oh;   // fused (n, c) plane index
        const Lerp ly = bilinear_src(sh, oy, h, ac), lx = bilinear_src(sw, ox, w, ac);
        const float* b = x + nc * h * w;
        y[i] = ly.w0 * (lx.w0 * __ldg(b + ly.i0 * w + lx.i0) + lx.w1 * __ldg(b + ly.i0 * w + lx.i1)) +
               ly.w1 * (lx.w0 * __ldg(b + ly.i1 * w + lx.i0) + lx.w1 * __ldg(b + ly.i1 * w + lx.i1));
    }
}

__global__ void __launch_bounds__(256) resize_nchw_bwd_kernel(const float* __restrict__ dy, int oh, int ow,
                                                               float* __restrict__ dx, int h, int w, int64_t total,
                                                               float sh, float sw, int ac) {
    pdl_prologue();
    for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
         i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        const int sx = static_cast<int>(i % w);
        int64_t p = i / w;
        const int sy = static_cast<int>(p % h);
        const int64_t nc = p / h;
        int ylo, yhi, xlo, xhi;
        bilinear_dst_range(sh, sy, h, oh, ac, ylo, yhi);
        bilinear_dst_range(sw, sx, w, ow, ac, xlo, xhi);
        const float* b = dy + nc * oh * ow;
        float acc = 0.f;
        for (int oy = ylo; oy <= yhi; ++oy) {
            const Lerp ly = bilinear_src(sh, oy, h, ac);
            const float wy = (ly.i0 == sy ? ly.w0 : 0.f) + (ly.i1 == sy ? ly.w1 : 0.f);
            if (wy == 0.f) continue;
            for (int ox = xlo; ox <= xhi; ++ox) {
                const Lerp lx = bilinear_src(sw, ox, w, ac);
                const float wx = (lx.i0 == sx ? lx.w0 : 0.f) + (lx.i1 == sx ? lx.w1 : 0.f);
                if (wx != 0.f) acc = fmaf(wy * wx, __ldg(b + static_cast<int64_t>(oy) * ow + ox), acc);
            }
        }
        dx[i] = acc;
    }
}

// ------------------------------------------------------------------------------------------------------------
// softmax cross-entropy, one thread per pixel, class vector in registers
template <int NCMAX>
__global__ void __launch_bounds__(256) softmax_ce_kernel(const float* __restrict__ logits, const int64_t* __restrict__ tgt,
                                                          int nc, int64_t hw, int64_t total, int64_t ignore_index,
                                                          float grad_scale, double* loss_sum,
                                                          unsigned long long* valid_count, float* __restrict__ dlogits,
                                                          int* status, const float* __restrict__ g_dev,
                                                          const unsigned long long* __restrict__ count_dev) {
    pdl_prologue();
    // backward form: the upstream gradient and the valid-pixel count live on the device (no host read)
    if (g_dev) grad_scale *= __ldg(g_dev) / static_cast<float>(__ldg(count_dev));
    double lsum = 0.0;
    unsigned int lcount = 0;
    for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
         i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        const int64_t n = i / hw, p = i - n * hw;
        const float* q = logits + n * nc * hw + p;
        const int64_t t = __ldcs(tgt + i);
        float v[NCMAX];
        float m = -INFINITY;
#pragma unroll
        for (int c = 0; c < NCMAX; ++c) {
            v[c] = c < nc ? __ldcs(q + c * hw) : -INFINITY;
            m = fmaxf(m, v[c]);
        }
        float s = 0.f;
#pragma unroll
        for (int c = 0; c < NCMAX; ++c) {
            v[c] = c < nc ? __expf(v[c] - m) : 0.f;
            s += v[c];
        }
        const bool ignored = (t == ignore_index);
        const bool bad = !ignored && (t < 0 || t >= nc);
        if (bad && status) *status = 1;
        const bool valid = !ignored && !bad;
        float xt = 0.f;
        if (valid) {
            xt = __ldg(q + t * hw);
            lsum += static_cast<double>(__logf(s) + m - xt);
            ++lcount;
        }
        if (dlogits) {
            const float inv = valid ? grad_scale / s : 0.f;
            float* d = dlogits + n * nc * hw + p;
#pragma unroll
            for (int c = 0; c < NCMAX; ++c)
                if (c < nc) __stcs(d + c * hw, v[c] * inv - ((valid && c == t) ? grad_scale : 0.f));
        }
    }
    lsum = warp_sum(lsum);
    lcount = __reduce_add_sync(0xffffffffu, lcount);
    __shared__ double ssum[8];
    __shared__ unsigned int scnt[8];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) {
        ssum[warp] = lsum;
        scnt[warp] = lcount;
    }
    __syncthreads();
    if (threadIdx.x == 0 && loss_sum) {
        double a = 0.0;
        unsigned long long b = 0;
        for (int k = 0; k < (blockDim.x >> 5); ++k) {
            a += ssum[k];
            b += scnt[k];
        }
        atomicAdd(loss_sum, a);
        if (valid_count) atomicAdd(valid_count, b);
    }
}

// any class count: three passes over the class dimension (max, sum, gradient)
__global__ void __launch_bounds__(256) softmax_ce_generic_kernel(const float* __restrict__ logits,
                                                                  const int64_t* __restrict__ tgt, int nc, int64_t hw,
                                                                  int64_t total, int64_t ignore_index, float grad_scale,
                                                                  double* loss_sum, unsigned long long* valid_count,
                                                                  float* __restrict__ dlogits, int* status,
                                                                  const float* __restrict__ g_dev,
                                                                  const unsigned long long* __restrict__ count_dev) {
    pdl_prologue();
    if (g_dev) grad_scale *= __ldg(g_dev) / static_cast<float>(__ldg(count_dev));
    double lsum = 0.0;
    unsigned int lcount = 0;
    for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
         i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        const int64_t n = i / hw, p = i - n * hw;
        const float* q = logits + n * nc * hw + p;
        const int64_t t = tgt[i];
        float m = -INFINITY;
        for (int c = 0; c < nc; ++c) m = fmaxf(m, q[c * hw]);
        float s = 0.f;
        for (int c = 0; c < nc; ++c) s += __expf(q[c * hw] - m);
        const bool ignored = (t == ignore_index);
        const bool bad = !ignored && (t < 0 || t >= nc);
        if (bad && status) *status = 1;
        const bool valid = !ignored && !bad;
        if (valid) {
            lsum += static_cast<double>(__logf(s) + m - q[t * hw]);
            ++lcount;
        }
        if (dlogits) {
            const float inv = valid ? grad_scale / s : 0.f;
            float* d = dlogits + n * nc * hw + p;
            for (int c = 0; c < nc; ++c)
                d[c * hw] = __expf(q[c * hw] - m) * inv - ((valid && c == t) ? grad_scale : 0.f);
        }
    }
    lsum = warp_sum(lsum);
    lcount = __reduce_add_sync(0xffffffffu, lcount);
    if ((threadIdx.x & 31) == 0 && loss_sum) {
        atomicAdd(loss_sum, lsum);
        if (valid_count) atomicAdd(valid_count, static_cast<unsigned long long>(lcount));
    }
}

// ------------------------------------------------------------------------------------------------------------
// argmax over the class planes; lowest index wins ties (strict >), NaN counts as the maximum like torch.max
__device__ __forceinline__ int argmax_planes(const float* q, int nc, int64_t hw) {
    float best = __ldcs(q);
    int bi = 0;
    for (int c = 1; c < nc; ++c) {
        const float v = __ldcs(q + c * hw);
        if (v > best || (v != v && best == best)) {
            best = v;
            bi = c;
        }
    }
    return bi;
}

__global__ void __launch_bounds__(256) argmax_kernel(const float* __restrict__ logits, int nc, int64_t hw, int64_t total,
                                                      int64_t* __restrict__ pred) {
    pdl_prologue();
    for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
         i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        const int64_t n = i / hw, p = i - n * hw;
        pred[i] = argmax_planes(logits + n * nc * hw + p, nc, hw);
    }
}

// Confusion histogram: every warp owns a private nc x nc uint32 histogram in shared memory (no inter-warp
// contention), flushed with one 64-bit global atomic per non-zero bin at the end of the block.
constexpr int kHistThreads = 256;
constexpr int kHistWarps = kHistThreads / 32;

template <bool FUSED_ARGMAX>
__global__ void __launch_bounds__(kHistThreads) confusion_kernel(const float* __restrict__ logits,
                                                                  const int64_t* __restrict__ pred,
                                                                  const int64_t* __restrict__ tgt, int nc, int64_t hw,
                                                                  int64_t total, int copies,
                                                                  unsigned long long* __restrict__ cm) {
    pdl_prologue();
    extern __shared__ unsigned int hist[];
    const int bins = nc * nc;
    for (int b = threadIdx.x; b < bins * copies; b += blockDim.x) hist[b] = 0;
    __syncthreads();
    unsigned int* mine = hist + ((threadIdx.x >> 5) % copies) * bins;
    for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
         i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        const int64_t t = __ldcs(tgt + i);
        int64_t pr;
        if (FUSED_ARGMAX) {
            const int64_t n = i / hw, p = i - n * hw;
            pr = argmax_planes(logits + n * nc * hw + p, nc, hw);
        } else {
            pr = __ldcs(pred + i);
        }
        if (t >= 0 && t < nc && pr >= 0 && pr < nc) atomicAdd(mine + static_cast<int>(t) * nc + static_cast<int>(pr), 1u);
    }
    __syncthreads();
    for (int b = threadIdx.x; b < bins; b += blockDim.x) {
        unsigned long long s = 0;
        for (int k = 0; k < copies; ++k) s += hist[k * bins + b];
        if (s) atomicAdd(cm + b, s);
    }
}

int hist_launch_geometry(int nc, int64_t total, int* copies, size_t* smem, int* grid) {
    B200SEG_CHECK(nc > 0, "confusion: nc = %d", nc);
    const size_t bins = static_cast<size_t>(nc) * nc;
    B200SEG_CHECK(bins * 4 <= 48 * 1024, "confusion: %d classes need %zu bytes of shared memory (max 110 classes)", nc,
                  bins * 4);
    int cp = static_cast<int>((48 * 1024) / (bins * 4));
    if (cp > kHistWarps) cp = kHistWarps;
    *copies = cp;
    *smem = bins * 4 * cp;
    int sms = sm_count();
    if (sms <= 0) sms = 148;
    // each thread keeps its 32-bit bins far from overflow: at most 2^31 items per block
    int64_t blocks = ceil_div64(total, static_cast<int64_t>(kHistThreads) * 16);
    if (blocks > static_cast<int64_t>(sms) * 8) blocks = static_cast<int64_t>(sms) * 8;
    if (blocks < 1) blocks = 1;
    *grid = static_cast<int>(blocks);
    return 0;
}

}  // namespace
}  // namespace b200seg

extern "C" {

int b200seg_logits_upsample_fwd(const float* x, int c_total, int n, int h, int w, int nc, float* y, int oh, int ow,
                                int align_corners, b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    B200SEG_CHECK(x && y && n > 0 && h > 0 && w > 0 && oh > 0 && ow > 0, "logits_upsample_fwd: bad arguments");
    B200SEG_CHECK(nc > 0 && nc <= c_total && c_total % 4 == 0, "logits_upsample_fwd: nc %d vs padded %d", nc, c_total);
    B200SEG_CHECK((reinterpret_cast<uintptr_t>(x) & 15) == 0, "logits_upsample_fwd: input not 16-byte aligned");
    const int64_t total = static_cast<int64_t>(n) * oh * ow;
    B200SEG_CUDA(launch_pdl(logits_up_fwd_kernel, elementwise_grid(total), 256, 0, static_cast<cudaStream_t>(stream), 
        x, c_total, h, w, nc, y, oh, ow, total, bilinear_scale(h, oh, align_corners), bilinear_scale(w, ow, align_corners),
        align_corners));
    return 0;
}

int b200seg_logits_upsample_bwd(const float* dy, int n, int nc, int oh, int ow, void* dx, int c_total, int h, int w,
                                int align_corners, b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    B200SEG_CHECK(dy && dx && n > 0 && h > 0 && w > 0 && oh > 0 && ow > 0, "logits_upsample_bwd: bad arguments");
    B200SEG_CHECK(nc > 0 && nc <= c_total && c_total % 8 == 0 && c_total <= 64,
                  "logits_upsample_bwd: nc %d, padded %d (multiple of 8, at most 64)", nc, c_total);
    B200SEG_CHECK((reinterpret_cast<uintptr_t>(dx) & 15) == 0, "logits_upsample_bwd: output not 16-byte aligned");
    const int64_t total = static_cast<int64_t>(n) * h * w;
    const float sh = bilinear_scale(h, oh, align_corners), sw = bilinear_scale(w, ow, align_corners);
    const int grid = static_cast<int>(ceil_div64(total, 128));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    __nv_bfloat16* o = static_cast<__nv_bfloat16*>(dx);
    // x window of a source pixel: at most 2/scale + 3 destination columns
    const bool tiled = sw > 0.f && (2.f / sw + 6.f) <= static_cast<float>(kWinMax) && h <= 65535 && n <= 65535;
    if (tiled) {
        B200SEG_CUDA(launch_pdl(logits_up_bwd_tiled_kernel, dim3(ceil_div(w, 128), h, n), dim3(128, 4), 128 * c_total * 2, st, 
            dy, nc, oh, ow, o, c_total, h, w, sh, sw, align_corners));
        return 0;
    }
    if (c_total <= 8)
        B200SEG_CUDA(launch_pdl(logits_up_bwd_kernel<8>, grid, 128, 0, st, dy, nc, oh, ow, o, c_total, h, w, total, sh, sw, align_corners));
    else if (c_total <= 24)
        B200SEG_CUDA(launch_pdl(logits_up_bwd_kernel<24>, grid, 128, 0, st, dy, nc, oh, ow, o, c_total, h, w, total, sh, sw, align_corners));
    else if (c_total <= 32)
        B200SEG_CUDA(launch_pdl(logits_up_bwd_kernel<32>, grid, 128, 0, st, dy, nc, oh, ow, o, c_total, h, w, total, sh, sw, align_corners));
    else
        B200SEG_CUDA(launch_pdl(logits_up_bwd_kernel<64>, grid, 128, 0, st, dy, nc, oh, ow, o, c_total, h, w, total, sh, sw, align_corners));
    return 0;
}

int b200seg_resize_nchw_fwd(const float* x, int64_t planes, int h, int w, float* y, int oh, int ow, int align_corners,
                            b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    B200SEG_CHECK(x && y && planes > 0 && h > 0 && w > 0 && oh > 0 && ow > 0, "resize_nchw_fwd: bad arguments");
    const int64_t total = planes * oh * ow;
    B200SEG_CUDA(launch_pdl(resize_nchw_fwd_kernel, elementwise_grid(total), 256, 0, static_cast<cudaStream_t>(stream), 
        x, h, w, y, oh, ow, total, bilinear_scale(h, oh, align_corners), bilinear_scale(w, ow, align_corners),
        align_corners));
    return 0;
}

int b200seg_resize_nchw_bwd(const float* dy, int64_t planes, int oh, int ow, float* dx, int h, int w, int align_corners,
                            b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    B200SEG_CHECK(dy && dx && planes > 0 && h > 0 && w > 0 && oh > 0 && ow > 0, "resize_nchw_bwd: bad arguments");
    const int64_t total = planes * h * w;
    B200SEG_CUDA(launch_pdl(resize_nchw_bwd_kernel, elementwise_grid(total), 256, 0, static_cast<cudaStream_t>(stream), 
        dy, oh, ow, dx, h, w, total, bilinear_scale(h, oh, align_corners), bilinear_scale(w, ow, align_corners),
        align_corners));
    return 0;
}

}  // extern "C"

namespace b200seg {
namespace {
// ------------------------------------------------------------------------------------------------------------
// Cross-entropy straight from the LOW-RESOLUTION class map (the head's fp32 NHWC [n][h][w][ct] output before its final
// bilinear upsample, models/deeplabv3plus.py:40-43 / unet.py:52-55 / hrnet.py:400-404): the full-resolution logits are
// interpolated on the fly, never read back from HBM (forward: 12.6 + 34 MB instead of 352 + 34 MB), and the backward
// emits the gradient of the low-resolution map directly — softmax recomputed per pixel, scattered through the bilinear
// weights into a shared-memory tile, flushed with fp32 atomics — instead of writing 352 MB of dlogits and reading them
// again in the upsample adjoint.  One CTA = a 16 x 64 tile of target pixels; its source footprint (FH x FW low-res
// pixels) is staged in shared memory.  Each thread owns 4 consecutive pixels of a row and pre-reduces their
// contributions per source column in registers (they share 1-2 source columns), so the shared-memory atomics are a
// quarter of the naive count.
constexpr int kLrTH = 16, kLrTW = 64;

template <int NCMAX, bool BWD>
__global__ void __launch_bounds__(256, BWD ? 2 : 4) lowres_ce_kernel(const float* __restrict__ lr, int ct, int h, int w, int nc,
                                                         const int64_t* __restrict__ tgt, int H, int W, float sh, float sw,
                                                         int ac, int64_t ignore_index, double* loss_sum,
                                                         unsigned long long* valid_count, int* status,
                                                         float* __restrict__ dlr, const float* __restrict__ g_dev,
                                                         const unsigned long long* __restrict__ count_dev, int FH, int FW) {
    pdl_prologue();
    extern __shared__ __align__(16) float lsm[];
    float* L = lsm;                               // [FH][FW][ct] staged class map
    float* G = lsm + FH * FW * ct;                // [FH][FW][ct] gradient accumulator (backward only)
    const int n = blockIdx.z, ty0 = blockIdx.y * kLrTH, tx0 = blockIdx.x * kLrTW;
    const int fy0 = bilinear_src(sh, ty0, h, ac).i0, fx0 = bilinear_src(sw, tx0, w, ac).i0;
    const float* base = lr + static_cast<int64_t>(n) * h * w * ct;
    const int cells = FH * FW;
    for (int i = threadIdx.x; i < cells * (ct >> 2); i += blockDim.x) {
        const int q = i % (ct >> 2), cell = i / (ct >> 2);
        const int yy = fy0 + cell / FW, xx = fx0 + cell % FW;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (yy < h && xx < w) v = __ldg(reinterpret_cast<const float4*>(base + (static_cast<int64_t>(yy) * w + xx) * ct) + q);
        reinterpret_cast<float4*>(L)[i] = v;
        if (BWD) reinterpret_cast<float4*>(G)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    __syncthreads();
    float gscale = 1.f;
    if (BWD) gscale = __ldg(g_dev) / static_cast<float>(__ldg(count_dev));
    const int ry = threadIdx.x >> 4, cg = threadIdx.x & 15;
    const int oy = ty0 + ry;
    double lsum = 0.0;
    unsigned int lcount = 0;
    if (oy < H) {
        const Lerp ly = bilinear_src(sh, oy, h, ac);
        const int y0 = ly.i0 - fy0, y1 = ly.i1 - fy0;
        float acc0[NCMAX], acc1[NCMAX];
        int cur = -1;
        auto flush = [&]() {
            if (cur < 0) return;
            float* g00 = G + (y0 * FW + cur) * ct;
            float* g10 = G + (y1 * FW + cur) * ct;
#pragma unroll
            for (int c = 0; c < NCMAX; ++c) {
                if (c < nc) {
                    atomicAdd(g00 + c, ly.w0 * acc0[c]);
                    if (ly.w1 != 0.f) atomicAdd(g10 + c, ly.w1 * acc0[c]);
                    if (cur + 1 < FW) {
                        atomicAdd(g00 + ct + c, ly.w0 * acc1[c]);
                        if (ly.w1 != 0.f) atomicAdd(g10 + ct + c, ly.w1 * acc1[c]);
                    }
                }
            }
        };
#pragma unroll 1
        for (int j = 0; j < 4; ++j) {
            const int ox = tx0 + cg * 4 + j;
            if (ox >= W) break;
            const Lerp lx = bilinear_src(sw, ox, w, ac);
            const int x0 = lx.i0 - fx0, x1 = lx.i1 - fx0;
            const float* p00 = L + (y0 * FW + x0) * ct;
            const float* p01 = L + (y0 * FW + x1) * ct;
            const float* p10 = L + (y1 * FW + x0) * ct;
            const float* p11 = L + (y1 * FW + x1) * ct;
            float v[NCMAX];
            float m = -INFINITY;
#pragma unroll
            for (int c = 0; c < NCMAX; ++c) {
                // same expression as logits_up_fwd_kernel, so the loss is the loss of the logits the model returned
                v[c] = c < nc ? ly.w0 * (lx.w0 * p00[c] + lx.w1 * p01[c]) + ly.w1 * (lx.w0 * p10[c] + lx.w1 * p11[c]) : -INFINITY;
                m = fmaxf(m, v[c]);
            }
            float s = 0.f;
#pragma unroll
            for (int c = 0; c < NCMAX; ++c) {
                v[c] = c < nc ? __expf(v[c] - m) : 0.f;
                s += v[c];
            }
            const int64_t t = __ldcs(tgt + (static_cast<int64_t>(n) * H + oy) * W + ox);
            const bool ignored = (t == ignore_index);
            const bool bad = !ignored && (t < 0 || t >= nc);
            if (bad && status) *status = 1;
            const bool valid = !ignored && !bad;
            if (!BWD) {
                if (valid) {
                    // log-softmax at the target: the interpolated logit is recomputed rather than indexed dynamically
                    const int tc = static_cast<int>(t);
                    const float xt = ly.w0 * (lx.w0 * p00[tc] + lx.w1 * p01[tc]) + ly.w1 * (lx.w0 * p10[tc] + lx.w1 * p11[tc]);
                    lsum += static_cast<double>(__logf(s) + m - xt);
                    ++lcount;
                }
                continue;
            }
            if (x0 != cur) {
                flush();
                cur = x0;
#pragma unroll
                for (int c = 0; c < NCMAX; ++c) acc0[c] = acc1[c] = 0.f;
            }
            const float inv = valid ? gscale / s : 0.f;
            const float a0 = x1 == x0 ? lx.w0 + lx.w1 : lx.w0, a1 = x1 == x0 ? 0.f : lx.w1;
#pragma unroll
            for (int c = 0; c < NCMAX; ++c) {
                const float d = v[c] * inv - ((valid && c == t) ? gscale : 0.f);
                acc0[c] = fmaf(a0, d, acc0[c]);
                acc1[c] = fmaf(a1, d, acc1[c]);
            }
        }
        if (BWD) flush();
    }
    if (!BWD) {
        lsum = warp_sum(lsum);
        lcount = __reduce_add_sync(0xffffffffu, lcount);
        __shared__ double ssum[8];
        __shared__ unsigned int scnt[8];
        const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
        if (lane == 0) {
            ssum[warp] = lsum;
            scnt[warp] = lcount;
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            double a = 0.0;
            unsigned long long b = 0;
            for (int k = 0; k < 8; ++k) {
                a += ssum[k];
                b += scnt[k];
            }
            atomicAdd(loss_sum, a);
            atomicAdd(valid_count, b);
        }
        return;
    }
    __syncthreads();
    float* out = dlr + static_cast<int64_t>(n) * h * w * ct;
    for (int i = threadIdx.x; i < cells * ct; i += blockDim.x) {
        const int c = i % ct, cell = i / ct;
        const int yy = fy0 + cell / FW, xx = fx0 + cell % FW;
        const float v = G[i];
        if (c < nc && yy < h && xx < w && v != 0.f) atomicAdd(out + (static_cast<int64_t>(yy) * w + xx) * ct + c, v);
    }
}

// out = bf16(a [+ b]) over n elements (n % 4 == 0): the fp32 class-map gradient becomes the bf16 slice the conv
// backward consumes; b carries the gradient that reached the full-resolution logits some other way (may be null)
__global__ void cast_add_bf16_kernel(const float4* __restrict__ a, const uint2* __restrict__ b, uint2* __restrict__ out,
                                     int64_t n4) {
    pdl_prologue();
    for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < n4;
         i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        float4 v = __ldg(a + i);
        if (b) {
            const uint2 r = __ldg(b + i);
            const __nv_bfloat162 lo = *reinterpret_cast<const __nv_bfloat162*>(&r.x);
            const __nv_bfloat162 hi = *reinterpret_cast<const __nv_bfloat162*>(&r.y);
            v.x += __low2float(lo); v.y += __high2float(lo); v.z += __low2float(hi); v.w += __high2float(hi);
        }
        const __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
        uint2 o;
        o.x = *reinterpret_cast<const unsigned int*>(&lo);
        o.y = *reinterpret_cast<const unsigned int*>(&hi);
        out[i] = o;
    }
}

int launch_lowres_ce(const float* lr, int n, int h, int w, int ct, int nc, const int64_t* targets, int H, int W,
                     int align_corners, int64_t ignore_index, double* loss_sum, int64_t* valid_count, int* status, float* dlr,
                     const float* g_dev, const int64_t* count_dev, cudaStream_t st) {
    const float sh = bilinear_scale(h, H, align_corners), sw = bilinear_scale(w, W, align_corners);
    // footprint of a 16 x 64 target tile in the class map (+1 for the second lerp tap, +1 for rounding)
    const int FH = static_cast<int>(kLrTH * sh) + 3 < h ? static_cast<int>(kLrTH * sh) + 3 : h;
    const int FW = static_cast<int>(kLrTW * sw) + 3 < w ? static_cast<int>(kLrTW * sw) + 3 : w;
    const size_t smem = static_cast<size_t>(FH) * FW * ct * sizeof(float) * (dlr ? 2 : 1);
    B200SEG_CHECK(smem <= 48 * 1024, "lowres_ce: class-map footprint of %zu bytes per tile exceeds shared memory", smem);
    B200SEG_CHECK(n <= 65535 && ceil_div(H, kLrTH) <= 65535, "lowres_ce: grid too large");
    dim3 grid(ceil_div(W, kLrTW), ceil_div(H, kLrTH), n);
    unsigned long long* vc = reinterpret_cast<unsigned long long*>(valid_count);
    const unsigned long long* cd = reinterpret_cast<const unsigned long long*>(count_dev);
#define B200SEG_LRCE(NC, B)                                                                                              \
    B200SEG_CUDA(launch_pdl(lowres_ce_kernel<NC, B>, grid, 256, smem, st, lr, ct, h, w, nc, targets, H, W, sh, sw,           \
                            align_corners, ignore_index, loss_sum, vc, status, dlr, g_dev, cd, FH, FW))
    if (dlr) {
        if (nc <= 8) B200SEG_LRCE(8, true);
        else if (nc <= 24) B200SEG_LRCE(24, true);
        else B200SEG_LRCE(32, true);
    } else {
        if (nc <= 8) B200SEG_LRCE(8, false);
        else if (nc <= 24) B200SEG_LRCE(24, false);
        else B200SEG_LRCE(32, false);
    }
#undef B200SEG_LRCE
    return 0;
}

// loss = sum / count as fp32 (NaN when a target was out of range): the scalar tail of nn.CrossEntropyLoss's mean
__global__ void ce_finalize_kernel(const double* __restrict__ loss_sum, const unsigned long long* __restrict__ count,
                                   const int* __restrict__ status, float* __restrict__ loss) {
    pdl_prologue();
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        const double c = static_cast<double>(*count);
        *loss = (*status != 0) ? __int_as_float(0x7fc00000) : static_cast<float>(*loss_sum / c);
    }
}

int launch_softmax_ce(const float* logits, const int64_t* targets, int n, int nc, int64_t hw, int64_t ignore_index,
                      float grad_scale, double* loss_sum, int64_t* valid_count, float* dlogits, int* status,
                      const float* g_dev, const int64_t* count_dev, cudaStream_t st) {
    const int64_t total = static_cast<int64_t>(n) * hw;
    const int grid = elementwise_grid(total);
    unsigned long long* vc = reinterpret_cast<unsigned long long*>(valid_count);
    const unsigned long long* cd = reinterpret_cast<const unsigned long long*>(count_dev);
    if (nc <= 8)
        B200SEG_CUDA(launch_pdl(softmax_ce_kernel<8>, grid, 256, 0, st, logits, targets, nc, hw, total, ignore_index, grad_scale,
                                loss_sum, vc, dlogits, status, g_dev, cd));
    else if (nc <= 24)
        B200SEG_CUDA(launch_pdl(softmax_ce_kernel<24>, grid, 256, 0, st, logits, targets, nc, hw, total, ignore_index, grad_scale,
                                loss_sum, vc, dlogits, status, g_dev, cd));
    else if (nc <= 32)
        B200SEG_CUDA(launch_pdl(softmax_ce_kernel<32>, grid, 256, 0, st, logits, targets, nc, hw, total, ignore_index, grad_scale,
                                loss_sum, vc, dlogits, status, g_dev, cd));
    else
        B200SEG_CUDA(launch_pdl(softmax_ce_generic_kernel, grid, 256, 0, st, logits, targets, nc, hw, total, ignore_index,
                                grad_scale, loss_sum, vc, dlogits, status, g_dev, cd));
    return 0;
}
}  // namespace
}  // namespace b200seg

extern "C" {

int b200seg_softmax_ce(const float* logits, const int64_t* targets, int n, int nc, int64_t hw, int64_t ignore_index,
                       float grad_scale, double* loss_sum, int64_t* valid_count, float* dlogits, int* status,
                       b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    B200SEG_CHECK(logits && targets && loss_sum && status && n > 0 && nc > 0 && hw > 0, "softmax_ce: bad arguments");
    return launch_softmax_ce(logits, targets, n, nc, hw, ignore_index, grad_scale, loss_sum, valid_count, dlogits, status,
                             nullptr, nullptr, static_cast<cudaStream_t>(stream));
}

int b200seg_lowres_ce_fwd(const float* lr, int n, int h, int w, int c_total, int nc, const int64_t* targets, int th, int tw,
                          int align_corners, int64_t ignore_index, double* loss_sum, int64_t* valid_count, int* status,
                          b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    B200SEG_CHECK(lr && targets && loss_sum && valid_count && status, "lowres_ce_fwd: null pointer");
    B200SEG_CHECK(n > 0 && h > 0 && w > 0 && th >= h && tw >= w && nc > 0 && nc <= c_total && nc <= 32 && c_total % 4 == 0,
                  "lowres_ce_fwd: bad geometry (class map %dx%d, targets %dx%d, nc %d of %d)", h, w, th, tw, nc, c_total);
    B200SEG_CHECK((reinterpret_cast<uintptr_t>(lr) & 15) == 0, "lowres_ce_fwd: class map not 16-byte aligned");
    return launch_lowres_ce(lr, n, h, w, c_total, nc, targets, th, tw, align_corners, ignore_index, loss_sum, valid_count,
                            status, nullptr, nullptr, nullptr, static_cast<cudaStream_t>(stream));
}

int b200seg_lowres_ce_bwd(const float* lr, int n, int h, int w, int c_total, int nc, const int64_t* targets, int th, int tw,
                          int align_corners, int64_t ignore_index, const float* grad_out, const int64_t* valid_count,
                          float* dlr, b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    B200SEG_CHECK(lr && targets && grad_out && valid_count && dlr, "lowres_ce_bwd: null pointer");
    B200SEG_CHECK(n > 0 && h > 0 && w > 0 && th >= h && tw >= w && nc > 0 && nc <= c_total && nc <= 32 && c_total % 4 == 0,
                  "lowres_ce_bwd: bad geometry (class map %dx%d, targets %dx%d, nc %d of %d)", h, w, th, tw, nc, c_total);
    B200SEG_CHECK(((reinterpret_cast<uintptr_t>(lr) | reinterpret_cast<uintptr_t>(dlr)) & 15) == 0,
                  "lowres_ce_bwd: pointers not 16-byte aligned");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    B200SEG_CUDA(cudaMemsetAsync(dlr, 0, sizeof(float) * static_cast<size_t>(n) * h * w * c_total, st));
    return launch_lowres_ce(lr, n, h, w, c_total, nc, targets, th, tw, align_corners, ignore_index, nullptr, nullptr, nullptr,
                            dlr, grad_out, valid_count, st);
}

int b200seg_cast_add_bf16(const float* a, const void* b, void* out, int64_t count, b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    B200SEG_CHECK(a && out && count > 0 && count % 4 == 0, "cast_add_bf16: null pointer or count %lld not a multiple of 4",
                  static_cast<long long>(count));
    B200SEG_CHECK((reinterpret_cast<uintptr_t>(a) & 15) == 0 && (reinterpret_cast<uintptr_t>(out) & 7) == 0 &&
                      (reinterpret_cast<uintptr_t>(b) & 7) == 0, "cast_add_bf16: misaligned pointer");
    const int64_t n4 = count / 4;
    const int blocks = static_cast<int>(n4 / 256 + 1 < 148 * 8 ? n4 / 256 + 1 : 148 * 8);
    B200SEG_CUDA(launch_pdl(cast_add_bf16_kernel, blocks, 256, 0, static_cast<cudaStream_t>(stream),
                            reinterpret_cast<const float4*>(a), static_cast<const uint2*>(b), static_cast<uint2*>(out), n4));
    return 0;
}

int b200seg_ce_finalize(const double* loss_sum, const int64_t* valid_count, const int* status, float* loss,
                        b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    B200SEG_CHECK(loss_sum && valid_count && status && loss, "ce_finalize: null pointer");
    B200SEG_CUDA(launch_pdl(ce_finalize_kernel, 1, 32, 0, static_cast<cudaStream_t>(stream), loss_sum,
                            reinterpret_cast<const unsigned long long*>(valid_count), status, loss));
    return 0;
}

int b200seg_softmax_ce_bwd(const float* logits, const int64_t* targets, int n, int nc, int64_t hw, int64_t ignore_index,
                           const float* grad_out, const int64_t* valid_count, float* dlogits, b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    B200SEG_CHECK(logits && targets && grad_out && valid_count && dlogits && n > 0 && nc > 0 && hw > 0,
                  "softmax_ce_bwd: bad arguments");
    return launch_softmax_ce(logits, targets, n, nc, hw, ignore_index, 1.0f, nullptr, nullptr, dlogits, nullptr, grad_out,
                             valid_count, static_cast<cudaStream_t>(stream));
}

int b200seg_argmax(const float* logits, int n, int nc, int64_t hw, int64_t* pred, b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    B200SEG_CHECK(logits && pred && n > 0 && nc > 0 && hw > 0, "argmax: bad arguments");
    const int64_t total = static_cast<int64_t>(n) * hw;
    B200SEG_CUDA(launch_pdl(argmax_kernel, elementwise_grid(total), 256, 0, static_cast<cudaStream_t>(stream), logits, nc, hw, total, pred));
    return 0;
}

int b200seg_confusion(const int64_t* pred, const int64_t* targets, int64_t count, int nc, int64_t* cm,
                      b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    B200SEG_CHECK(pred && targets && cm && count >= 0, "confusion: bad arguments");
    if (count == 0) return 0;
    int copies, grid;
    size_t smem;
    if (hist_launch_geometry(nc, count, &copies, &smem, &grid)) return 1;
    B200SEG_CUDA(launch_pdl(confusion_kernel<false>, grid, kHistThreads, smem, static_cast<cudaStream_t>(stream), 
        nullptr, pred, targets, nc, 1, count, copies, reinterpret_cast<unsigned long long*>(cm)));
    return 0;
}

int b200seg_argmax_confusion(const float* logits, const int64_t* targets, int n, int nc, int64_t hw, int64_t* cm,
                             b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    B200SEG_CHECK(logits && targets && cm && n > 0 && hw > 0, "argmax_confusion: bad arguments");
    int copies, grid;
    size_t smem;
    const int64_t total = static_cast<int64_t>(n) * hw;
    if (hist_launch_geometry(nc, total, &copies, &smem, &grid)) return 1;
    B200SEG_CUDA(launch_pdl(confusion_kernel<true>, grid, kHistThreads, smem, static_cast<cudaStream_t>(stream), 
        logits, nullptr, targets, nc, hw, total, copies, reinterpret_cast<unsigned long long*>(cm)));
    return 0;
}

}  // extern "C"
