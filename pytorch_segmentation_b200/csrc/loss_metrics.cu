// Class-map resampling, pixel-wise softmax cross-entropy and the mIoU confusion histogram.
// All HBM-bound.  Logits at the drop-in boundary are NCHW fp32 (what the reference's forward returns), so the
// per-pixel kernels put consecutive threads on consecutive pixels of one class plane (coalesced 128-byte rows)
// and walk the class dimension in registers.
#include "common.cuh"
#include "pdl.cuh"
#include "vec.cuh"

namespace b200seg {
namespace {

// Targets at the boundary are int64 (what the reference's post_fetch_fn hands on: ``segs.long()``, utils/datasets.py:213)
// or the uint8 / int32 masks the dataset holds before that cast (utils/datasets.py:161-194) — B200SEG_TARGET_*: reading the
// narrow type directly saves the 8x wider copy (33.5 MB -> 4.2 MB per 16 x 512 x 512 batch, host -> device included).
__device__ __forceinline__ int64_t load_target(const void* t, int kind, int64_t i) {
    if (kind == B200SEG_TARGET_I64) return __ldcs(static_cast<const long long*>(t) + i);
    if (kind == B200SEG_TARGET_I32) return __ldcs(static_cast<const int*>(t) + i);
    return __ldcs(static_cast<const unsigned char*>(t) + i);
}

// One bilinear tap set, the same instruction sequence wherever full-resolution logits are formed from the class map
// (logits_up_fwd_kernel writes them, class_map_confusion_kernel takes their argmax without writing them): explicit
// round-to-nearest intrinsics, so the compiler cannot contract the two sites differently and the fused eval path is
// bit-identical to argmax over the written logits.
__device__ __forceinline__ float lerp2d(float wy0, float wy1, float wx0, float wx1, float a, float b, float c, float d) {
    const float top = __fmaf_rn(wx1, b, __fmul_rn(wx0, a));
    const float bot = __fmaf_rn(wx1, d, __fmul_rn(wx0, c));
    return __fmaf_rn(wy1, bot, __fmul_rn(wy0, top));
}

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
            const float v[4] = {lerp2d(ly.w0, ly.w1, lx.w0, lx.w1, a.x, bb.x, cc.x, d.x),
                                lerp2d(ly.w0, ly.w1, lx.w0, lx.w1, a.y, bb.y, cc.y, d.y),
                                lerp2d(ly.w0, ly.w1, lx.w0, lx.w1, a.z, bb.z, cc.z, d.z),
                                lerp2d(ly.w0, ly.w1, lx.w0, lx.w1, a.w, bb.w, cc.w, d.w)};
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
__global__ void __launch_bounds__(256) softmax_ce_kernel(const float* __restrict__ logits, const void* __restrict__ tgt, int tk,
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
        const int64_t t = load_target(tgt, tk, i);
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
                                                                  const void* __restrict__ tgt, int tk, int nc, int64_t hw,
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
        const int64_t t = load_target(tgt, tk, i);
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
                                                                  const void* __restrict__ tgt, int tk, int nc, int64_t hw,
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
        const int64_t t = load_target(tgt, tk, i);
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


// Evaluation straight from the LOW-RESOLUTION class map (test.py:28-46 on the output of models/deeplabv3plus.py:40-43):
// prediction = argmax over classes of the bilinear upsample, confusion histogram against the targets — the
// full-resolution logits (352 MB for 16 x 21 x 512 x 512) are neither read nor needed: 11 MB class map + targets in.
// One thread per target pixel: its four source cells' class vectors come from L1/L2 (every cell serves ~16 pixels),
// the interpolation is lerp2d — the very instruction sequence logits_up_fwd_kernel stores — so the prediction is
// BIT-IDENTICAL to argmax over the written logits.  Tie rule (= argmax_planes = ``outputs.max(1)[1]``'s first
// maximum): strict '>' in ascending class order, lowest index wins; the first NaN wins over everything.
__global__ void __launch_bounds__(kHistThreads) class_map_confusion_kernel(const float* __restrict__ x, int ct, int h, int w,
                                                                            int nc, const void* __restrict__ tgt, int tk,
                                                                            int oh, int ow, int64_t total, float sh, float sw,
                                                                            int ac, int copies,
                                                                            unsigned long long* __restrict__ cm,
                                                                            int64_t* __restrict__ pred_out) {
    pdl_prologue();
    extern __shared__ unsigned int hist[];
    const int bins = nc * nc;
    for (int b = threadIdx.x; b < bins * copies; b += blockDim.x) hist[b] = 0;
    __syncthreads();
    unsigned int* mine = hist + ((threadIdx.x >> 5) % copies) * bins;
    for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < total;
         i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        const int ox = static_cast<int>(i % ow);
        const int64_t p = i / ow;
        const int oy = static_cast<int>(p % oh);
        const int n = static_cast<int>(p / oh);
        const Lerp ly = bilinear_src(sh, oy, h, ac), lx = bilinear_src(sw, ox, w, ac);
        const float* b = x + static_cast<int64_t>(n) * h * w * ct;
        const float4* p00 = reinterpret_cast<const float4*>(b + (static_cast<int64_t>(ly.i0) * w + lx.i0) * ct);
        const float4* p01 = reinterpret_cast<const float4*>(b + (static_cast<int64_t>(ly.i0) * w + lx.i1) * ct);
        const float4* p10 = reinterpret_cast<const float4*>(b + (static_cast<int64_t>(ly.i1) * w + lx.i0) * ct);
        const float4* p11 = reinterpret_cast<const float4*>(b + (static_cast<int64_t>(ly.i1) * w + lx.i1) * ct);
        float best = -INFINITY;
        int bi = 0;
        for (int c4 = 0; c4 * 4 < nc; ++c4) {
            const float4 a = __ldg(p00 + c4), bb = __ldg(p01 + c4), cc = __ldg(p10 + c4), d = __ldg(p11 + c4);
            const float v[4] = {lerp2d(ly.w0, ly.w1, lx.w0, lx.w1, a.x, bb.x, cc.x, d.x),
                                lerp2d(ly.w0, ly.w1, lx.w0, lx.w1, a.y, bb.y, cc.y, d.y),
                                lerp2d(ly.w0, ly.w1, lx.w0, lx.w1, a.z, bb.z, cc.z, d.z),
                                lerp2d(ly.w0, ly.w1, lx.w0, lx.w1, a.w, bb.w, cc.w, d.w)};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int c = c4 * 4 + j;
                if (c < nc && (v[j] > best || (v[j] != v[j] && best == best))) {
                    best = v[j];
                    bi = c;
                }
            }
        }
        if (pred_out) pred_out[i] = bi;
        if (tgt) {
            const int64_t t = load_target(tgt, tk, i);
            if (t >= 0 && t < nc) atomicAdd(mine + static_cast<int>(t) * nc + bi, 1u);
        }
    }
    __syncthreads();
    if (cm)
        for (int b = threadIdx.x; b < bins; b += blockDim.x) {
            unsigned long long sacc = 0;
            for (int k = 0; k < copies; ++k) sacc += hist[k * bins + b];
            if (sacc) atomicAdd(cm + b, sacc);
        }
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
// emits the gradient of the class map directly instead of writing 352 MB of dlogits and reading them again in the
// upsample adjoint.
//
// One CTA = a 32 x 128 tile of target pixels, its class-map footprint (FH x FW cells) staged in shared memory with an
// odd cell stride.  One thread = a 4 x 4 patch of pixels: the patch reads an NCELL x NCELL window of cells (3 for an
// upsample of x3 or more, 4 down to x1.5), so per class it loads 9-16 values for 16 pixels, and both the interpolation
// and its adjoint are two small dense contractions with per-thread constant weight tables (zeros where a pixel does not
// touch a cell) — no divergent indexing, no shared-memory float atomics (those are CAS loops on this architecture).
// Passes over the classes: max, sum (+ target logit), and in backward a third that forms (softmax - onehot) * g / count,
// contracts it to the patch's cell window, and exchanges the windows through shared memory: each footprint cell has a
// K x K group of slots, one per patch whose window covers it (the covering patches are K or fewer consecutive patch
// columns / rows, so a patch's slot is its distance from the first of them), and one thread per cell adds the group up
// (a deterministic gather) into the staged tile, which is flushed with one coalesced fp32 atomic per element (tiles
// overlap by the interpolation halo).
constexpr int kCmP = 4, kCmPX = 32, kCmPY = 8;                 // patch edge, patches per tile in x / y
constexpr int kCmTW = kCmP * kCmPX, kCmTH = kCmP * kCmPY;      // 128 x 32 target pixels per CTA

// exchange geometry: K = most patch columns (rows) whose window can cover one cell, KS = floats per cell group
template <int NCELL> struct CmExchange;
template <> struct CmExchange<3> { static constexpr int K = 4, KS = 20; };    // 16 slots + 4 pad: conflict-free LDS.128
template <> struct CmExchange<4> { static constexpr int K = 3, KS = 9; };

// MODE: 0 = loss accumulators only, 1 = gradient only (scaled by *g_dev / *count_dev), 2 = both in one pass over the
// tile with the gradient left UNSCALED (the pixel count is only known when every tile is done; the caller scales) — a
// training step then interpolates the logits three times per class instead of five.
template <int NCELL, int MODE>
__global__ void __launch_bounds__(256, 2) class_map_ce_kernel(const float* __restrict__ lr, int ct, int h, int w, int nc,
                                                              const void* __restrict__ tgt, int tk, int H, int W, float sh,
                                                              float sw, int ac, int64_t ignore_index, double* loss_sum,
                                                              unsigned long long* valid_count, int* status,
                                                              float* __restrict__ dlr, const float* __restrict__ g_dev,
                                                              const unsigned long long* __restrict__ count_dev, int FH,
                                                              int FW, int nbuf) {
    pdl_prologue();
    constexpr bool BWD = MODE != 0, FWD = MODE != 1;
    constexpr int P = kCmP, K = CmExchange<NCELL>::K, KS = CmExchange<NCELL>::KS;
    extern __shared__ __align__(16) float lsm[];
    __shared__ int bxs[kCmPX], bys[kCmPY];
    __shared__ int xlo[160], ylo[64];                         // FW <= 160, FH <= 64 (checked by the launcher)
    const int ctp = ct | 1;                                   // ct % 4 == 0 -> odd stride, conflict-free across cells
    const int cells = FH * FW;
    float* L = lsm;                                           // [cells][ctp]; in backward overwritten class by class with the gradient
    float* T = lsm + ((cells * ctp + 3) & ~3);                // [nbuf][cells][KS] exchange slots (backward)
    const int n = blockIdx.z, ty0 = blockIdx.y * kCmTH, tx0 = blockIdx.x * kCmTW;
    const int fy0 = bilinear_src(sh, ty0, h, ac).i0, fx0 = bilinear_src(sw, tx0, w, ac).i0;
    const float* base = lr + static_cast<int64_t>(n) * h * w * ct;
    const int ctq = ct >> 2;
    for (int i = threadIdx.x; i < cells * ctq; i += blockDim.x) {
        const int q = i % ctq, cell = i / ctq;
        const int yy = fy0 + cell / FW, xx = fx0 + cell % FW;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (yy < h && xx < w) v = __ldg(reinterpret_cast<const float4*>(base + (static_cast<int64_t>(yy) * w + xx) * ct) + q);
        float* d = L + cell * ctp + 4 * q;
        d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w;
    }
    const int pxi = threadIdx.x & (kCmPX - 1), pyi = threadIdx.x / kCmPX;
    // per-thread weight tables: WX[j][cell] = weight of cell (window-relative) in pixel column j of the patch
    float WX[P][NCELL], WY[P][NCELL];
    int bx, by;
    {
        const Lerp l0 = bilinear_src(sw, min(tx0 + pxi * P, W - 1), w, ac);
        bx = l0.i0 - fx0;
#pragma unroll
        for (int j = 0; j < P; ++j) {
            const Lerp l = bilinear_src(sw, min(tx0 + pxi * P + j, W - 1), w, ac);
            const int r0 = l.i0 - fx0 - bx, r1 = l.i1 - fx0 - bx;
#pragma unroll
            for (int cx = 0; cx < NCELL; ++cx) WX[j][cx] = (cx == r0 ? l.w0 : 0.f) + (cx == r1 ? l.w1 : 0.f);
        }
        const Lerp m0 = bilinear_src(sh, min(ty0 + pyi * P, H - 1), h, ac);
        by = m0.i0 - fy0;
#pragma unroll
        for (int i = 0; i < P; ++i) {
            const Lerp l = bilinear_src(sh, min(ty0 + pyi * P + i, H - 1), h, ac);
            const int r0 = l.i0 - fy0 - by, r1 = l.i1 - fy0 - by;
#pragma unroll
            for (int cy = 0; cy < NCELL; ++cy) WY[i][cy] = (cy == r0 ? l.w0 : 0.f) + (cy == r1 ? l.w1 : 0.f);
        }
    }
    if (BWD) {
        if (pyi == 0) bxs[pxi] = bx;
        if (pxi == 0) bys[pyi] = by;
    }
    // window offsets into L (cells past the footprint edge carry zero weight; clamp them onto staged data)
    int off[NCELL][NCELL];
#pragma unroll
    for (int cy = 0; cy < NCELL; ++cy)
#pragma unroll
        for (int cx = 0; cx < NCELL; ++cx) off[cy][cx] = (min(by + cy, FH - 1) * FW + min(bx + cx, FW - 1)) * ctp;
    // targets of the patch: class code per pixel, 255 = not counted (ignored, out of range, outside the image)
    unsigned int tcode[P];
    bool bad = false;
#pragma unroll
    for (int i = 0; i < P; ++i) {
        tcode[i] = 0;
        const int oy = ty0 + pyi * P + i;
#pragma unroll
        for (int j = 0; j < P; ++j) {
            const int ox = tx0 + pxi * P + j;
            unsigned int code = 255u;
            if (oy < H && ox < W) {
                const int64_t t = load_target(tgt, tk, (static_cast<int64_t>(n) * H + oy) * W + ox);
                if (t != ignore_index) {
                    if (t < 0 || t >= nc) bad = true;
                    else code = static_cast<unsigned int>(t);
                }
            }
            tcode[i] |= code << (8 * j);
        }
    }
    if (FWD && bad && status) *status = 1;
    __syncthreads();

    // interpolated logits of the patch for class c: rows of the window contracted with WX, then with WY
    auto patch_logits = [&](int c, float (&v)[P][P]) {
        float Hrow[NCELL][P];
#pragma unroll
        for (int cy = 0; cy < NCELL; ++cy) {
            float C[NCELL];
#pragma unroll
            for (int cx = 0; cx < NCELL; ++cx) C[cx] = L[off[cy][cx] + c];
#pragma unroll
            for (int j = 0; j < P; ++j) {
                float a = WX[j][0] * C[0];
#pragma unroll
                for (int cx = 1; cx < NCELL; ++cx) a = fmaf(WX[j][cx], C[cx], a);
                Hrow[cy][j] = a;
            }
        }
#pragma unroll
        for (int i = 0; i < P; ++i)
#pragma unroll
            for (int j = 0; j < P; ++j) {
                float a = WY[i][0] * Hrow[0][j];
#pragma unroll
                for (int cy = 1; cy < NCELL; ++cy) a = fmaf(WY[i][cy], Hrow[cy][j], a);
                v[i][j] = a;
            }
    };

    float m[P][P], s[P][P], vt[P][P];
#pragma unroll
    for (int i = 0; i < P; ++i)
#pragma unroll
        for (int j = 0; j < P; ++j) m[i][j] = -INFINITY, s[i][j] = 0.f, vt[i][j] = 0.f;
#pragma unroll 1
    for (int c = 0; c < nc; ++c) {
        float v[P][P];
        patch_logits(c, v);
#pragma unroll
        for (int i = 0; i < P; ++i)
#pragma unroll
            for (int j = 0; j < P; ++j) m[i][j] = fmaxf(m[i][j], v[i][j]);
    }
#pragma unroll 1
    for (int c = 0; c < nc; ++c) {
        float v[P][P];
        patch_logits(c, v);
#pragma unroll
        for (int i = 0; i < P; ++i)
#pragma unroll
            for (int j = 0; j < P; ++j) {
                s[i][j] += __expf(v[i][j] - m[i][j]);
                if (FWD) vt[i][j] = (((tcode[i] >> (8 * j)) & 255u) == static_cast<unsigned int>(c)) ? v[i][j] : vt[i][j];
            }
    }
    if (FWD) {
        double lsum = 0.0;
        unsigned int lcount = 0;
#pragma unroll
        for (int i = 0; i < P; ++i)
#pragma unroll
            for (int j = 0; j < P; ++j)
                if (((tcode[i] >> (8 * j)) & 255u) != 255u) {
                    lsum += static_cast<double>(__logf(s[i][j]) + m[i][j] - vt[i][j]);
                    ++lcount;
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
        if (!BWD) return;
    }

    // ---- backward ----
    // first patch column / row whose window covers footprint column X / row Y (window origins are monotonic)
    for (int X = threadIdx.x; X < FW; X += blockDim.x) {
        int lo = 0;
        while (lo < kCmPX - 1 && bxs[lo] + NCELL <= X) ++lo;
        xlo[X] = lo;
    }
    for (int Y = threadIdx.x; Y < FH; Y += blockDim.x) {
        int lo = 0;
        while (lo < kCmPY - 1 && bys[lo] + NCELL <= Y) ++lo;
        ylo[Y] = lo;
    }
    for (int i = threadIdx.x; i < nbuf * cells * KS; i += blockDim.x) T[i] = 0.f;   // slots nobody owns stay zero
    const float gscale = MODE == 2 ? 1.f : __ldg(g_dev) / static_cast<float>(__ldg(count_dev));
#pragma unroll
    for (int i = 0; i < P; ++i)
#pragma unroll
        for (int j = 0; j < P; ++j) s[i][j] = (((tcode[i] >> (8 * j)) & 255u) != 255u) ? gscale / s[i][j] : 0.f;
    __syncthreads();
    // this patch's slot for each cell of its window (-1: the cell lies outside the footprint, its weights are zero)
    int wofs[NCELL][NCELL];
#pragma unroll
    for (int cy = 0; cy < NCELL; ++cy)
#pragma unroll
        for (int cx = 0; cx < NCELL; ++cx) {
            wofs[cy][cx] = -1;
            if (by + cy < FH && bx + cx < FW) {
                const int ky = pyi - ylo[by + cy], kx = pxi - xlo[bx + cx];
                if (ky >= 0 && ky < K && kx >= 0 && kx < K) wofs[cy][cx] = ((by + cy) * FW + bx + cx) * KS + ky * K + kx;
            }
        }
    int buf = 0;
#pragma unroll 1
    for (int c = 0; c < nc; ++c) {
        float v[P][P];
        patch_logits(c, v);
        float Hacc[P][NCELL];
#pragma unroll
        for (int i = 0; i < P; ++i) {
#pragma unroll
            for (int j = 0; j < P; ++j) {
                const bool hit = ((tcode[i] >> (8 * j)) & 255u) == static_cast<unsigned int>(c);
                v[i][j] = __expf(v[i][j] - m[i][j]) * s[i][j] - (hit ? gscale : 0.f);
            }
#pragma unroll
            for (int cx = 0; cx < NCELL; ++cx) {
                float a = WX[0][cx] * v[i][0];
#pragma unroll
                for (int j = 1; j < P; ++j) a = fmaf(WX[j][cx], v[i][j], a);
                Hacc[i][cx] = a;
            }
        }
        float* Tb = T + buf * cells * KS;
#pragma unroll
        for (int cy = 0; cy < NCELL; ++cy)
#pragma unroll
            for (int cx = 0; cx < NCELL; ++cx) {
                float a = WY[0][cy] * Hacc[0][cx];
#pragma unroll
                for (int i = 1; i < P; ++i) a = fmaf(WY[i][cy], Hacc[i][cx], a);
                if (wofs[cy][cx] >= 0) Tb[wofs[cy][cx]] = a;
            }
        __syncthreads();
        for (int cell = threadIdx.x; cell < cells; cell += blockDim.x) {
            float a;
            if (NCELL == 3) {
                const float4* q = reinterpret_cast<const float4*>(Tb + cell * KS);
                const float4 a0 = q[0], a1 = q[1], a2 = q[2], a3 = q[3];
                a = ((a0.x + a0.y) + (a0.z + a0.w)) + ((a1.x + a1.y) + (a1.z + a1.w)) +
                    (((a2.x + a2.y) + (a2.z + a2.w)) + ((a3.x + a3.y) + (a3.z + a3.w)));
            } else {
                a = 0.f;
#pragma unroll
                for (int k = 0; k < K * K; ++k) a += Tb[cell * KS + k];
            }
            L[cell * ctp + c] = a;        // class c of the staged map is dead from here on: reuse it for the gradient
        }
        if (nbuf == 2) buf ^= 1;
        else __syncthreads();
    }
    __syncthreads();
    float* out = dlr + static_cast<int64_t>(n) * h * w * ct;
    for (int i = threadIdx.x; i < cells * ct; i += blockDim.x) {
        const int c = i % ct, cell = i / ct;
        const int yy = fy0 + cell / FW, xx = fx0 + cell % FW;
        if (c < nc && yy < h && xx < w) {
            const float v = L[cell * ctp + c];
            if (v != 0.f) atomicAdd(out + (static_cast<int64_t>(yy) * w + xx) * ct + c, v);
        }
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

// y *= *g / *count, in place: the scale the fused loss + gradient pass could not know (mean over the valid pixels times
// the upstream gradient), both read on the device
__global__ void scale_by_ratio_kernel(float4* __restrict__ y, int64_t n4, const float* __restrict__ g,
                                      const unsigned long long* __restrict__ count) {
    pdl_prologue();
    const float sc = __ldg(g) / static_cast<float>(__ldg(count));
    for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < n4;
         i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        float4 v = y[i];
        v.x *= sc; v.y *= sc; v.z *= sc; v.w *= sc;
        y[i] = v;
    }
}

int launch_lowres_ce(const float* lr, int n, int h, int w, int ct, int nc, const void* targets, int tk, int H, int W,
                     int align_corners, int64_t ignore_index, double* loss_sum, int64_t* valid_count, int* status, float* dlr,
                     const float* g_dev, const int64_t* count_dev, cudaStream_t st) {
    // dlr == nullptr: loss only; dlr with g_dev: gradient only; dlr without g_dev: both, gradient unscaled
    const int mode = !dlr ? 0 : (g_dev ? 1 : 2);
    const float sh = bilinear_scale(h, H, align_corners), sw = bilinear_scale(w, W, align_corners);
    // cells a 4-pixel run can touch: floor(3 * scale) + 3
    const float smax = sh > sw ? sh : sw;
    const int ncell = static_cast<int>(3.f * smax) + 3;
    B200SEG_CHECK(ncell <= 4, "lowres_ce: upsample factor below 1.5 (class map %dx%d, targets %dx%d)", h, w, H, W);
    // footprint of a 32 x 128 target tile in the class map (+1 for the second lerp tap, +1 for rounding)
    const int FH = static_cast<int>(kCmTH * sh) + 3 < h ? static_cast<int>(kCmTH * sh) + 3 : h;
    const int FW = static_cast<int>(kCmTW * sw) + 3 < w ? static_cast<int>(kCmTW * sw) + 3 : w;
    B200SEG_CHECK(FH <= 64 && FW <= 160, "lowres_ce: footprint %dx%d too large", FH, FW);
    // patch columns covering one cell: window origins advance by >= floor(4 * scale) cells per patch
    const float smin = sh < sw ? sh : sw;
    const int kmax = ncell == 3 ? 4 : 3, ks = ncell == 3 ? 20 : 9;
    B200SEG_CHECK(smin > 0.f && static_cast<int>((ncell - 1) / (4.f * smin)) + 2 <= kmax,
                  "lowres_ce: upsample factor not supported (class map %dx%d, targets %dx%d)", h, w, H, W);
    const size_t lbytes = sizeof(float) * ((static_cast<size_t>(FH) * FW * (ct | 1) + 3) & ~size_t(3));
    const size_t tbytes = sizeof(float) * static_cast<size_t>(FH) * FW * ks;
    const size_t limit = 110 * 1024;
    int nbuf = dlr ? (lbytes + 2 * tbytes <= limit ? 2 : 1) : 0;
    const size_t smem = lbytes + nbuf * tbytes;
    B200SEG_CHECK(smem <= limit, "lowres_ce: class-map footprint of %zu bytes per tile exceeds shared memory", smem);
    B200SEG_CHECK(n <= 65535 && ceil_div(H, kCmTH) <= 65535, "lowres_ce: grid too large");
    dim3 grid(ceil_div(W, kCmTW), ceil_div(H, kCmTH), n);
    unsigned long long* vc = reinterpret_cast<unsigned long long*>(valid_count);
    const unsigned long long* cd = reinterpret_cast<const unsigned long long*>(count_dev);
#define B200SEG_LRCE(NCELL, B)                                                                                             \
    do {                                                                                                                   \
        static bool attr_set = false;                                                                                      \
        if (!attr_set) {                                                                                                   \
            B200SEG_CUDA(cudaFuncSetAttribute(class_map_ce_kernel<NCELL, B>, cudaFuncAttributeMaxDynamicSharedMemorySize,  \
                                              110 * 1024));                                                                \
            attr_set = true;                                                                                               \
        }                                                                                                                  \
        B200SEG_CUDA(launch_pdl(class_map_ce_kernel<NCELL, B>, grid, 256, smem, st, lr, ct, h, w, nc, targets, tk, H, W, sh,\
                                sw, align_corners, ignore_index, loss_sum, vc, status, dlr, g_dev, cd, FH, FW, nbuf));     \
    } while (0)
    if (mode == 1) {
        if (ncell == 3) B200SEG_LRCE(3, 1);
        else B200SEG_LRCE(4, 1);
    } else if (mode == 2) {
        if (ncell == 3) B200SEG_LRCE(3, 2);
        else B200SEG_LRCE(4, 2);
    } else {
        if (ncell == 3) B200SEG_LRCE(3, 0);
        else B200SEG_LRCE(4, 0);
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

int launch_softmax_ce(const float* logits, const void* targets, int tk, int n, int nc, int64_t hw, int64_t ignore_index,
                      float grad_scale, double* loss_sum, int64_t* valid_count, float* dlogits, int* status,
                      const float* g_dev, const int64_t* count_dev, cudaStream_t st) {
    const int64_t total = static_cast<int64_t>(n) * hw;
    const int grid = elementwise_grid(total);
    unsigned long long* vc = reinterpret_cast<unsigned long long*>(valid_count);
    const unsigned long long* cd = reinterpret_cast<const unsigned long long*>(count_dev);
    if (nc <= 8)
        B200SEG_CUDA(launch_pdl(softmax_ce_kernel<8>, grid, 256, 0, st, logits, targets, tk, nc, hw, total, ignore_index, grad_scale,
                                loss_sum, vc, dlogits, status, g_dev, cd));
    else if (nc <= 24)
        B200SEG_CUDA(launch_pdl(softmax_ce_kernel<24>, grid, 256, 0, st, logits, targets, tk, nc, hw, total, ignore_index, grad_scale,
                                loss_sum, vc, dlogits, status, g_dev, cd));
    else if (nc <= 32)
        B200SEG_CUDA(launch_pdl(softmax_ce_kernel<32>, grid, 256, 0, st, logits, targets, tk, nc, hw, total, ignore_index, grad_scale,
                                loss_sum, vc, dlogits, status, g_dev, cd));
    else
        B200SEG_CUDA(launch_pdl(softmax_ce_generic_kernel, grid, 256, 0, st, logits, targets, tk, nc, hw, total, ignore_index,
                                grad_scale, loss_sum, vc, dlogits, status, g_dev, cd));
    return 0;
}
}  // namespace
}  // namespace b200seg

extern "C" {

#define B200SEG_CHECK_TK(what)                                                                                       \
    B200SEG_CHECK(target_kind == B200SEG_TARGET_I64 || target_kind == B200SEG_TARGET_I32 || target_kind == B200SEG_TARGET_U8, \
                  what ": target_kind %d", target_kind)

int b200seg_softmax_ce(const float* logits, const void* targets, int target_kind, int n, int nc, int64_t hw,
                       int64_t ignore_index, float grad_scale, double* loss_sum, int64_t* valid_count, float* dlogits,
                       int* status, b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    B200SEG_CHECK(logits && targets && loss_sum && status && n > 0 && nc > 0 && hw > 0, "softmax_ce: bad arguments");
    B200SEG_CHECK_TK("softmax_ce");
    return launch_softmax_ce(logits, targets, target_kind, n, nc, hw, ignore_index, grad_scale, loss_sum, valid_count, dlogits, status,
                             nullptr, nullptr, static_cast<cudaStream_t>(stream));
}

int b200seg_lowres_ce_fwd(const float* lr, int n, int h, int w, int c_total, int nc, const void* targets, int target_kind,
                          int th, int tw, int align_corners, int64_t ignore_index, double* loss_sum, int64_t* valid_count, int* status,
                          b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    B200SEG_CHECK(lr && targets && loss_sum && valid_count && status, "lowres_ce_fwd: null pointer");
    B200SEG_CHECK(n > 0 && h > 0 && w > 0 && th >= h && tw >= w && nc > 0 && nc <= c_total && nc <= 32 && c_total % 4 == 0,
                  "lowres_ce_fwd: bad geometry (class map %dx%d, targets %dx%d, nc %d of %d)", h, w, th, tw, nc, c_total);
    B200SEG_CHECK((reinterpret_cast<uintptr_t>(lr) & 15) == 0, "lowres_ce_fwd: class map not 16-byte aligned");
    B200SEG_CHECK_TK("lowres_ce_fwd");
    return launch_lowres_ce(lr, n, h, w, c_total, nc, targets, target_kind, th, tw, align_corners, ignore_index, loss_sum, valid_count,
                            status, nullptr, nullptr, nullptr, static_cast<cudaStream_t>(stream));
}

int b200seg_lowres_ce_bwd(const float* lr, int n, int h, int w, int c_total, int nc, const void* targets, int target_kind,
                          int th, int tw, int align_corners, int64_t ignore_index, const float* grad_out, const int64_t* valid_count,
                          float* dlr, b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    B200SEG_CHECK(lr && targets && grad_out && valid_count && dlr, "lowres_ce_bwd: null pointer");
    B200SEG_CHECK(n > 0 && h > 0 && w > 0 && th >= h && tw >= w && nc > 0 && nc <= c_total && nc <= 32 && c_total % 4 == 0,
                  "lowres_ce_bwd: bad geometry (class map %dx%d, targets %dx%d, nc %d of %d)", h, w, th, tw, nc, c_total);
    B200SEG_CHECK(((reinterpret_cast<uintptr_t>(lr) | reinterpret_cast<uintptr_t>(dlr)) & 15) == 0,
                  "lowres_ce_bwd: pointers not 16-byte aligned");
    B200SEG_CHECK_TK("lowres_ce_bwd");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    B200SEG_CUDA(cudaMemsetAsync(dlr, 0, sizeof(float) * static_cast<size_t>(n) * h * w * c_total, st));
    return launch_lowres_ce(lr, n, h, w, c_total, nc, targets, target_kind, th, tw, align_corners, ignore_index, nullptr, nullptr, nullptr,
                            dlr, grad_out, valid_count, st);
}

int b200seg_lowres_ce_fused(const float* lr, int n, int h, int w, int c_total, int nc, const void* targets, int target_kind,
                            int th, int tw, int align_corners, int64_t ignore_index, double* loss_sum, int64_t* valid_count,
                            int* status, float* dlr, b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    B200SEG_CHECK(lr && targets && loss_sum && valid_count && status && dlr, "lowres_ce_fused: null pointer");
    B200SEG_CHECK(n > 0 && h > 0 && w > 0 && th >= h && tw >= w && nc > 0 && nc <= c_total && nc <= 32 && c_total % 4 == 0,
                  "lowres_ce_fused: bad geometry (class map %dx%d, targets %dx%d, nc %d of %d)", h, w, th, tw, nc, c_total);
    B200SEG_CHECK(((reinterpret_cast<uintptr_t>(lr) | reinterpret_cast<uintptr_t>(dlr)) & 15) == 0,
                  "lowres_ce_fused: pointers not 16-byte aligned");
    B200SEG_CHECK_TK("lowres_ce_fused");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    B200SEG_CUDA(cudaMemsetAsync(dlr, 0, sizeof(float) * static_cast<size_t>(n) * h * w * c_total, st));
    return launch_lowres_ce(lr, n, h, w, c_total, nc, targets, target_kind, th, tw, align_corners, ignore_index, loss_sum, valid_count,
                            status, dlr, nullptr, nullptr, st);
}

int b200seg_scale_by_ratio(float* y, int64_t count, const float* grad_out, const int64_t* valid_count, b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    B200SEG_CHECK(y && grad_out && valid_count && count > 0 && count % 4 == 0 && (reinterpret_cast<uintptr_t>(y) & 15) == 0,
                  "scale_by_ratio: null / misaligned pointer or count %lld not a multiple of 4", static_cast<long long>(count));
    B200SEG_CUDA(launch_pdl(scale_by_ratio_kernel, elementwise_grid(count / 4), 256, 0, static_cast<cudaStream_t>(stream),
                            reinterpret_cast<float4*>(y), count / 4, grad_out,
                            reinterpret_cast<const unsigned long long*>(valid_count)));
    return 0;
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

int b200seg_softmax_ce_bwd(const float* logits, const void* targets, int target_kind, int n, int nc, int64_t hw,
                           int64_t ignore_index, const float* grad_out, const int64_t* valid_count, float* dlogits, b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    B200SEG_CHECK(logits && targets && grad_out && valid_count && dlogits && n > 0 && nc > 0 && hw > 0,
                  "softmax_ce_bwd: bad arguments");
    B200SEG_CHECK_TK("softmax_ce_bwd");
    return launch_softmax_ce(logits, targets, target_kind, n, nc, hw, ignore_index, 1.0f, nullptr, nullptr, dlogits, nullptr, grad_out,
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

int b200seg_confusion(const int64_t* pred, const void* targets, int target_kind, int64_t count, int nc, int64_t* cm,
                      b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    B200SEG_CHECK(pred && targets && cm && count >= 0, "confusion: bad arguments");
    B200SEG_CHECK_TK("confusion");
    if (count == 0) return 0;
    int copies, grid;
    size_t smem;
    if (hist_launch_geometry(nc, count, &copies, &smem, &grid)) return 1;
    B200SEG_CUDA(launch_pdl(confusion_kernel<false>, grid, kHistThreads, smem, static_cast<cudaStream_t>(stream), 
        nullptr, pred, targets, target_kind, nc, 1, count, copies, reinterpret_cast<unsigned long long*>(cm)));
    return 0;
}

int b200seg_argmax_confusion(const float* logits, const void* targets, int target_kind, int n, int nc, int64_t hw,
                             int64_t* cm, b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    B200SEG_CHECK(logits && targets && cm && n > 0 && hw > 0, "argmax_confusion: bad arguments");
    B200SEG_CHECK_TK("argmax_confusion");
    int copies, grid;
    size_t smem;
    const int64_t total = static_cast<int64_t>(n) * hw;
    if (hist_launch_geometry(nc, total, &copies, &smem, &grid)) return 1;
    B200SEG_CUDA(launch_pdl(confusion_kernel<true>, grid, kHistThreads, smem, static_cast<cudaStream_t>(stream), 
        logits, nullptr, targets, target_kind, nc, hw, total, copies, reinterpret_cast<unsigned long long*>(cm)));
    return 0;
}

int b200seg_class_map_confusion(const float* lr, int n, int h, int w, int c_total, int nc, const void* targets,
                                int target_kind, int th, int tw, int align_corners, int64_t* cm, int64_t* pred,
                                b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    B200SEG_CHECK(lr && (pred || (targets && cm)), "class_map_confusion: null pointer");
    B200SEG_CHECK(n > 0 && h > 0 && w > 0 && th > 0 && tw > 0 && nc > 0 && nc <= c_total && c_total % 4 == 0,
                  "class_map_confusion: bad geometry (class map %dx%d, targets %dx%d, nc %d of %d)", h, w, th, tw, nc, c_total);
    B200SEG_CHECK((reinterpret_cast<uintptr_t>(lr) & 15) == 0, "class_map_confusion: class map not 16-byte aligned");
    if (targets) B200SEG_CHECK_TK("class_map_confusion");
    int copies, grid;
    size_t smem;
    const int64_t total = static_cast<int64_t>(n) * th * tw;
    if (hist_launch_geometry(nc, total, &copies, &smem, &grid)) return 1;
    const float sh = bilinear_scale(h, th, align_corners), sw = bilinear_scale(w, tw, align_corners);
    B200SEG_CUDA(launch_pdl(class_map_confusion_kernel, grid, kHistThreads, smem, static_cast<cudaStream_t>(stream), lr,
                            c_total, h, w, nc, cm ? targets : nullptr, target_kind, th, tw, total, sh, sw, align_corners,
                            copies, reinterpret_cast<unsigned long long*>(cm), pred));
    return 0;
}

}  // extern "C"
