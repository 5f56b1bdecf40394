// Multi-tensor kernels for the parameter side of a training step: the SGD(momentum) update of ALL
// parameters and the bf16 re-packing of ALL conv weights, each in one or two launches instead of one per tensor.
// The per-tensor descriptors travel BY VALUE in the kernel parameters (gradient pointers change from step to step
// in eager mode, so there is no device-side table to keep in sync; a CUDA-graph capture records them as constants).
// Both are pure HBM streams: fp32 p/g/m in, p/m out for the update (20 B per element), 4 B in / 4 B out for the pack.
#include "common.cuh"
#include "pdl.cuh"
#include "vec.cuh"

#include <string.h>

namespace b200seg {
namespace {

// ---------------------------------------------------------------------------------------------------------------
// SGD with momentum — torch.optim.SGD semantics (the reference's external Trainer default, train.py:61-68):
//   g' = grad_scale * g + weight_decay * p ;  m = first ? g' : momentum * m + (1 - dampening) * g'
//   step = nesterov ? g' + momentum * m : m ;  p -= lr * step          (momentum == 0: step = g', m untouched)
// ---------------------------------------------------------------------------------------------------------------
constexpr int kSgdTensors = 96;
constexpr int kSgdElemsPerBlock = 256 * 16;

struct SgdChunk {
    b200seg_sgd_desc t[kSgdTensors];
    int blk_start[kSgdTensors + 1];
    int n;
    float lr, momentum, dampening, weight_decay, grad_scale;
    int nesterov, first;
    const float* hyper;      // optional device vector {lr, momentum, dampening, weight_decay, grad_scale}: read at run
                             // time, so a CUDA-graph replay sees the values of the moment (LR schedules)
};

struct SgdHyper {
    float lr, momentum, dampening, weight_decay, grad_scale;
};

__device__ __forceinline__ float sgd_one(float p, float g, float* m, const SgdHyper& h, bool use_m, bool first, bool nesterov) {
    g = fmaf(h.weight_decay, p, h.grad_scale * g);
    float step = g;
    if (use_m) {
        const float mm = first ? g : fmaf(h.momentum, *m, (1.f - h.dampening) * g);
        *m = mm;
        step = nesterov ? fmaf(h.momentum, mm, g) : mm;
    }
    return fmaf(-h.lr, step, p);
}

__global__ void __launch_bounds__(256) sgd_multi_kernel(const __grid_constant__ SgdChunk c) {
    pdl_prologue();
    // which tensor does this block belong to (block ranges are a prefix sum; n <= 96: binary search)
    int lo = 0, hi = c.n - 1;
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (c.blk_start[mid] <= static_cast<int>(blockIdx.x)) lo = mid; else hi = mid - 1;
    }
    const b200seg_sgd_desc& t = c.t[lo];
    const int64_t begin = static_cast<int64_t>(blockIdx.x - c.blk_start[lo]) * kSgdElemsPerBlock;
    const int64_t end = begin + kSgdElemsPerBlock < t.numel ? begin + kSgdElemsPerBlock : t.numel;
    float* p = t.p;
    const float* g = t.g;
    float* m = t.m;
    // whether a momentum buffer exists is structural (fixed at launch / capture); the VALUES may come from the device
    const bool use_m = c.momentum != 0.f;
    const bool first = c.first != 0, nesterov = c.nesterov != 0;
    SgdHyper h = {c.lr, c.momentum, c.dampening, c.weight_decay, c.grad_scale};
    if (c.hyper) {
        h.lr = __ldg(c.hyper + 0);
        h.momentum = __ldg(c.hyper + 1);
        h.dampening = __ldg(c.hyper + 2);
        h.weight_decay = __ldg(c.hyper + 3);
        h.grad_scale = __ldg(c.hyper + 4);
    }
    const bool aligned = ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) |
                           (use_m ? reinterpret_cast<uintptr_t>(m) : 0)) & 15) == 0;
    if (aligned) {
        const int64_t vend = begin + ((end - begin) & ~int64_t(3));
        for (int64_t i = begin + threadIdx.x * 4; i < vend; i += 256 * 4) {
            float4 pv = *reinterpret_cast<const float4*>(p + i);
            const float4 gv = __ldcs(reinterpret_cast<const float4*>(g + i));
            float4 mv = make_float4(0.f, 0.f, 0.f, 0.f);
            if (use_m && !c.first) mv = *reinterpret_cast<const float4*>(m + i);
            pv.x = sgd_one(pv.x, gv.x, &mv.x, h, use_m, first, nesterov);
            pv.y = sgd_one(pv.y, gv.y, &mv.y, h, use_m, first, nesterov);
            pv.z = sgd_one(pv.z, gv.z, &mv.z, h, use_m, first, nesterov);
            pv.w = sgd_one(pv.w, gv.w, &mv.w, h, use_m, first, nesterov);
            *reinterpret_cast<float4*>(p + i) = pv;
            if (use_m) *reinterpret_cast<float4*>(m + i) = mv;
        }
        for (int64_t i = vend + threadIdx.x; i < end; i += 256) {
            float mv = (use_m && !c.first) ? m[i] : 0.f;
            p[i] = sgd_one(p[i], g[i], &mv, h, use_m, first, nesterov);
            if (use_m) m[i] = mv;
        }
    } else {
        for (int64_t i = begin + threadIdx.x; i < end; i += 256) {
            float mv = (use_m && !c.first) ? m[i] : 0.f;
            p[i] = sgd_one(p[i], g[i], &mv, h, use_m, first, nesterov);
            if (use_m) m[i] = mv;
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------
// Adam / AdamW — torch.optim.Adam(W) semantics, amsgrad off (the reference's ``--adam`` switch, train.py:96, handed to
// the external Trainer):
//   g' = grad_scale * g  (+ weight_decay * p when the decay is coupled: Adam)
//   decoupled (AdamW):  p *= 1 - lr * weight_decay
//   m = m + (1 - beta1) * (g' - m) ;  v = beta2 * v + (1 - beta2) * g'^2
//   p -= (lr / bc1) * m / (sqrt(v) / sqrt(bc2) + eps)          bc_i = 1 - beta_i^step
// hyper = {lr, beta1, beta2, eps, weight_decay, grad_scale, lr/bc1, sqrt(bc2), 1-beta1, 1-beta2, 1-lr*weight_decay}
// lives in DEVICE memory so a CUDA-graph replay sees this step's learning rate and bias corrections; the derived
// entries are formed on the host in double precision like torch.optim's Python scalars (1 - 0.999 in fp32 is off by 1e-5).
// ---------------------------------------------------------------------------------------------------------------
constexpr int kAdamTensors = 64;

struct AdamChunk {
    b200seg_adam_desc t[kAdamTensors];
    int blk_start[kAdamTensors + 1];
    int n;
    int decoupled, first;
    const float* hyper;
};

struct AdamHyper {
    float lr, beta1, beta2, eps, wd, gs, step_size, sbc2, omb1, omb2, decay;
};

__device__ __forceinline__ float adam_one(float p, float g, float* m, float* v, const AdamHyper& h, bool decoupled) {
    g *= h.gs;
    if (decoupled) p = p * h.decay;
    else g = fmaf(h.wd, p, g);
    const float mm = fmaf(h.omb1, g - *m, *m);
    const float vv = fmaf(h.beta2, *v, h.omb2 * g * g);
    *m = mm;
    *v = vv;
    const float denom = __fdiv_rn(__fsqrt_rn(vv), h.sbc2) + h.eps;
    return p - h.step_size * __fdiv_rn(mm, denom);
}

__global__ void __launch_bounds__(256) adam_multi_kernel(const __grid_constant__ AdamChunk c) {
    pdl_prologue();
    int lo = 0, hi = c.n - 1;
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (c.blk_start[mid] <= static_cast<int>(blockIdx.x)) lo = mid; else hi = mid - 1;
    }
    const b200seg_adam_desc& t = c.t[lo];
    const int64_t begin = static_cast<int64_t>(blockIdx.x - c.blk_start[lo]) * kSgdElemsPerBlock;
    const int64_t end = begin + kSgdElemsPerBlock < t.numel ? begin + kSgdElemsPerBlock : t.numel;
    AdamHyper h;
    h.lr = __ldg(c.hyper + 0); h.beta1 = __ldg(c.hyper + 1); h.beta2 = __ldg(c.hyper + 2); h.eps = __ldg(c.hyper + 3);
    h.wd = __ldg(c.hyper + 4); h.gs = __ldg(c.hyper + 5); h.step_size = __ldg(c.hyper + 6); h.sbc2 = __ldg(c.hyper + 7);
    h.omb1 = __ldg(c.hyper + 8); h.omb2 = __ldg(c.hyper + 9); h.decay = __ldg(c.hyper + 10);
    const bool dec = c.decoupled != 0, first = c.first != 0;
    float* p = t.p;
    const float* g = t.g;
    float* m = t.m;
    float* v = t.v;
    const bool aligned = ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(m) |
                           reinterpret_cast<uintptr_t>(v)) & 15) == 0;
    const int64_t vend = aligned ? begin + ((end - begin) & ~int64_t(3)) : begin;
    for (int64_t i = begin + threadIdx.x * 4; i < vend; i += 256 * 4) {
        float4 pv = *reinterpret_cast<const float4*>(p + i);
        const float4 gv = __ldcs(reinterpret_cast<const float4*>(g + i));
        float4 mv = make_float4(0.f, 0.f, 0.f, 0.f), vv = mv;
        if (!first) {
            mv = *reinterpret_cast<const float4*>(m + i);
            vv = *reinterpret_cast<const float4*>(v + i);
        }
        pv.x = adam_one(pv.x, gv.x, &mv.x, &vv.x, h, dec);
        pv.y = adam_one(pv.y, gv.y, &mv.y, &vv.y, h, dec);
        pv.z = adam_one(pv.z, gv.z, &mv.z, &vv.z, h, dec);
        pv.w = adam_one(pv.w, gv.w, &mv.w, &vv.w, h, dec);
        *reinterpret_cast<float4*>(p + i) = pv;
        *reinterpret_cast<float4*>(m + i) = mv;
        *reinterpret_cast<float4*>(v + i) = vv;
    }
    for (int64_t i = vend + threadIdx.x; i < end; i += 256) {
        float mv = first ? 0.f : m[i], vv = first ? 0.f : v[i];
        p[i] = adam_one(p[i], g[i], &mv, &vv, h, dec);
        m[i] = mv;
        v[i] = vv;
    }
}

// dst <- src for many fp32 tensors at once (gradients into their slots of a flat all-reduce bucket): descriptors are
// b200seg_sgd_desc with p = destination, g = source, m unused
struct CopyChunk {
    b200seg_sgd_desc t[kSgdTensors];
    int blk_start[kSgdTensors + 1];
    int n;
};

__global__ void __launch_bounds__(256) copy_multi_kernel(const __grid_constant__ CopyChunk c) {
    pdl_prologue();
    int lo = 0, hi = c.n - 1;
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (c.blk_start[mid] <= static_cast<int>(blockIdx.x)) lo = mid; else hi = mid - 1;
    }
    const b200seg_sgd_desc& t = c.t[lo];
    const int64_t begin = static_cast<int64_t>(blockIdx.x - c.blk_start[lo]) * kSgdElemsPerBlock;
    const int64_t end = begin + kSgdElemsPerBlock < t.numel ? begin + kSgdElemsPerBlock : t.numel;
    float* dst = t.p;
    const float* src = t.g;
    if (((reinterpret_cast<uintptr_t>(dst) | reinterpret_cast<uintptr_t>(src)) & 15) == 0) {
        const int64_t vend = begin + ((end - begin) & ~int64_t(3));
        for (int64_t i = begin + threadIdx.x * 4; i < vend; i += 256 * 4)
            *reinterpret_cast<float4*>(dst + i) = __ldcs(reinterpret_cast<const float4*>(src + i));
        for (int64_t i = vend + threadIdx.x; i < end; i += 256) dst[i] = src[i];
    } else {
        for (int64_t i = begin + threadIdx.x; i < end; i += 256) dst[i] = src[i];
    }
}

// ---------------------------------------------------------------------------------------------------------------
// weight packing, all conv weights of a model in one go (same element mapping as pack_weights_kernel /
// pack_weights_im2col_kernel in pack.cu / im2col.cu)
// ---------------------------------------------------------------------------------------------------------------
constexpr int kPackTensors = 48;
constexpr int kPackTile = 32;            // a block packs a TK (k) x 32 (c) x taps tile of one weight
constexpr int kPackMaxTaps = 9;
constexpr int kPackSmemFloats = kPackTile * (kPackTile * kPackMaxTaps + 1);
// k rows per tile: as many as fit the staging buffer (32 for 3x3, 256 for 1x1 — a 32 x 32 tile of a 1x1 weight is 4
// elements per thread, and 12 M of the model's 39 M conv weights are 1x1)
__host__ __device__ inline int pack_tile_k(int taps) {
    const int rows = kPackSmemFloats / (kPackTile * taps + 1);
    const int tk = rows / kPackTile * kPackTile;
    return tk > 256 ? 256 : tk;
}

struct PackChunk {
    b200seg_pack_desc t[kPackTensors];
    int blk_start[kPackTensors + 1];
    int n;
};

// OIHW is [k][c][tap] with the taps contiguous, both packings want them outermost.  Reading one element per packed
// element strides the loads by `taps` floats (9x sector over-fetch: 335 us for DeepLabV3+'s 39 M weights); instead a
// block stages a 32 x 32 x taps tile through shared memory: 288-float contiguous runs in, 64-byte runs out
// (wf[tap][k][c..c+31], wd[tap'][c][k..k+31]).  kTaps = 1 / 9 make every index split a compile-time constant (the
// run-time divisions of the generic form, three per packed pair, were half of this kernel's time); 0 = any r x s.
template <int kTaps>
__device__ __forceinline__ void pack_tile(const b200seg_pack_desc& t, int local, float* tile) {
    const float* __restrict__ w = t.w;
    const int taps = kTaps > 0 ? kTaps : t.r * t.s;
    __nv_bfloat16* wf = static_cast<__nv_bfloat16*>(t.w_fprop);
    __nv_bfloat16* wd = static_cast<__nv_bfloat16*>(t.w_dgrad);
    const int tk = pack_tile_k(taps);
    const int tiles_c = (t.c_rows + kPackTile - 1) / kPackTile;
    const int k0 = (local / tiles_c) * tk, c0 = (local % tiles_c) * kPackTile;
    const int run = kPackTile * taps;                                  // contiguous floats per k row of the tile
    const int pitch = run + 1;
    // stage: float4 loads when every k row of the tile starts 16-byte aligned and lies inside the tensor
    const bool vec = ((static_cast<long long>(t.c) * taps) & 3) == 0 && c0 + kPackTile <= t.c &&
                     (reinterpret_cast<uintptr_t>(w) & 15) == 0;
    if (vec) {
        const int run4 = run >> 2;
        const float* base = w + (static_cast<size_t>(k0) * t.c + c0) * taps;
        const size_t k_pitch = static_cast<size_t>(t.c) * taps;
        for (int i = threadIdx.x; i < tk * run4; i += 256) {
            const int kk = i / run4, j = (i - kk * run4) << 2;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (k0 + kk < t.k) v = __ldcs(reinterpret_cast<const float4*>(base + kk * k_pitch + j));
            float* d = tile + kk * pitch + j;
            d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w;
        }
    } else {
        for (int i = threadIdx.x; i < tk * run; i += 256) {
            const int kk = i / run, j = i - kk * run;                  // j = cc * taps + tap
            const int cc = j / taps;
            float v = 0.f;
            if (k0 + kk < t.k && c0 + cc < t.c) v = __ldg(w + (static_cast<size_t>(k0 + kk) * t.c + c0) * taps + j);
            tile[kk * pitch + j] = v;
        }
    }
    __syncthreads();
    // fprop packing: [tap][k_rows][c], 32 consecutive c per (tap, k): a thread writes a bf16 pair
    if (wf) {
        const bool pair = (t.c & 1) == 0;
        constexpr int kHalf = kPackTile / 2;
        for (int i = threadIdx.x; i < taps * tk * kHalf; i += 256) {
            const int cc = (i % kHalf) * 2, kk = (i / kHalf) % tk, tap = i / (kHalf * tk);
            if (k0 + kk >= t.k_rows || c0 + cc >= t.c) continue;
            __nv_bfloat16* d = wf + (static_cast<size_t>(tap) * t.k_rows + k0 + kk) * t.c + c0 + cc;
            const float a = tile[kk * pitch + cc * taps + tap], b = tile[kk * pitch + (cc + 1) * taps + tap];
            if (pair && c0 + cc + 1 < t.c) {
                *reinterpret_cast<__nv_bfloat162*>(d) = __floats2bfloat162_rn(a, b);
            } else {
                d[0] = __float2bfloat16(a);
                if (c0 + cc + 1 < t.c) d[1] = __float2bfloat16(b);
            }
        }
    }
    // dgrad packing: [tap mirrored][c_rows][k_cols], consecutive k per (tap, c): a bf16 pair per thread
    if (wd) {
        const bool pair = (t.k_cols & 1) == 0;
        const int hk = tk / 2;
        for (int i = threadIdx.x; i < taps * kPackTile * hk; i += 256) {
            const int kk = (i % hk) * 2, cc = (i / hk) % kPackTile, tap = i / (hk * kPackTile);
            if (c0 + cc >= t.c_rows || k0 + kk >= t.k_cols) continue;
            __nv_bfloat16* d = wd + (static_cast<size_t>(taps - 1 - tap) * t.c_rows + c0 + cc) * t.k_cols + k0 + kk;
            const float a = tile[kk * pitch + cc * taps + tap], b = tile[(kk + 1) * pitch + cc * taps + tap];
            if (pair && k0 + kk + 1 < t.k_cols) {
                *reinterpret_cast<__nv_bfloat162*>(d) = __floats2bfloat162_rn(a, b);
            } else {
                d[0] = __float2bfloat16(a);
                if (k0 + kk + 1 < t.k_cols) d[1] = __float2bfloat16(b);
            }
        }
    }
}

__global__ void __launch_bounds__(256) pack_multi_kernel(const __grid_constant__ PackChunk c) {
    pdl_prologue();
    __shared__ __align__(16) float tile[kPackSmemFloats];
    int lo = 0, hi = c.n - 1;
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (c.blk_start[mid] <= static_cast<int>(blockIdx.x)) lo = mid; else hi = mid - 1;
    }
    const b200seg_pack_desc& t = c.t[lo];
    const int local = blockIdx.x - c.blk_start[lo];
    if (t.im2col) {
        // patch-gather packing of a few-channel stem: tiny, one element per thread
        const float* __restrict__ w = t.w;
        const int taps = t.r * t.s;
        __nv_bfloat16* wf = static_cast<__nv_bfloat16*>(t.w_fprop);
        const long long total = static_cast<long long>(t.k_rows) * t.c_pad;
        for (long long i = static_cast<long long>(local) * 2048 + threadIdx.x; i < static_cast<long long>(local + 1) * 2048 && i < total;
             i += 256) {
            const int j = static_cast<int>(i % t.c_pad), kk = static_cast<int>(i / t.c_pad);
            float v = 0.f;
            if (kk < t.k && j < taps * t.c) v = __ldg(w + (static_cast<size_t>(kk) * t.c + j % t.c) * taps + j / t.c);
            wf[i] = __float2bfloat16(v);
        }
        return;
    }
    const int taps = t.r * t.s;
    if (taps == 9) pack_tile<9>(t, local, tile);
    else if (taps == 1) pack_tile<1>(t, local, tile);
    else pack_tile<0>(t, local, tile);
}

}  // namespace
}  // namespace b200seg

extern "C" {

int b200seg_sgd_multi(const b200seg_sgd_desc* descs, int n, float lr, float momentum, float dampening, float weight_decay,
                      int nesterov, int first_step, float grad_scale, const float* hyper, b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    B200SEG_CHECK(descs && n > 0, "sgd_multi: no tensors");
    B200SEG_CHECK(!(nesterov && (momentum <= 0.f || dampening != 0.f)),
                  "sgd_multi: Nesterov momentum requires a momentum and zero dampening");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    for (int base = 0; base < n;) {
        SgdChunk c;
        memset(&c, 0, sizeof(c));
        c.lr = lr; c.momentum = momentum; c.dampening = dampening; c.weight_decay = weight_decay;
        c.grad_scale = grad_scale; c.nesterov = nesterov; c.first = first_step;
        c.hyper = hyper;
        int blocks = 0;
        while (base < n && c.n < kSgdTensors) {
            const b200seg_sgd_desc& d = descs[base];
            B200SEG_CHECK(d.p && d.g && d.numel >= 0 && (momentum == 0.f || d.m), "sgd_multi: tensor %d has null pointers", base);
            ++base;
            if (d.numel == 0) continue;
            c.t[c.n] = d;
            c.blk_start[c.n] = blocks;
            blocks += static_cast<int>(ceil_div64(d.numel, kSgdElemsPerBlock));
            ++c.n;
        }
        c.blk_start[c.n] = blocks;
        if (c.n == 0) continue;
        B200SEG_CUDA(launch_pdl(sgd_multi_kernel, blocks, 256, 0, st, c));
    }
    return 0;
}

int b200seg_adam_multi(const b200seg_adam_desc* descs, int n, const float* hyper, int decoupled, int first_step,
                       b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    B200SEG_CHECK(descs && n > 0, "adam_multi: no tensors");
    B200SEG_CHECK(hyper, "adam_multi: the hyper-parameter vector lives in device memory (11 floats)");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    for (int base = 0; base < n;) {
        AdamChunk c;
        memset(&c, 0, sizeof(c));
        c.decoupled = decoupled;
        c.first = first_step;
        c.hyper = hyper;
        int blocks = 0;
        while (base < n && c.n < kAdamTensors) {
            const b200seg_adam_desc& d = descs[base];
            B200SEG_CHECK(d.p && d.g && d.m && d.v && d.numel >= 0, "adam_multi: tensor %d has null pointers", base);
            ++base;
            if (d.numel == 0) continue;
            c.t[c.n] = d;
            c.blk_start[c.n] = blocks;
            blocks += static_cast<int>(ceil_div64(d.numel, kSgdElemsPerBlock));
            ++c.n;
        }
        c.blk_start[c.n] = blocks;
        if (c.n == 0) continue;
        B200SEG_CUDA(launch_pdl(adam_multi_kernel, blocks, 256, 0, st, c));
    }
    return 0;
}

int b200seg_copy_multi(const b200seg_sgd_desc* descs, int n, b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    B200SEG_CHECK(descs && n > 0, "copy_multi: no tensors");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    for (int base = 0; base < n;) {
        CopyChunk c;
        memset(&c, 0, sizeof(c));
        int blocks = 0;
        while (base < n && c.n < kSgdTensors) {
            const b200seg_sgd_desc& d = descs[base];
            B200SEG_CHECK(d.p && d.g && d.numel >= 0, "copy_multi: tensor %d has null pointers", base);
            ++base;
            if (d.numel == 0) continue;
            c.t[c.n] = d;
            c.blk_start[c.n] = blocks;
            blocks += static_cast<int>(ceil_div64(d.numel, kSgdElemsPerBlock));
            ++c.n;
        }
        c.blk_start[c.n] = blocks;
        if (c.n == 0) continue;
        B200SEG_CUDA(launch_pdl(copy_multi_kernel, blocks, 256, 0, st, c));
    }
    return 0;
}

int b200seg_pack_weights_multi(const b200seg_pack_desc* descs, int n, b200seg_stream_t stream) {
    using namespace b200seg;
    clear_error();
    B200SEG_CHECK(descs && n > 0, "pack_weights_multi: no tensors");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    for (int base = 0; base < n;) {
        PackChunk c;
        memset(&c, 0, sizeof(c));
        int blocks = 0;
        while (base < n && c.n < kPackTensors) {
            const b200seg_pack_desc& d = descs[base];
            B200SEG_CHECK(d.w && d.w_fprop && d.k > 0 && d.c > 0 && d.r > 0 && d.s > 0 && d.k_rows >= d.k,
                          "pack_weights_multi: tensor %d: bad descriptor", base);
            int nblk;
            if (d.im2col) {
                B200SEG_CHECK(d.c_pad % 8 == 0 && d.c_pad >= d.r * d.s * d.c, "pack_weights_multi: tensor %d: bad c_pad", base);
                nblk = static_cast<int>(ceil_div64(static_cast<long long>(d.k_rows) * d.c_pad, 2048));
            } else {
                B200SEG_CHECK(d.r * d.s <= kPackMaxTaps, "pack_weights_multi: tensor %d: more than %d taps", base, kPackMaxTaps);
                B200SEG_CHECK(!d.w_dgrad || (d.c_rows >= d.c && d.k_cols >= d.k), "pack_weights_multi: tensor %d: bad dgrad dims",
                              base);
                const int krows = d.w_dgrad && d.k_cols > d.k_rows ? d.k_cols : d.k_rows;
                const int crows = d.w_dgrad && d.c_rows > d.c ? d.c_rows : d.c;
                b200seg_pack_desc dd = d;
                if (!d.w_dgrad) {
                    dd.c_rows = d.c;
                    dd.k_cols = d.k_rows;
                }
                c.t[c.n] = dd;
                nblk = ceil_div(krows, pack_tile_k(d.r * d.s)) * ceil_div(crows, kPackTile);
                c.blk_start[c.n] = blocks;
                blocks += nblk;
                ++c.n;
                ++base;
                continue;
            }
            c.t[c.n] = d;
            c.blk_start[c.n] = blocks;
            blocks += nblk;
            ++c.n;
            ++base;
        }
        c.blk_start[c.n] = blocks;
        B200SEG_CUDA(launch_pdl(pack_multi_kernel, blocks, 256, 0, st, c));
    }
    return 0;
}

}  // extern "C"
