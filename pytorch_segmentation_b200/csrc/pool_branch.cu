// The ASPP image-pooling branch after the global average pool (models/aspp.py:11-12): a 1x1 ConvNormAct on an
// N x C x 1 x 1 map, i.e. a [N x C] * [C x K] product with N = batch (16) rows followed by BatchNorm over those
// N samples.  16.8 MFLOP against a 2 MB fp32 weight: bandwidth/latency bound, and a 128-row tcgen05 tile would be
// 87 % padding — so it runs on CUDA cores, in fp32 straight from the fp32 master weight (no bf16 packing).
// fp32 also matters numerically: BatchNorm over N samples divides by the *between-image* spread of the pooled
// features, which is small against their mean, so bf16 rounding of the pooled vector is amplified ~10x
// (DESIGN.md "precision").
#include "common.cuh"
#include "pdl.cuh"
#include "vec.cuh"

namespace b200seg {
namespace {

constexpr int kRowChunk = 16;

// y[n][k] = sum_c x[n][c] * w[k][c];  one CTA (8 warps) per output channel k, threads stride over c — a warp per
// channel walked c = 2048 in 64 dependent-latency steps; here it is 8, and the partial sums meet in shared memory
__global__ void __launch_bounds__(256) linear_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                                          float* __restrict__ y, int n, int c, int k) {
    pdl_prologue();
    __shared__ float red[8][kRowChunk];
    const int kk = blockIdx.x, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const float* wr = w + static_cast<size_t>(kk) * c;
    for (int n0 = 0; n0 < n; n0 += kRowChunk) {
        float acc[kRowChunk];
#pragma unroll
        for (int i = 0; i < kRowChunk; ++i) acc[i] = 0.f;
        for (int cc = threadIdx.x; cc < c; cc += 256) {
            const float wv = __ldg(wr + cc);
#pragma unroll
            for (int i = 0; i < kRowChunk; ++i)
                if (n0 + i < n) acc[i] = fmaf(__ldg(x + static_cast<size_t>(n0 + i) * c + cc), wv, acc[i]);
        }
#pragma unroll
        for (int i = 0; i < kRowChunk; ++i) {
            const float s = warp_sum(acc[i]);
            if (lane == 0) red[warp][i] = s;
        }
        __syncthreads();
        if (threadIdx.x < kRowChunk && n0 + threadIdx.x < n) {
            float s = 0.f;
#pragma unroll
            for (int r = 0; r < 8; ++r) s += red[r][threadIdx.x];
            y[static_cast<size_t>(n0 + threadIdx.x) * k + kk] = s;
        }
        __syncthreads();
    }
}

// dx[n][c] = sum_k dy[n][k] * w[k][c];  one thread per input channel c (coalesced over c for every k); the k range is
// split over blockIdx.y (16 CTAs of 128 threads could not pull a 2 MB weight at speed) and partial sums meet in dx
// (zeroed by the caller) through atomics.
constexpr int kBwdDataKChunk = 16;
__global__ void __launch_bounds__(128) linear_bwd_data_kernel(const float* __restrict__ dy, const float* __restrict__ w,
                                                               float* __restrict__ dx, int n, int c, int k) {
    pdl_prologue();
    const int cc = blockIdx.x * blockDim.x + threadIdx.x;
    if (cc >= c) return;
    const int k_begin = blockIdx.y * kBwdDataKChunk;
    const int k_end = min(k, k_begin + kBwdDataKChunk);
    for (int n0 = 0; n0 < n; n0 += kRowChunk) {
        float acc[kRowChunk];
#pragma unroll
        for (int i = 0; i < kRowChunk; ++i) acc[i] = 0.f;
        for (int kk = k_begin; kk < k_end; ++kk) {
            const float wv = __ldg(w + static_cast<size_t>(kk) * c + cc);
#pragma unroll
            for (int i = 0; i < kRowChunk; ++i)
                if (n0 + i < n) acc[i] = fmaf(__ldg(dy + static_cast<size_t>(n0 + i) * k + kk), wv, acc[i]);
        }
#pragma unroll
        for (int i = 0; i < kRowChunk; ++i)
            if (n0 + i < n) atomicAdd(dx + static_cast<size_t>(n0 + i) * c + cc, acc[i]);
    }
}

// dw[k][c] (+)= sum_n dy[n][k] * x[n][c]
__global__ void __launch_bounds__(256) linear_bwd_weight_kernel(const float* __restrict__ dy, const float* __restrict__ x,
                                                                 float* __restrict__ dw, int n, int c, int k,
                                                                 int accumulate) {
    pdl_prologue();
    const size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x;
    if (i >= static_cast<size_t>(k) * c) return;
    const int kk = static_cast<int>(i / c), cc = static_cast<int>(i % c);
    float acc = 0.f;
    for (int r = 0; r < n; ++r) acc = fmaf(__ldg(dy + static_cast<size_t>(r) * k + kk), __ldg(x + static_cast<size_t>(r) * c + cc), acc);
    dw[i] = accumulate ? dw[i] + acc : acc;
}

// BatchNorm (+ReLU) over the n rows of x[n][k]; one thread per channel, two-pass statistics.
__global__ void bn_rows_fwd_kernel(const float* __restrict__ x, int n, int k, const float* __restrict__ gamma,
                                   const float* __restrict__ beta, float eps, float momentum, float* running_mean,
                                   float* running_var, int training, int relu, float* __restrict__ y,
                                   float* __restrict__ save_mean, float* __restrict__ save_invstd) {
    pdl_prologue();
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= k) return;
    float mean, invstd;
    if (training) {
        float s = 0.f;
        for (int r = 0; r < n; ++r) s += x[static_cast<size_t>(r) * k + j];
        mean = s / n;
        float v = 0.f;
        for (int r = 0; r < n; ++r) {
            const float d = x[static_cast<size_t>(r) * k + j] - mean;
            v = fmaf(d, d, v);
        }
        const float var = v / n;
        invstd = rsqrtf(var + eps);
        if (running_mean) running_mean[j] = fmaf(momentum, mean - running_mean[j], running_mean[j]);
        if (running_var) {
            const float unb = n > 1 ? v / (n - 1) : var;
            running_var[j] = fmaf(momentum, unb - running_var[j], running_var[j]);
        }
    } else {
        mean = running_mean[j];
        invstd = rsqrtf(running_var[j] + eps);
    }
    save_mean[j] = mean;
    save_invstd[j] = invstd;
    const float g = gamma ? gamma[j] : 1.f, b = beta ? beta[j] : 0.f;
    for (int r = 0; r < n; ++r) {
        float v = fmaf((x[static_cast<size_t>(r) * k + j] - mean) * invstd, g, b);
        if (relu) v = fmaxf(v, 0.f);
        y[static_cast<size_t>(r) * k + j] = v;
    }
}

__global__ void bn_rows_bwd_kernel(const float* __restrict__ dy, const float* __restrict__ x, int n, int k,
                                   const float* __restrict__ gamma, const float* __restrict__ beta,
                                   const float* __restrict__ save_mean, const float* __restrict__ save_invstd, int training,
                                   int relu, float* __restrict__ dx, float* __restrict__ dgamma,
                                   float* __restrict__ dbeta) {
    pdl_prologue();
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= k) return;
    const float mean = save_mean[j], invstd = save_invstd[j];
    const float g = gamma ? gamma[j] : 1.f, b = beta ? beta[j] : 0.f;
    float s0 = 0.f, s1 = 0.f;
    for (int r = 0; r < n; ++r) {
        const float xh = (x[static_cast<size_t>(r) * k + j] - mean) * invstd;
        float dz = dy[static_cast<size_t>(r) * k + j];
        if (relu && fmaf(xh, g, b) <= 0.f) dz = 0.f;
        s0 += dz;
        s1 = fmaf(dz, xh, s1);
    }
    dgamma[j] = s1;
    dbeta[j] = s0;
    const float inv_n = 1.f / n;
    for (int r = 0; r < n; ++r) {
        const float xh = (x[static_cast<size_t>(r) * k + j] - mean) * invstd;
        float dz = dy[static_cast<size_t>(r) * k + j];
        if (relu && fmaf(xh, g, b) <= 0.f) dz = 0.f;
        const float v = training ? dz - s0 * inv_n - xh * s1 * inv_n : dz;
        dx[static_cast<size_t>(r) * k + j] = v * g * invstd;
    }
}

}  // namespace
}  // namespace b200seg

extern "C" {

int b200seg_linear_f32_fwd(const float* x, const float* w, float* y, int n, int c, int k, b200seg_stream_t st) {
    using namespace b200seg;
    clear_error();
    B200SEG_CHECK(x && w && y && n > 0 && c > 0 && k > 0, "linear_f32_fwd: bad arguments");
    B200SEG_CUDA(launch_pdl(linear_fwd_kernel, k, 256, 0, static_cast<cudaStream_t>(st), x, w, y, n, c, k));
    return 0;
}

int b200seg_linear_f32_bwd_data(const float* dy, const float* w, float* dx, int n, int c, int k, b200seg_stream_t st) {
    using namespace b200seg;
    clear_error();
    B200SEG_CHECK(dy && w && dx && n > 0 && c > 0 && k > 0, "linear_f32_bwd_data: bad arguments");
    B200SEG_CUDA(cudaMemsetAsync(dx, 0, sizeof(float) * static_cast<size_t>(n) * c, static_cast<cudaStream_t>(st)));
    B200SEG_CUDA(launch_pdl(linear_bwd_data_kernel, dim3(ceil_div(c, 128), ceil_div(k, kBwdDataKChunk)), 128, 0,
                            static_cast<cudaStream_t>(st), dy, w, dx, n, c, k));
    return 0;
}

int b200seg_linear_f32_bwd_weight(const float* dy, const float* x, float* dw, int n, int c, int k, int accumulate,
                                  b200seg_stream_t st) {
    using namespace b200seg;
    clear_error();
    B200SEG_CHECK(dy && x && dw && n > 0 && c > 0 && k > 0, "linear_f32_bwd_weight: bad arguments");
    const size_t total = static_cast<size_t>(k) * c;
    B200SEG_CUDA(launch_pdl(linear_bwd_weight_kernel, static_cast<unsigned>((total + 255) / 256), 256, 0, static_cast<cudaStream_t>(st), 
        dy, x, dw, n, c, k, accumulate));
    return 0;
}

int b200seg_bn_rows_fwd(const float* x, int n, int k, const float* gamma, const float* beta, float eps, float momentum,
                        float* running_mean, float* running_var, int training, int relu, float* y, float* save_mean,
                        float* save_invstd, b200seg_stream_t st) {
    using namespace b200seg;
    clear_error();
    B200SEG_CHECK(x && y && save_mean && save_invstd && n > 0 && k > 0, "bn_rows_fwd: bad arguments");
    B200SEG_CHECK(training || (running_mean && running_var), "bn_rows_fwd: eval mode needs running statistics");
    B200SEG_CUDA(launch_pdl(bn_rows_fwd_kernel, ceil_div(k, 128), 128, 0, static_cast<cudaStream_t>(st), 
        x, n, k, gamma, beta, eps, momentum, running_mean, running_var, training, relu, y, save_mean, save_invstd));
    return 0;
}

int b200seg_bn_rows_bwd(const float* dy, const float* x, int n, int k, const float* gamma, const float* beta,
                        const float* save_mean, const float* save_invstd, int training, int relu, float* dx, float* dgamma,
                        float* dbeta, b200seg_stream_t st) {
    using namespace b200seg;
    clear_error();
    B200SEG_CHECK(dy && x && save_mean && save_invstd && dx && dgamma && dbeta && n > 0 && k > 0,
                  "bn_rows_bwd: bad arguments");
    B200SEG_CUDA(launch_pdl(bn_rows_bwd_kernel, ceil_div(k, 128), 128, 0, static_cast<cudaStream_t>(st), 
        dy, x, n, k, gamma, beta, save_mean, save_invstd, training, relu, dx, dgamma, dbeta));
    return 0;
}

}  // extern "C"
