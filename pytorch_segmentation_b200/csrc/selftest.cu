// Standalone GPU self-test of the tcgen05 conv kernels against a straightforward host loop.
// Development tool (not part of the library): `make selftest && build/selftest`.  Exits non-zero on mismatch.
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "../../include/b200seg.h"

#define CK(x)                                                                         \
    do {                                                                              \
        cudaError_t e = (x);                                                          \
        if (e != cudaSuccess) {                                                       \
            printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); \
            exit(2);                                                                  \
        }                                                                             \
    } while (0)
#define AB(x)                                                          \
    do {                                                               \
        if ((x) != 0) {                                                \
            printf("b200seg error: %s (%s:%d)\n", b200seg_last_error(), __FILE__, __LINE__); \
            exit(3);                                                   \
        }                                                              \
    } while (0)

static uint32_t rng_state = 12345;
static float frand() {
    rng_state = rng_state * 1664525u + 1013904223u;
    return ((rng_state >> 8) & 0xFFFF) / 65536.0f - 0.5f;
}
static float bf16_round(float v) { return __bfloat162float(__float2bfloat16(v)); }

struct ConvCase {
    int n, h, w, c, k, r, dil, c_total, c_off, y_total, y_off, fp32_out, relu, use_scale;
};

static int run_fprop(const ConvCase& cs) {
    const int taps = cs.r * cs.r;
    const size_t npix = (size_t)cs.n * cs.h * cs.w;
    std::vector<float> x(npix * cs.c_total), w((size_t)cs.k * cs.c * taps), scale(cs.k), bias(cs.k);
    for (auto& v : x) v = bf16_round(frand());
    for (auto& v : w) v = bf16_round(frand() * 0.2f);
    for (auto& v : scale) v = 1.f + frand();
    for (auto& v : bias) v = frand();
    std::vector<__nv_bfloat16> xb(x.size());
    for (size_t i = 0; i < x.size(); ++i) xb[i] = __float2bfloat16(x[i]);
    const int k_rows = (cs.k + 31) / 32 * 32;
    float* d_w;
    __nv_bfloat16 *d_x, *d_wf;
    void* d_y;
    float *d_scale, *d_bias;
    CK(cudaMalloc(&d_x, xb.size() * 2));
    CK(cudaMalloc(&d_w, w.size() * 4));
    CK(cudaMalloc(&d_wf, (size_t)taps * k_rows * cs.c * 2));
    const size_t ybytes = npix * cs.y_total * (cs.fp32_out ? 4 : 2);
    CK(cudaMalloc(&d_y, ybytes));
    CK(cudaMemset(d_y, 0, ybytes));
    CK(cudaMalloc(&d_scale, cs.k * 4));
    CK(cudaMalloc(&d_bias, cs.k * 4));
    CK(cudaMemcpy(d_x, xb.data(), xb.size() * 2, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d_w, w.data(), w.size() * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d_scale, scale.data(), cs.k * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d_bias, bias.data(), cs.k * 4, cudaMemcpyHostToDevice));
    AB(b200seg_pack_weights(d_w, cs.k, cs.c, cs.r, cs.r, d_wf, k_rows, nullptr, 0, 0, 0));
    b200seg_conv_src src = {d_x, cs.c_total, cs.c_off, cs.c, d_wf, k_rows, cs.r, cs.r, cs.dil};
    b200seg_conv_out out = {d_y, cs.y_total, cs.y_off, cs.k, cs.fp32_out, cs.use_scale ? d_scale : nullptr, d_bias, cs.relu, nullptr};
    AB(b200seg_conv_fprop(cs.n, cs.h, cs.w, 1, &src, &out, 0));
    CK(cudaDeviceSynchronize());
    std::vector<uint8_t> yraw(ybytes);
    CK(cudaMemcpy(yraw.data(), d_y, ybytes, cudaMemcpyDeviceToHost));
    // host reference on a pixel subsample (all pixels if small)
    double max_err = 0, max_ref = 0;
    const size_t step = npix > 4096 ? npix / 1531 : 1;
    for (size_t p = 0; p < npix; p += step) {
        const int n = p / ((size_t)cs.h * cs.w), yy = (p / cs.w) % cs.h, xx = p % cs.w;
        for (int k = 0; k < cs.k; ++k) {
            double acc = 0;
            for (int r = 0; r < cs.r; ++r)
                for (int s = 0; s < cs.r; ++s) {
                    const int iy = yy + (r - cs.r / 2) * cs.dil, ix = xx + (s - cs.r / 2) * cs.dil;
                    if (iy < 0 || iy >= cs.h || ix < 0 || ix >= cs.w) continue;
                    const float* xp = &x[(((size_t)n * cs.h + iy) * cs.w + ix) * cs.c_total + cs.c_off];
                    for (int c = 0; c < cs.c; ++c) acc += (double)xp[c] * w[((size_t)k * cs.c + c) * taps + r * cs.r + s];
                }
            if (cs.use_scale) acc *= scale[k];
            acc += bias[k];
            if (cs.relu && acc < 0) acc = 0;
            const size_t off = p * cs.y_total + cs.y_off + k;
            const double got = cs.fp32_out ? reinterpret_cast<float*>(yraw.data())[off]
                                           : __bfloat162float(reinterpret_cast<__nv_bfloat16*>(yraw.data())[off]);
            max_err = fmax(max_err, fabs(got - acc));
            max_ref = fmax(max_ref, fabs(acc));
        }
    }
    // untouched channels of the output buffer must still be zero
    int dirty = 0;
    for (size_t p = 0; p < npix && !dirty; p += step)
        for (int ch = 0; ch < cs.y_total; ++ch) {
            if (ch >= cs.y_off && ch < cs.y_off + cs.k) continue;
            const size_t off = p * cs.y_total + ch;
            const float got = cs.fp32_out ? reinterpret_cast<float*>(yraw.data())[off]
                                          : __bfloat162float(reinterpret_cast<__nv_bfloat16*>(yraw.data())[off]);
            if (got != 0.f) dirty = 1;
        }
    const double rel = max_err / fmax(max_ref, 1e-20);
    const int ok = rel < (cs.fp32_out ? 2e-3 : 8e-3) && !dirty;
    printf("fprop n%d %dx%d c%d(+%d/%d) k%d r%d dil%d fp32=%d relu=%d : max_err %.4g max_ref %.4g rel %.3g dirty %d %s\n",
           cs.n, cs.h, cs.w, cs.c, cs.c_off, cs.c_total, cs.k, cs.r, cs.dil, cs.fp32_out, cs.relu, max_err, max_ref, rel,
           dirty, ok ? "OK" : "FAIL");
    cudaFree(d_x); cudaFree(d_w); cudaFree(d_wf); cudaFree(d_y); cudaFree(d_scale); cudaFree(d_bias);
    return ok ? 0 : 1;
}

// dgrad through the same kernel with the mirrored/transposed packing, two sources summed
static int run_dgrad(int n, int h, int w, int c, int k, int dil) {
    const int taps = 9;
    const size_t npix = (size_t)n * h * w;
    std::vector<float> dy1(npix * k), dy2(npix * k), w1((size_t)k * c * taps), w2((size_t)k * c);
    for (auto& v : dy1) v = bf16_round(frand());
    for (auto& v : dy2) v = bf16_round(frand());
    for (auto& v : w1) v = bf16_round(frand() * 0.2f);
    for (auto& v : w2) v = bf16_round(frand() * 0.2f);
    std::vector<__nv_bfloat16> b1(dy1.size()), b2(dy2.size());
    for (size_t i = 0; i < dy1.size(); ++i) b1[i] = __float2bfloat16(dy1[i]), b2[i] = __float2bfloat16(dy2[i]);
    const int c_rows = (c + 31) / 32 * 32;
    __nv_bfloat16 *d1, *d2, *wd1, *wd2, *dx;
    float *dw1, *dw2;
    CK(cudaMalloc(&d1, b1.size() * 2)); CK(cudaMalloc(&d2, b2.size() * 2));
    CK(cudaMalloc(&dw1, w1.size() * 4)); CK(cudaMalloc(&dw2, w2.size() * 4));
    CK(cudaMalloc(&wd1, (size_t)taps * c_rows * k * 2)); CK(cudaMalloc(&wd2, (size_t)c_rows * k * 2));
    CK(cudaMalloc(&dx, npix * c * 2));
    CK(cudaMemcpy(d1, b1.data(), b1.size() * 2, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d2, b2.data(), b2.size() * 2, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dw1, w1.data(), w1.size() * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dw2, w2.data(), w2.size() * 4, cudaMemcpyHostToDevice));
    AB(b200seg_pack_weights(dw1, k, c, 3, 3, nullptr, 0, wd1, c_rows, k, 0));
    AB(b200seg_pack_weights(dw2, k, c, 1, 1, nullptr, 0, wd2, c_rows, k, 0));
    b200seg_conv_src src[2] = {{d1, k, 0, k, wd1, c_rows, 3, 3, dil}, {d2, k, 0, k, wd2, c_rows, 1, 1, 1}};
    b200seg_conv_out out = {dx, c, 0, c, 0, nullptr, nullptr, 0, nullptr};
    AB(b200seg_conv_dgrad(n, h, w, 2, src, &out, 0));
    CK(cudaDeviceSynchronize());
    std::vector<__nv_bfloat16> got(npix * c);
    CK(cudaMemcpy(got.data(), dx, got.size() * 2, cudaMemcpyDeviceToHost));
    double max_err = 0, max_ref = 0;
    const size_t step = npix > 2048 ? npix / 701 : 1;
    for (size_t p = 0; p < npix; p += step) {
        const int nn = p / ((size_t)h * w), yy = (p / w) % h, xx = p % w;
        for (int cc = 0; cc < c; ++cc) {
            double acc = 0;
            for (int r = 0; r < 3; ++r)
                for (int s = 0; s < 3; ++s) {
                    const int oy = yy - (r - 1) * dil, ox = xx - (s - 1) * dil;
                    if (oy < 0 || oy >= h || ox < 0 || ox >= w) continue;
                    const float* g = &dy1[(((size_t)nn * h + oy) * w + ox) * k];
                    for (int kk = 0; kk < k; ++kk) acc += (double)g[kk] * w1[((size_t)kk * c + cc) * 9 + r * 3 + s];
                }
            const float* g2 = &dy2[p * k];
            for (int kk = 0; kk < k; ++kk) acc += (double)g2[kk] * w2[(size_t)kk * c + cc];
            max_err = fmax(max_err, fabs(__bfloat162float(got[p * c + cc]) - acc));
            max_ref = fmax(max_ref, fabs(acc));
        }
    }
    const double rel = max_err / fmax(max_ref, 1e-20);
    const int ok = rel < 8e-3;
    printf("dgrad(2 src) n%d %dx%d c%d k%d dil%d : max_err %.4g max_ref %.4g rel %.3g %s\n", n, h, w, c, k, dil, max_err,
           max_ref, rel, ok ? "OK" : "FAIL");
    cudaFree(d1); cudaFree(d2); cudaFree(dw1); cudaFree(dw2); cudaFree(wd1); cudaFree(wd2); cudaFree(dx);
    return ok ? 0 : 1;
}

static int run_wgrad(int n, int h, int w, int c, int k, int k_out, int r, int dil) {
    const int taps = r * r;
    const size_t npix = (size_t)n * h * w;
    std::vector<float> x(npix * c), dy(npix * k);
    for (auto& v : x) v = bf16_round(frand());
    for (auto& v : dy) v = bf16_round(frand());
    std::vector<__nv_bfloat16> xb(x.size()), yb(dy.size());
    for (size_t i = 0; i < x.size(); ++i) xb[i] = __float2bfloat16(x[i]);
    for (size_t i = 0; i < dy.size(); ++i) yb[i] = __float2bfloat16(dy[i]);
    __nv_bfloat16 *d_x, *d_dy;
    float* d_dw;
    CK(cudaMalloc(&d_x, xb.size() * 2)); CK(cudaMalloc(&d_dy, yb.size() * 2));
    CK(cudaMalloc(&d_dw, (size_t)k_out * c * taps * 4));
    CK(cudaMemcpy(d_x, xb.data(), xb.size() * 2, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d_dy, yb.data(), yb.size() * 2, cudaMemcpyHostToDevice));
    b200seg_wgrad_desc d;
    memset(&d, 0, sizeof(d));
    d.n = n; d.h = h; d.w = w;
    d.dy = d_dy; d.dy_c_total = k; d.dy_c_offset = 0; d.k = k;
    d.x = d_x; d.x_c_total = c; d.x_c_offset = 0; d.c = c;
    d.r = r; d.s = r; d.dil = dil; d.dw = d_dw; d.k_out = k_out; d.accumulate = 0;
    d.workspace_bytes = b200seg_conv_wgrad_workspace(&d);
    if (!d.workspace_bytes) { printf("wgrad workspace query failed: %s\n", b200seg_last_error()); return 1; }
    CK(cudaMalloc(&d.workspace, d.workspace_bytes));
    AB(b200seg_conv_wgrad(&d, 0));
    CK(cudaDeviceSynchronize());
    std::vector<float> got((size_t)k_out * c * taps);
    CK(cudaMemcpy(got.data(), d_dw, got.size() * 4, cudaMemcpyDeviceToHost));
    double max_err = 0, max_ref = 0;
    const int kstep = k_out > 16 ? k_out / 13 : 1, cstep = c > 16 ? c / 11 : 1;
    for (int kk = 0; kk < k_out; kk += kstep)
        for (int cc = 0; cc < c; cc += cstep)
            for (int t = 0; t < taps; ++t) {
                const int dyo = (t / r - r / 2) * dil, dxo = (t % r - r / 2) * dil;
                double acc = 0;
                for (int nn = 0; nn < n; ++nn)
                    for (int yy = 0; yy < h; ++yy)
                        for (int xx = 0; xx < w; ++xx) {
                            const int iy = yy + dyo, ix = xx + dxo;
                            if (iy < 0 || iy >= h || ix < 0 || ix >= w) continue;
                            acc += (double)dy[(((size_t)nn * h + yy) * w + xx) * k + kk] *
                                   x[(((size_t)nn * h + iy) * w + ix) * c + cc];
                        }
                max_err = fmax(max_err, fabs(got[((size_t)kk * c + cc) * taps + t] - acc));
                max_ref = fmax(max_ref, fabs(acc));
            }
    const double rel = max_err / fmax(max_ref, 1e-20);
    const int ok = rel < 2e-3;
    printf("wgrad n%d %dx%d c%d k%d(k_out %d) r%d dil%d ws %zu : max_err %.4g max_ref %.4g rel %.3g %s\n", n, h, w, c, k,
           k_out, r, dil, d.workspace_bytes, max_err, max_ref, rel, ok ? "OK" : "FAIL");
    cudaFree(d_x); cudaFree(d_dy); cudaFree(d_dw); cudaFree(d.workspace);
    return ok ? 0 : 1;
}

int main(int argc, char** argv) {
    int sms = 0, maj = 0, min = 0;
    AB(b200seg_device_info(&sms, &maj, &min));
    printf("device: %d SMs, sm_%d%d\n", sms, maj, min);
    int fails = 0;
    const ConvCase cases[] = {
        // n  h   w   c    k   r dil ctot coff ytot yoff fp32 relu scale
        {1, 8, 16, 64, 32, 1, 1, 64, 0, 32, 0, 1, 0, 0},           // smallest: one tile, one k-iteration, fp32 out
        {2, 32, 32, 128, 256, 3, 6, 128, 0, 256, 0, 0, 1, 1},      // ASPP-like atrous
        {2, 32, 32, 256, 256, 3, 18, 256, 0, 1280, 512, 0, 1, 1},  // rate 18 into a concat slice
        {3, 10, 10, 160, 64, 3, 1, 160, 0, 64, 0, 0, 1, 1},        // UNet-like ragged map
        {2, 20, 20, 352, 128, 3, 1, 352, 0, 128, 0, 0, 1, 1},      // C not a multiple of 64
        {1, 64, 64, 384, 32, 3, 1, 384, 0, 32, 0, 1, 0, 0},        // cls_conv-like narrow N, fp32, bias only
        {16, 1, 1, 256, 256, 1, 1, 256, 0, 256, 0, 0, 1, 1},       // pooled branch: batch in the tile
        {2, 16, 16, 64, 128, 1, 1, 192, 64, 384, 256, 0, 1, 1},    // 1x1 on an input slice, into an output slice
        {1, 40, 40, 128, 352, 3, 2, 128, 0, 352, 0, 0, 0, 0},      // N tiled 2 x 192
    };
    for (const auto& c : cases) fails += run_fprop(c);
    fails += run_dgrad(2, 32, 32, 128, 64, 6);
    fails += run_dgrad(2, 16, 24, 88, 32, 1);
    fails += run_wgrad(2, 32, 32, 128, 256, 256, 3, 6);
    fails += run_wgrad(2, 32, 32, 256, 64, 64, 1, 1);
    fails += run_wgrad(1, 64, 64, 384, 32, 21, 3, 1);
    fails += run_wgrad(3, 10, 10, 160, 64, 64, 3, 1);
    fails += run_wgrad(2, 32, 32, 64, 64, 64, 3, 18);
    printf(fails ? "SELFTEST FAILED (%d)\n" : "SELFTEST PASSED\n", fails);
    return fails ? 1 : 0;
}
