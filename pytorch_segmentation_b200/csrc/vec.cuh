// 128-bit bf16x8 helpers and small reductions shared by the bandwidth-bound kernels.
#pragma once
#include <cuda_bf16.h>
#include <stdint.h>

namespace b200seg {

// 8 bf16 packed in a uint4 <-> 8 floats
__device__ __forceinline__ void unpack8(const uint4& v, float (&f)[8]) {
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        f[2 * i] = __uint_as_float(w[i] << 16);
        f[2 * i + 1] = __uint_as_float(w[i] & 0xFFFF0000u);
    }
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
    uint4 o;
    __nv_bfloat162 t;
    t = __floats2bfloat162_rn(f[0], f[1]);
    o.x = *reinterpret_cast<uint32_t*>(&t);
    t = __floats2bfloat162_rn(f[2], f[3]);
    o.y = *reinterpret_cast<uint32_t*>(&t);
    t = __floats2bfloat162_rn(f[4], f[5]);
    o.z = *reinterpret_cast<uint32_t*>(&t);
    t = __floats2bfloat162_rn(f[6], f[7]);
    o.w = *reinterpret_cast<uint32_t*>(&t);
    return o;
}
__device__ __forceinline__ uint4 ldg128(const void* p) { return __ldg(reinterpret_cast<const uint4*>(p)); }
// streaming (evict-first) variants for data touched once
__device__ __forceinline__ uint4 ldcs128(const void* p) { return __ldcs(reinterpret_cast<const uint4*>(p)); }
__device__ __forceinline__ void st128(void* p, const uint4& v) { *reinterpret_cast<uint4*>(p) = v; }
__device__ __forceinline__ void load8f(const float* p, float (&f)[8]) {
    const float4 a = __ldg(reinterpret_cast<const float4*>(p));
    const float4 b = __ldg(reinterpret_cast<const float4*>(p) + 1);
    f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w;
    f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// PyTorch's bilinear source index (aten/native/UpSample.h area_pixel_compute_source_index, cubic = false).
__host__ __device__ __forceinline__ float bilinear_scale(int in, int out, int align_corners) {
    if (align_corners) return out > 1 ? static_cast<float>(in - 1) / static_cast<float>(out - 1) : 0.f;
    return static_cast<float>(in) / static_cast<float>(out);
}
// (group, x, y, n) of flat element i of an [n][H][W][groups] enumeration.  64-bit division is a ~100-instruction
// subroutine on this machine and three of them per element made the gather kernels ALU-bound; every shape the
// models produce fits 32 bits (uniform branch).
struct Idx4 {
    int g, x, y, n;
};
__device__ __forceinline__ Idx4 decode4(int64_t i, int groups, int W, int H, int64_t total) {
    Idx4 r;
    if (total < (int64_t(1) << 31)) {
        const unsigned u = static_cast<unsigned>(i);
        const unsigned p = u / static_cast<unsigned>(groups);
        r.g = static_cast<int>(u - p * groups);
        const unsigned q = p / static_cast<unsigned>(W);
        r.x = static_cast<int>(p - q * W);
        const unsigned n = q / static_cast<unsigned>(H);
        r.y = static_cast<int>(q - n * H);
        r.n = static_cast<int>(n);
    } else {
        r.g = static_cast<int>(i % groups);
        int64_t p = i / groups;
        r.x = static_cast<int>(p % W);
        p /= W;
        r.y = static_cast<int>(p % H);
        r.n = static_cast<int>(p / H);
    }
    return r;
}
// i = q * d + rem, with 32-bit arithmetic when `total` (an upper bound of i) allows
__device__ __forceinline__ int64_t divmod_small(int64_t i, int d, int64_t total, int& rem) {
    if (total < (int64_t(1) << 31)) {
        const unsigned u = static_cast<unsigned>(i), q = u / static_cast<unsigned>(d);
        rem = static_cast<int>(u - q * d);
        return q;
    }
    const int64_t q = i / d;
    rem = static_cast<int>(i - q * d);
    return q;
}
struct Lerp {
    int i0, i1;
    float w0, w1;
};
__device__ __forceinline__ Lerp bilinear_src(float scale, int dst, int in, int align_corners) {
    float src = align_corners ? scale * dst : fmaxf(scale * (dst + 0.5f) - 0.5f, 0.f);
    Lerp l;
    l.i0 = min(static_cast<int>(src), in - 1);
    l.i1 = l.i0 + (l.i0 < in - 1 ? 1 : 0);
    l.w1 = src - l.i0;
    l.w0 = 1.f - l.w1;
    return l;
}
// Inclusive range of destination indices that can read source index `i` (superset; callers re-check
// with bilinear_src).
__device__ __forceinline__ void bilinear_dst_range(float scale, int i, int in, int out, int align_corners, int& lo,
                                                   int& hi) {
    if (scale <= 0.f) {
        lo = 0;
        hi = out - 1;
        return;
    }
    const float inv = 1.f / scale;
    const float off = align_corners ? 0.f : 0.5f;
    // src = scale*(dst+off)-off in (i-1, i+1)  =>  dst in ((i-1+off)/scale-off, (i+1+off)/scale-off)
    lo = static_cast<int>(floorf((i - 1 + off) * inv - off)) - 1;
    hi = static_cast<int>(ceilf((i + 1 + off) * inv - off)) + 1;
    if (!align_corners && i == 0) lo = 0;   // clamped negative sources all land on 0
    lo = max(lo, 0);
    hi = min(hi, out - 1);
}

}  // namespace b200seg
