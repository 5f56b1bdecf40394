// Error state, device queries and TMA descriptor encoding for libb200seg.
#include "common.cuh"

#include <cudaTypedefs.h>
#include <string.h>

#include <mutex>

namespace b200seg {

static thread_local char g_err[1024] = "";

int set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return 1;
}
void clear_error() { g_err[0] = 0; }

int sm_count() {
    static int cached[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 0;
    if (!cached[dev]) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return 0;
        cached[dev] = n;
    }
    return cached[dev];
}

// SMs the tensor-core kernels size their grids for.  They hold one CTA per SM (227 KB of shared memory) and stride
// their tiles statically, so a co-running kernel that takes a few SMs (NCCL's all-reduce under data parallelism)
// would push the last CTAs into a second wave and nearly double the launch; B200SEG_SM_RESERVE=<n> leaves n SMs free.
int gemm_sm_count() {
    static const int reserve = [] {
        const char* e = getenv("B200SEG_SM_RESERVE");
        const int v = e ? atoi(e) : 0;
        return v > 0 && v < 64 ? v : 0;
    }();
    const int n = sm_count();
    return n > reserve ? n - reserve : n;
}

// cuTensorMapEncodeTiled is a driver-API symbol; fetch it through the runtime so the library does not link
// against libcuda.so (absent on build machines without a driver).
static PFN_cuTensorMapEncodeTiled_v12000 encode_fn() {
    static PFN_cuTensorMapEncodeTiled_v12000 fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(p);
    });
    return fn;
}

int encode_tmap_bf16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                     const uint32_t* box) {
    auto fn = encode_fn();
    B200SEG_CHECK(fn != nullptr, "cuTensorMapEncodeTiled not available from the driver");
    B200SEG_CHECK((reinterpret_cast<uintptr_t>(base) & 15) == 0, "TMA base pointer %p is not 16-byte aligned", base);
    cuuint64_t gdim[5];
    cuuint64_t gstr[5];
    cuuint32_t bdim[5];
    cuuint32_t estr[5];
    for (int i = 0; i < rank; ++i) {
        gdim[i] = dims[i];
        bdim[i] = box[i];
        estr[i] = 1;
        if (i > 0) {
            gstr[i - 1] = strides_bytes[i - 1];
            B200SEG_CHECK((gstr[i - 1] & 15) == 0, "TMA stride %llu of dim %d is not a multiple of 16 bytes",
                          (unsigned long long)gstr[i - 1], i);
        }
        B200SEG_CHECK(box[i] >= 1 && box[i] <= 256, "TMA box dim %d = %u out of range", i, box[i]);
    }
    CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base), gdim, gstr, bdim,
                    estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    B200SEG_CHECK(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with CUresult %d (rank %d dims %llu,%llu,%llu,%llu)",
                  (int)r, rank, (unsigned long long)dims[0], (unsigned long long)(rank > 1 ? dims[1] : 0),
                  (unsigned long long)(rank > 2 ? dims[2] : 0), (unsigned long long)(rank > 3 ? dims[3] : 0));
    return 0;
}

int check_slice(const char* what, const void* p, int c_total, int c_off, int c) {
    B200SEG_CHECK(p != nullptr, "%s: null pointer", what);
    B200SEG_CHECK(c > 0 && c % 8 == 0 && c_off % 8 == 0 && c_total % 8 == 0 && c_off + c <= c_total,
                  "%s: bad channel slice (c=%d offset=%d total=%d; all must be multiples of 8)", what, c, c_off, c_total);
    B200SEG_CHECK((reinterpret_cast<uintptr_t>(p) & 15) == 0, "%s: pointer not 16-byte aligned", what);
    return 0;
}

int elementwise_grid(int64_t total) {
    int sms = sm_count();
    if (sms <= 0) sms = 148;
    int64_t blocks = ceil_div64(total, 256);
    const int64_t cap = static_cast<int64_t>(sms) * 16;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return static_cast<int>(blocks);
}

}  // namespace b200seg

extern "C" {

const char* b200seg_last_error(void) { return b200seg::g_err; }

int b200seg_version(void) { return 100; }

int b200seg_device_info(int* sm_count, int* cc_major, int* cc_minor) {
    int dev = 0;
    B200SEG_CUDA(cudaGetDevice(&dev));
    int major = 0, minor = 0, sms = 0;
    B200SEG_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
    B200SEG_CUDA(cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, dev));
    B200SEG_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    if (sm_count) *sm_count = sms;
    if (cc_major) *cc_major = major;
    if (cc_minor) *cc_minor = minor;
    B200SEG_CHECK(major == 10, "libb200seg is built for sm_100a only; device %d is sm_%d%d", dev, major, minor);
    return 0;
}

}  // extern "C"
