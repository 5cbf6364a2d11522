// Library-wide state of libpsb200_kernels.so: last-error text, launch counter, TMA
// tensor-map cache (cuTensorMapEncodeTiled is resolved at run time through
// cudaGetDriverEntryPoint so the library loads on a machine without libcuda).
#include <stdarg.h>
#include <string.h>

#include <atomic>
#include <mutex>
#include <unordered_map>

#include "../../include/psb200.h"
#include "host_common.h"

namespace psb {

static thread_local char g_err[512] = "";
static std::atomic<uint64_t> g_launches{0};

int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

void count_launch(uint64_t n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

int after_launch(const char* what) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(PSB_ERR_CUDA, "%s launch failed: %s", what, cudaGetErrorString(e));
    count_launch();
    return PSB_OK;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) ==
                cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    });
    return fn;
}

struct MapKey {
    uintptr_t base;
    uint64_t rows, cols, ld;
    uint32_t box_rows;
    bool operator==(const MapKey& o) const {
        return base == o.base && rows == o.rows && cols == o.cols && ld == o.ld &&
               box_rows == o.box_rows;
    }
};
struct MapKeyHash {
    size_t operator()(const MapKey& k) const {
        size_t h = k.base * 0x9E3779B97F4A7C15ull;
        h ^= (k.rows + 0x9E3779B97F4A7C15ull + (h << 6) + (h >> 2));
        h ^= (k.cols + 0x9E3779B97F4A7C15ull + (h << 6) + (h >> 2));
        h ^= (k.ld + 0x9E3779B97F4A7C15ull + (h << 6) + (h >> 2));
        h ^= (k.box_rows + 0x9E3779B97F4A7C15ull + (h << 6) + (h >> 2));
        return h;
    }
};

const CUtensorMap* tensor_map_bf16_2d(const void* base, uint64_t rows, uint64_t cols,
                                      uint64_t ld_elems, uint32_t box_rows) {
    static std::mutex mu;
    // node-based container: addresses of mapped values stay valid across inserts
    static std::unordered_map<MapKey, CUtensorMap, MapKeyHash> cache;
    MapKey key{reinterpret_cast<uintptr_t>(base), rows, cols, ld_elems, box_rows};
    std::lock_guard<std::mutex> lock(mu);
    auto it = cache.find(key);
    if (it != cache.end()) return &it->second;

    EncodeTiledFn enc = encode_fn();
    if (!enc) {
        fail(PSB_ERR_DRIVER, "cuTensorMapEncodeTiled is not available from the driver");
        return nullptr;
    }
    if ((reinterpret_cast<uintptr_t>(base) & 15) || (ld_elems % 8) || box_rows == 0 ||
        box_rows > 256 || rows == 0 || cols == 0) {
        fail(PSB_ERR_ARG,
             "tensor map: base %p must be 16-byte aligned, ld %llu a multiple of 8, "
             "box_rows in 1..256 (got %u), rows/cols non-zero",
             base, (unsigned long long)ld_elems, box_rows);
        return nullptr;
    }
    CUtensorMap m;
    cuuint64_t gdim[2] = {cols, rows};
    cuuint64_t gstride[1] = {ld_elems * 2};
    cuuint32_t box[2] = {64, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim, gstride,
                     box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        fail(PSB_ERR_DRIVER, "cuTensorMapEncodeTiled failed with CUresult %d (rows %llu cols %llu ld %llu)",
             (int)r, (unsigned long long)rows, (unsigned long long)cols, (unsigned long long)ld_elems);
        return nullptr;
    }
    auto ins = cache.emplace(key, m);
    return &ins.first->second;
}

}  // namespace psb

extern "C" {

const char* psb_last_error(void) { return psb::g_err; }
uint64_t psb_launch_count(void) { return psb::g_launches.load(std::memory_order_relaxed); }
int psb_version(void) { return 100; }

}  // extern "C"
