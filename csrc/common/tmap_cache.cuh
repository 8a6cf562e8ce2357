// Host-side cache of TMA tensor maps.  A training step issues ~2,000 GEMM / attention calls whose operands are the
// same few hundred (pointer, shape, box) combinations step after step; encoding a CUtensorMap for each call again
// was ~2,000 driver calls per step for descriptors that never change.  The descriptor depends
// only on the key below, so a cached copy stays valid even when the memory behind the pointer is freed and handed
// out again with the same shape.
#pragma once
#include <cuda.h>
#include <cudaTypedefs.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <mutex>
#include <unordered_map>

namespace ftl {

struct TmapKey {
    const void* ptr;
    int64_t rows, cols, ld;
    int box_cols, box_rows, device;
    bool operator==(const TmapKey& o) const {
        return ptr == o.ptr && rows == o.rows && cols == o.cols && ld == o.ld && box_cols == o.box_cols &&
               box_rows == o.box_rows && device == o.device;
    }
};
struct TmapKeyHash {
    size_t operator()(const TmapKey& k) const {
        uint64_t h = reinterpret_cast<uintptr_t>(k.ptr) * 0x9E3779B97F4A7C15ull;
        h ^= (uint64_t)k.rows * 0xC2B2AE3D27D4EB4Full + ((uint64_t)k.cols << 17) + ((uint64_t)k.ld << 3);
        h ^= ((uint64_t)k.box_cols << 40) ^ ((uint64_t)k.box_rows << 48) ^ ((uint64_t)k.device << 56);
        return (size_t)(h ^ (h >> 29));
    }
};

class TmapCache {
  public:
    // Row-major bf16 matrix [rows, cols] with leading dimension ld (elements); box = {box_cols (contiguous), box_rows},
    // 128-byte swizzle.  Returns 0 or a cudaError_t value.
    int get(CUtensorMap* out, const void* ptr, int64_t rows, int64_t cols, int64_t ld, int box_cols, int box_rows) {
        std::call_once(once_, [this] { load(); });
        if (!encode_) return (int)cudaErrorNotSupported;
        int dev = 0;
        (void)cudaGetDevice(&dev);
        const TmapKey key{ptr, rows, cols, ld, box_cols, box_rows, dev};
        {
            std::lock_guard<std::mutex> lk(mu_);
            auto it = map_.find(key);
            if (it != map_.end()) {
                *out = it->second;
                return 0;
            }
        }
        cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
        cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
        cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
        cuuint32_t estr[2] = {1, 1};
        CUresult r = encode_(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                             CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                             CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) return (int)cudaErrorInvalidValue;
        std::lock_guard<std::mutex> lk(mu_);
        if (map_.size() >= kMaxEntries) map_.clear();   // shapes keep changing (sweeps, tests): start over
        map_.emplace(key, *out);
        return 0;
    }

  private:
    static constexpr size_t kMaxEntries = 1 << 15;
    void load() {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            encode_ = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(fn);
    }
    PFN_cuTensorMapEncodeTiled_v12000 encode_ = nullptr;
    std::once_flag once_;
    std::mutex mu_;
    std::unordered_map<TmapKey, CUtensorMap, TmapKeyHash> map_;
};

// one cache per translation unit that includes this header is fine (each .cu has its own shapes)
static TmapCache g_tmaps;

}  // namespace ftl
