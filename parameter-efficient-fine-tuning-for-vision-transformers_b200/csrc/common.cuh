// Host-side helpers shared by the C-ABI translation units: thread-local error string, launch counter,
// CUDA error checks, cached device properties and the cuTensorMapEncodeTiled entry point (resolved at run time
// through cudart so that the library has no link-time dependency on libcuda and can be dlopen'ed on a CPU box).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <unordered_map>
#include <utility>
#include "../../include/peft_b200.h"

namespace peft {

inline char* tls_error() {
  static thread_local char buf[512] = {0};
  return buf;
}
inline int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(tls_error(), 512, fmt, ap);
  va_end(ap);
  return code;
}
inline std::atomic<unsigned long long>& launch_counter() {
  static std::atomic<unsigned long long> c{0};
  return c;
}
// per-kernel-family launch counters (peft_kernel_count): lets the tests assert WHICH kernel served a call
enum KernelKind { KIND_GEMM_1CTA = 0, KIND_GEMM_2CTA = 1, KIND_ATTN_TC_FWD = 2, KIND_ATTN_TC_BWD = 3, KIND_ATTN_LEGACY = 4,
                  KIND_GEMM_TN = 5, KIND_GEMM_4CTA = 6, KIND_COUNT = 16 };
inline std::atomic<unsigned long long>& kind_counter(int k) {
  static std::atomic<unsigned long long> c[KIND_COUNT];
  return c[(k >= 0 && k < KIND_COUNT) ? k : KIND_COUNT - 1];
}

#define PEFT_CUDA_CHECK(expr)                                                                         \
  do {                                                                                                \
    cudaError_t _e = (expr);                                                                          \
    if (_e != cudaSuccess)                                                                            \
      return ::peft::fail(PEFT_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
  } while (0)

#define PEFT_LAUNCH_CHECK()                                                                           \
  do {                                                                                                \
    ::peft::launch_counter().fetch_add(1, std::memory_order_relaxed);                                 \
    cudaError_t _e = cudaGetLastError();                                                              \
    if (_e != cudaSuccess)                                                                            \
      return ::peft::fail(PEFT_ERR_CUDA, "kernel launch failed: %s (%s:%d)", cudaGetErrorString(_e), __FILE__, __LINE__); \
  } while (0)

#define PEFT_LAUNCH_CHECK_KIND(kind)                                                                  \
  do {                                                                                                \
    ::peft::kind_counter(kind).fetch_add(1, std::memory_order_relaxed);                               \
    PEFT_LAUNCH_CHECK();                                                                              \
  } while (0)

#define PEFT_REQUIRE(cond, ...)                                   \
  do {                                                            \
    if (!(cond)) return ::peft::fail(PEFT_ERR_INVALID, __VA_ARGS__); \
  } while (0)

inline int current_device() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0) return 0;
  return dev & 63;
}
// SM count of the CURRENT device (cached per device: one process may drive several GPUs)
inline int sm_count() {
  static std::atomic<int> n[64];
  const int dev = current_device();
  int v = n[dev].load(std::memory_order_relaxed);
  if (v == 0) {
    if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) v = 148;
    n[dev].store(v, std::memory_order_relaxed);
  }
  return v;
}
// One-time per-DEVICE configuration of a kernel (cudaFuncSetAttribute is per device, not per process).  The guarded action is
// idempotent, so two threads racing through needs() on the same device is harmless.
struct DeviceOnce {
  std::atomic<unsigned long long> mask{0};
  bool needs() const { return !(mask.load(std::memory_order_acquire) & (1ull << current_device())); }
  void done() { mask.fetch_or(1ull << current_device(), std::memory_order_release); }
};
// same for a configured maximum (dynamic shared memory raised on demand)
struct DeviceMax {
  std::atomic<int> v[64];
  int get() const { return v[current_device()].load(std::memory_order_acquire); }
  void set(int x) { v[current_device()].store(x, std::memory_order_release); }
};

// ---------------------------------------------------------------- programmatic dependent launch (PDL)
// A step is ~290 dependent launches on one stream; with full stream serialisation every boundary costs the drain of the
// previous grid + the launch latency + the prologue (barrier init, TMEM allocation, tensor-map fetch, cluster sync) of the
// next one.  Kernels launched through launch_pdl() carry cudaLaunchAttributeProgrammaticStreamSerialization: their CTAs may
// be scheduled once every CTA of the predecessor has executed pdl_trigger() (pdl_sm100.cuh) and an SM has room, run their
// prologue, and block in pdl_wait() until the predecessor grid has COMPLETED and its memory is visible.  Rules every PDL
// kernel in this library follows: no global-memory access of any kind before pdl_wait(), and every CTA calls it — so
// completion of a grid still implies completion of everything before it.  Captured by CUDA graphs as programmatic edges.
// WHERE the trigger sits matters (measured, r02): the tensor-core kernels (one CTA per SM, CTAs finish at different times)
// trigger at their very END — with an early trigger the grid-stride CTAs of a following LayerNorm / token reduction were
// placed on the SMs that happened to be free first (up to 8 per SM) instead of evenly, which cost 12 us per such launch
// (+0.7 ms per C3 step); the small kernels, whose CTAs are spread evenly, trigger at their start.  Effect on C3 (graph replay,
// same box): B = 256 28.07 -> 27.93 ms, B = 32 (the 8-GPU split) 4.94 -> 4.81 ms; GEMM launch floor 7.4 -> 6.4 us.
// PEFT_PDL=0 (or peft_set_pdl(0)) launches everything fully serialised; PEFT_PDL=<mask> enables single kernel families.
// kernel families (bit mask in PEFT_PDL / peft_set_pdl: A/B measurements per family)
enum { PDL_GEMM2 = 1, PDL_GEMM1 = 2, PDL_ATTN = 4, PDL_LN = 8, PDL_REDUCE = 16, PDL_MISC = 32, PDL_ALL = 63 };
inline std::atomic<int>& pdl_mode() {
  static std::atomic<int> m{[] { const char* e = getenv("PEFT_PDL"); return e ? atoi(e) : (int)PDL_ALL; }()};
  return m;
}
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(int family, void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                              Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = (pdl_mode().load(std::memory_order_relaxed) & family) ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
inline EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn == nullptr) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// 2-D row-major [rows, cols] tensor (bf16 or fp32) with row pitch ld (elements); box = box_rows x box_cols with
// box_cols * elt == 128 bytes, 128-byte swizzle.  Loads outside [rows, cols] are zero-filled and stores are clipped by
// the hardware, which is what handles every M/N/K tail.  Encoded maps are cached (the workspaces of the engine are
// persistent, so a step re-uses the same few hundred maps): the key holds every encoded field, so a hit is always valid.
struct TmapKey {
  const void* ptr; unsigned long long rows, cols, ld; unsigned box_rows, box_cols; int f32;
  bool operator==(const TmapKey& o) const {
    return ptr == o.ptr && rows == o.rows && cols == o.cols && ld == o.ld && box_rows == o.box_rows && box_cols == o.box_cols && f32 == o.f32;
  }
};
struct TmapKeyHash {
  size_t operator()(const TmapKey& k) const {
    size_t h = reinterpret_cast<size_t>(k.ptr) * 0x9E3779B97F4A7C15ull;
    h ^= (k.rows * 0x100000001B3ull) ^ (k.cols << 17) ^ (k.ld << 29) ^ ((size_t)k.box_rows << 7) ^ ((size_t)k.box_cols << 3) ^ (size_t)k.f32;
    return h;
  }
};
inline int make_tmap_2d(CUtensorMap* tm, const void* ptr, bool f32, unsigned long long rows, unsigned long long cols,
                        unsigned long long ld, unsigned box_rows, unsigned box_cols) {
  static std::mutex mu;
  static std::unordered_map<TmapKey, CUtensorMap, TmapKeyHash> cache;
  const TmapKey key{ptr, rows, cols, ld, box_rows, box_cols, f32 ? 1 : 0};
  {
    std::lock_guard<std::mutex> g(mu);
    auto it = cache.find(key);
    if (it != cache.end()) { *tm = it->second; return 0; }
  }
  EncodeTiledFn fn = encode_tiled_fn();
  const unsigned elt = f32 ? 4 : 2;
  if (fn == nullptr) return fail(PEFT_ERR_CUDA, "cuTensorMapEncodeTiled entry point unavailable (no CUDA driver?)");
  if ((reinterpret_cast<uintptr_t>(ptr) & 15) != 0) return fail(PEFT_ERR_INVALID, "TMA base pointer must be 16-byte aligned");
  if ((ld * elt) % 16 != 0) return fail(PEFT_ERR_INVALID, "TMA row pitch must be a multiple of 16 bytes (ld=%llu)", ld);
  if (box_cols * elt != 128 || box_rows == 0 || box_rows > 256) return fail(PEFT_ERR_INVALID, "TMA box must be 128 bytes wide and <= 256 rows");
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {ld * elt};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(tm, f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims,
                  strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(PEFT_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  {
    std::lock_guard<std::mutex> g(mu);
    if (cache.size() > 262144) cache.clear();     // ~34 MB of descriptors: far above what persistent workspaces produce; a
                                                  // caller that streams fresh pointers through the library just re-encodes
    cache.emplace(key, *tm);
  }
  return 0;
}
// 3-D bf16 view [batch, rows, cols] of a row-major [batch*rows, ld] buffer (batch stride = rows * ld): box =
// 1 x box_rows x 64 columns.  Used by the attention kernels: a (batch, head) tile never crosses into the next batch
// element because rows >= `rows` are out of bounds in dimension 1 (zero-filled on load, clipped on store).
inline int make_tmap_bf16_3d(CUtensorMap* tm, const void* ptr, unsigned long long batch, unsigned long long rows,
                             unsigned long long cols, unsigned long long ld, unsigned box_rows) {
  static std::mutex mu;
  static std::unordered_map<TmapKey, CUtensorMap, TmapKeyHash> cache;
  const TmapKey key{ptr, rows, cols, ld, box_rows, (unsigned)batch, 3};
  {
    std::lock_guard<std::mutex> g(mu);
    auto it = cache.find(key);
    if (it != cache.end()) { *tm = it->second; return 0; }
  }
  EncodeTiledFn fn = encode_tiled_fn();
  if (fn == nullptr) return fail(PEFT_ERR_CUDA, "cuTensorMapEncodeTiled entry point unavailable (no CUDA driver?)");
  if ((reinterpret_cast<uintptr_t>(ptr) & 15) != 0) return fail(PEFT_ERR_INVALID, "TMA base pointer must be 16-byte aligned");
  if ((ld * 2) % 16 != 0) return fail(PEFT_ERR_INVALID, "TMA row pitch must be a multiple of 16 bytes (ld=%llu)", ld);
  if (box_rows == 0 || box_rows > 256) return fail(PEFT_ERR_INVALID, "TMA box rows must be in [1, 256]");
  cuuint64_t dims[3] = {cols, rows, batch};
  cuuint64_t strides[2] = {ld * 2, rows * ld * 2};
  cuuint32_t box[3] = {64, box_rows, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(ptr), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(PEFT_ERR_CUDA, "cuTensorMapEncodeTiled (3-D) failed with CUresult %d", (int)r);
  {
    std::lock_guard<std::mutex> g(mu);
    if (cache.size() > 262144) cache.clear();     // ~34 MB of descriptors: far above what persistent workspaces produce; a
                                                  // caller that streams fresh pointers through the library just re-encodes
    cache.emplace(key, *tm);
  }
  return 0;
}
inline int make_tmap_bf16_2d(CUtensorMap* tm, const void* ptr, unsigned long long rows, unsigned long long cols,
                             unsigned long long ld, unsigned box_rows) {
  return make_tmap_2d(tm, ptr, false, rows, cols, ld, box_rows, 64);
}

}  // namespace peft
