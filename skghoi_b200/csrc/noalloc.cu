// Allocation-free variant of the split-K GEMM entry point (include/skghoi_b200.h: skg_gemm_bf16_splitk_static).
// No counterpart in the reference: added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
//
// skg_gemm_bf16_splitk reduces its K splits through a CALLER-owned scratch region (1024 tile counters + one fp32 tile per
// output tile, csrc/gemm.cu).  Here that region is a file-scope device array of SKG_GEMM_STATIC_SCRATCH_BYTES: every GPU
// that loads the library holds it until the process exits, every call on a device shares it.
//
//  * In host code the array's NAME is a host-side stand-in, not a device address.  Its device address is taken once per
//    device -- with that device current -- through cudaGetSymbolAddress, and that pointer is what the GEMM receives.
//  * The array keeps the last call's contents.  The kernel wants the region zero on entry and leaves it zero on exit; a call
//    that was aborted mid-way (or raced) could leave partial sums behind, so the part a launch uses is cleared on the
//    caller's stream in front of it (cudaMemsetAsync: a stream operation, no allocation).
//  * Calls that overlap on the GPU (two streams, two host threads on one device) RACE on the array: serialise them, or use
//    skg_gemm_bf16_splitk with one scratch region per stream.
//  * Nothing here calls cudaMalloc / cudaFree or an allocating library routine; neither does the launch path of csrc/gemm.cu
//    (tensor maps are encoded on the host into a host-side cache and passed as kernel parameters).
#include "common.cuh"
#include "skghoi_b200.h"
#include <mutex>

namespace skg {

static __device__ __align__(256) unsigned char g_splitk_static_scratch[SKG_GEMM_STATIC_SCRATCH_BYTES];

constexpr int kMaxDevices = 64;
constexpr int64_t kCounterBytes = 4096;   // the 1024 int tile counters at the head of a split-K scratch region (csrc/gemm.cu)

// device address of the array on the CURRENT device (looked up once per device)
static int static_scratch_address(void** out) {
  static std::mutex mu;
  static void* addr[kMaxDevices] = {};
  int dev = -1;
  SKG_CUDA(cudaGetDevice(&dev));
  SKG_REQUIRE(dev >= 0 && dev < kMaxDevices, "device ordinal %d outside [0, %d)", dev, kMaxDevices);
  std::lock_guard<std::mutex> lock(mu);
  if (!addr[dev]) {
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_splitk_static_scratch);
    if (e != cudaSuccess || !p) {
      cudaGetLastError();
      set_error("cudaGetSymbolAddress of the static split-K scratch failed on device %d: %s", dev, cudaGetErrorString(e));
      return SKG_ERR_CUDA;
    }
    addr[dev] = p;
  }
  *out = addr[dev];
  return SKG_OK;
}

}  // namespace skg

using namespace skg;

extern "C" int64_t skg_gemm_static_scratch_bytes(void) { return (int64_t)SKG_GEMM_STATIC_SCRATCH_BYTES; }

extern "C" int skg_gemm_bf16_splitk_static(const void* A, int64_t lda, const void* B, int64_t ldb, int M, int N, int K,
                                           const skg_epilogue_group* groups, int num_groups, int block_n, cudaStream_t stream) {
  SKG_REQUIRE(M >= 0 && N > 0 && K > 0, "bad GEMM shape M=%d N=%d K=%d", M, N, K);
  SKG_REQUIRE(block_n == 0 || block_n == 128 || block_n == 256, "block_n must be 0 (auto), 128 or 256");
  // scratch a launch can touch: the counters + one fp32 tile per output tile.  For any tile width the tiles cover at most
  // the output padded to 128 rows x 256 columns, which is the bound checked (and cleared) here.
  const int64_t rows_pad = ((int64_t)M + 127) / 128 * 128, cols_pad = ((int64_t)N + 255) / 256 * 256;
  const int64_t need = kCounterBytes + rows_pad * cols_pad * 4;
  SKG_REQUIRE(need <= (int64_t)SKG_GEMM_STATIC_SCRATCH_BYTES,
              "skg_gemm_bf16_splitk_static: M=%d N=%d needs %lld bytes of split-K scratch (output padded to %lld x %lld fp32 + %lld), "
              "the static array holds %lld: use skg_gemm_bf16_splitk with a caller-owned region",
              M, N, (long long)need, (long long)rows_pad, (long long)cols_pad, (long long)kCounterBytes,
              (long long)SKG_GEMM_STATIC_SCRATCH_BYTES);
  if (M == 0) return SKG_OK;
  void* scratch = nullptr;
  const int rc = static_scratch_address(&scratch);
  if (rc) return rc;
  SKG_CUDA(cudaMemsetAsync(scratch, 0, (size_t)need, stream));
  return skg_gemm_bf16_splitk(A, lda, B, ldb, M, N, K, groups, num_groups, block_n, scratch, (int64_t)SKG_GEMM_STATIC_SCRATCH_BYTES,
                              stream);
}
