// Shared helpers for the skghoi_b200 CUDA library (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>

#define SKG_OK 0
#define SKG_ERR_INVALID (-1)
#define SKG_ERR_CUDA (-2)
#define SKG_ERR_UNSUPPORTED (-3)

namespace skg {

// last error text, retrievable through skg_last_error() (no exceptions cross the C boundary)
void set_error(const char* fmt, ...);
int check_launch(const char* what);
void add_launches(long long n);   // adjusts skg_launch_count (graph capture / replay)

#define SKG_REQUIRE(cond, ...)                                  \
  do {                                                          \
    if (!(cond)) {                                              \
      skg::set_error(__VA_ARGS__);                              \
      return SKG_ERR_INVALID;                                   \
    }                                                           \
  } while (0)

#define SKG_CUDA(call)                                                              \
  do {                                                                              \
    cudaError_t e__ = (call);                                                       \
    if (e__ != cudaSuccess) {                                                       \
      skg::set_error("%s failed: %s", #call, cudaGetErrorString(e__));              \
      return SKG_ERR_CUDA;                                                          \
    }                                                                               \
  } while (0)

__host__ __device__ inline int64_t ceil_div64(int64_t a, int64_t b) { return (a + b - 1) / b; }
__host__ __device__ inline int ceil_div(int a, int b) { return (a + b - 1) / b; }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// number of SMs of the current device (cached)
int num_sms();

// ---- capacity mode (SKG_HEAD_CAPACITY): launches that read input addresses / true row counts from the segment table ----
int cast_pad_bf16_indirect(const unsigned long long* src_slot, int64_t ld_src, const int32_t* rows_dev, int64_t rows_cap, int cols,
                           void* dst, int64_t ld_dst, int cols_pad, cudaStream_t stream);
int global_pool_indirect(const unsigned long long* f_slot, int64_t planes, int hw, void* out_bf16, float* out_f32, cudaStream_t stream);
int pairs_spatial_encode_impl(const float* boxes, const unsigned long long* boxes_slot, const int32_t* box_off, const int32_t* n_h,
                              const int64_t* cell_off, const int64_t* pair_off, const int32_t* hum_off, const float* hw,
                              int num_images, int64_t total_cells, int num_features, float eps, int nan_guard, float* enc,
                              int64_t* x, int64_t* y, int64_t* x_keep, int64_t* y_keep, int32_t* cell2pair, int32_t* cell_hrow,
                              int32_t* cell_orow, int32_t* cell_img, int32_t* pair2cell, int32_t* nan_flag, cudaStream_t stream);

}  // namespace skg
