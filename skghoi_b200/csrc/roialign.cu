// Multi-scale RoIAlign of the interaction head on the device (reference: torchvision MultiScaleRoIAlign(['0','1','2','3'],
// 7, 2) built at models.py:158-162 and called at head.py:387).
//
//   level(box)  = clamp(floor(lvl0 + log2(sqrt(area) / s0) + 1e-6), k_min, k_max) - k_min        (poolers.LevelMapper)
//   roi_align   = aligned=False bilinear pooling, sampling_ratio x sampling_ratio samples per bin, on that level's map
//                 scaled by the level's spatial scale                                               (ops/roi_align)
//
// One CTA per (box, channel slab).  The 2 * pooled * sampling_ratio sample coordinates of a box (low / high index and the
// two weights per axis) are computed once into shared memory; a thread then owns consecutive outputs of the box's
// [C, pooled, pooled] block, so the stores are coalesced.  The block is written as fp32 in the reference layout
// ([sum n, C, 7, 7]) and / or as bf16 straight into the K-major operand of box_head's first GEMM (K = C * 49), which
// takes the 12.8 MB fp32 tensor and the cast kernel out of the path when the caller does not need the fp32 copy.
#include "common.cuh"
#include "skghoi_b200.h"
#include <cuda_bf16.h>

namespace skg {

constexpr int kRoiThreads = 256;
constexpr int kRoiMaxLevels = 8;
constexpr int kRoiMaxSamples = 64;      // pooled * sampling_ratio per axis

struct RoiLevels {
  const float* feat[kRoiMaxLevels];     // [B, C, H_l, W_l] fp32
  int h[kRoiMaxLevels], w[kRoiMaxLevels];
  float scale[kRoiMaxLevels];
  int num;
};

struct AxisSample {
  int lo, hi;
  float wl, wh;     // weights of hi (l = coordinate - lo) and lo (h = 1 - l); both 0 for a sample outside the map
};

// bilinear_interpolate's per-axis part (torchvision/csrc/ops/cpu/roi_align_common.h pre_calc_for_bilinear_interpolate)
__device__ __forceinline__ AxisSample axis_sample(float v, int size) {
  AxisSample s;
  if (v < -1.0f || v > (float)size) { s.lo = s.hi = 0; s.wl = s.wh = 0.f; return s; }
  if (v <= 0.f) v = 0.f;
  int lo = (int)v, hi;
  if (lo >= size - 1) { hi = lo = size - 1; v = (float)lo; } else { hi = lo + 1; }
  s.lo = lo; s.hi = hi;
  s.wl = __fsub_rn(v, (float)lo);
  s.wh = __fsub_rn(1.f, s.wl);
  return s;
}

__global__ void __launch_bounds__(kRoiThreads) roi_align_kernel(
    RoiLevels lv, int channels, const float4* __restrict__ boxes, const int32_t* __restrict__ box_off, int num_images,
    int pooled, int ratio, int k_min, int k_max, float s0, float lvl0, float* __restrict__ out_f32,
    __nv_bfloat16* __restrict__ out_bf16, int64_t ld_bf16, int chan_per_cta) {
  __shared__ AxisSample sy[kRoiMaxSamples], sx[kRoiMaxSamples];
  __shared__ int s_level, s_image;
  __shared__ bool s_outside_y[kRoiMaxSamples], s_outside_x[kRoiMaxSamples];
  const int64_t box = blockIdx.x;
  const float4 bx = boxes[box];
  const int ns = pooled * ratio;
  if (threadIdx.x == 0) {
    int img = 0;
    while (img + 1 < num_images && box >= box_off[img + 1]) ++img;
    s_image = img;
    int level = 0;
    if (lv.num > 1) {
      const float area = __fmul_rn(__fsub_rn(bx.z, bx.x), __fsub_rn(bx.w, bx.y));
      const float s = __fsqrt_rn(area);
      float t = floorf(__fadd_rn(__fadd_rn(lvl0, log2f(__fdiv_rn(s, s0))), 1e-6f));
      t = fminf(fmaxf(t, (float)k_min), (float)k_max);          // NaN (degenerate box) propagates like torch.clamp ...
      level = (t == t) ? (int)t - k_min : 0;                      // ... and its int64 cast is undefined there: level 0
      if (level < 0) level = 0;
      if (level >= lv.num) level = lv.num - 1;
    }
    s_level = level;
  }
  __syncthreads();
  const int level = s_level;
  const int H = lv.h[level], W = lv.w[level];
  const float scale = lv.scale[level];
  if (threadIdx.x < 2 * ns) {
    const bool is_y = threadIdx.x < ns;
    const int i = is_y ? threadIdx.x : threadIdx.x - ns;
    const float start = __fmul_rn(is_y ? bx.y : bx.x, scale), end = __fmul_rn(is_y ? bx.w : bx.z, scale);
    const float extent = fmaxf(__fsub_rn(end, start), 1.0f);     // aligned = False: at least one feature-map pixel
    const float bin = __fdiv_rn(extent, (float)pooled);
    const int p = i / ratio, k = i - p * ratio;
    // y = roi_start_h + ph * bin_size_h + (iy + .5f) * bin_size_h / roi_bin_grid_h
    const float v = __fadd_rn(__fadd_rn(start, __fmul_rn((float)p, bin)),
                              __fdiv_rn(__fmul_rn(__fadd_rn((float)k, 0.5f), bin), (float)ratio));
    const int size = is_y ? H : W;
    const AxisSample a = axis_sample(v, size);
    const bool outside = v < -1.0f || v > (float)size;
    if (is_y) { sy[i] = a; s_outside_y[i] = outside; } else { sx[i] = a; s_outside_x[i] = outside; }
  }
  __syncthreads();
  const int bins = pooled * pooled;
  const int c_begin = blockIdx.y * chan_per_cta;
  const int c_end = min(channels, c_begin + chan_per_cta);
  const float* fmap = lv.feat[level] + (int64_t)s_image * channels * H * W;
  const float inv_count = (float)(ratio * ratio);
  for (int o = c_begin * bins + threadIdx.x; o < c_end * bins; o += kRoiThreads) {
    const int c = o / bins, bin = o - c * bins;
    const int ph = bin / pooled, pw = bin - ph * pooled;
    const float* plane = fmap + (int64_t)c * H * W;
    float acc = 0.f;
    for (int iy = 0; iy < ratio; ++iy) {
      const AxisSample ay = sy[ph * ratio + iy];
      const bool oy = s_outside_y[ph * ratio + iy];
      for (int ix = 0; ix < ratio; ++ix) {
        const AxisSample ax = sx[pw * ratio + ix];
        if (oy || s_outside_x[pw * ratio + ix]) continue;         // empty sample: contributes 0
        const float w1 = __fmul_rn(ay.wh, ax.wh), w2 = __fmul_rn(ay.wh, ax.wl);
        const float w3 = __fmul_rn(ay.wl, ax.wh), w4 = __fmul_rn(ay.wl, ax.wl);
        const float v1 = __ldg(plane + ay.lo * W + ax.lo), v2 = __ldg(plane + ay.lo * W + ax.hi);
        const float v3 = __ldg(plane + ay.hi * W + ax.lo), v4 = __ldg(plane + ay.hi * W + ax.hi);
        acc = __fadd_rn(acc, __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(w1, v1), __fmul_rn(w2, v2)), __fmul_rn(w3, v3)),
                                       __fmul_rn(w4, v4)));
      }
    }
    acc = __fdiv_rn(acc, inv_count);
    if (out_f32) out_f32[box * (int64_t)channels * bins + o] = acc;
    if (out_bf16) out_bf16[box * ld_bf16 + o] = __float2bfloat16_rn(acc);
  }
}

}  // namespace skg

using namespace skg;

extern "C" int skg_multiscale_roi_align(const float* const* feats, const int32_t* feat_h, const int32_t* feat_w,
                                        const float* scales, int num_levels, int channels, int num_images,
                                        const float* boxes, const int32_t* box_off, int64_t num_boxes, int pooled,
                                        int sampling_ratio, int k_min, int k_max, float canonical_scale,
                                        int canonical_level, float* out_f32, void* out_bf16, int64_t ld_bf16,
                                        cudaStream_t stream) {
  if (num_boxes == 0) return SKG_OK;
  SKG_REQUIRE(feats && feat_h && feat_w && scales && boxes && box_off, "null pointer argument");
  SKG_REQUIRE(out_f32 || out_bf16, "at least one of out_f32 / out_bf16 must be set");
  SKG_REQUIRE(num_levels >= 1 && num_levels <= kRoiMaxLevels, "num_levels must be in [1, %d]", kRoiMaxLevels);
  SKG_REQUIRE(channels > 0 && num_images > 0 && pooled > 0 && sampling_ratio > 0, "bad sizes");
  SKG_REQUIRE(pooled * sampling_ratio <= kRoiMaxSamples,
              "pooled * sampling_ratio must be <= %d (adaptive sampling_ratio <= 0 is not supported)", kRoiMaxSamples);
  SKG_REQUIRE((reinterpret_cast<uintptr_t>(boxes) & 15u) == 0, "boxes must be 16-byte aligned (rows are read as float4)");
  SKG_REQUIRE(!out_bf16 || ld_bf16 >= (int64_t)channels * pooled * pooled, "ld_bf16 is smaller than a box's block");
  RoiLevels lv;
  lv.num = num_levels;
  for (int l = 0; l < num_levels; ++l) {
    SKG_REQUIRE(feats[l] && feat_h[l] > 0 && feat_w[l] > 0, "level %d: bad feature map", l);
    lv.feat[l] = feats[l]; lv.h[l] = feat_h[l]; lv.w[l] = feat_w[l]; lv.scale[l] = scales[l];
  }
  // enough CTAs to fill the SMs even for a single image's boxes
  int slabs = 1;
  while ((int64_t)slabs * num_boxes < 4 * (int64_t)num_sms() && slabs < channels && slabs < 16) slabs *= 2;
  const int chan_per_cta = (channels + slabs - 1) / slabs;
  SKG_REQUIRE(num_boxes <= 0x7fffffff, "too many boxes");
  dim3 grid((unsigned)num_boxes, (unsigned)((channels + chan_per_cta - 1) / chan_per_cta));
  roi_align_kernel<<<grid, kRoiThreads, 0, stream>>>(lv, channels, reinterpret_cast<const float4*>(boxes), box_off, num_images,
                                                    pooled, sampling_ratio, k_min, k_max, canonical_scale,
                                                    (float)canonical_level, out_f32,
                                                    reinterpret_cast<__nv_bfloat16*>(out_bf16), ld_bf16, chan_per_cta);
  return check_launch("roi_align_kernel");
}
