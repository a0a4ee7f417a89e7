// Detection pre-processing of the interaction head on the device (reference InteractionHead.preprocess,
// head.py:92-151): score filter, class-wise NMS, score ordering, humans first, caps.
//
// One CTA per image, everything in shared memory (an image carries at most a few hundred detections):
//   1. active = score >= box_score_thresh                                     (head.py:118-120)
//   2. stable descending rank by score among the active boxes                 (torchvision nms sorts stably)
//   3. class-wise NMS exactly as torchvision.ops.batched_nms does it for small inputs -- the "coordinate trick":
//      boxes are shifted by label * (max active coordinate + 1) in fp32 and a plain greedy NMS runs on the shifted
//      boxes with torchvision's IoU expression, inter / (Sa + Sb - inter) > thresh        (head.py:122-127)
//   4. survivors in score order (head.py:128-130 re-sorts an already sorted list), humans first, at most
//      max_human humans and max_object others                                 (head.py:132-141)
// The kept detections come back as GLOBAL row indices into the concatenated detection arrays, image b's at
// out_index[b * (max_human + max_object) ...], with (n, n_h) per image in out_counts: bit-exact with the reference
// (integer outputs), no host round trip inside (the reference syncs four times per image: three nonzero() + NMS).
#include "common.cuh"
#include "skghoi_b200.h"

namespace skg {

constexpr int kPreThreads = 256;

__device__ __forceinline__ bool nms_iou_above(float4 a, float4 b, float thresh) {   // torchvision devIoU, fp32
  const float left = fmaxf(a.x, b.x), right = fminf(a.z, b.z);
  const float top = fmaxf(a.y, b.y), bottom = fminf(a.w, b.w);
  const float width = fmaxf(__fsub_rn(right, left), 0.f), height = fmaxf(__fsub_rn(bottom, top), 0.f);
  const float inter = __fmul_rn(width, height);
  const float sa = __fmul_rn(__fsub_rn(a.z, a.x), __fsub_rn(a.w, a.y));
  const float sb = __fmul_rn(__fsub_rn(b.z, b.x), __fsub_rn(b.w, b.y));
  return __fdiv_rn(inter, __fsub_rn(__fadd_rn(sa, sb), inter)) > thresh;
}

__global__ void __launch_bounds__(kPreThreads) preprocess_kernel(
    const float4* __restrict__ boxes, const int64_t* __restrict__ labels, const float* __restrict__ scores,
    const int32_t* __restrict__ det_off, int max_n, float score_thresh, float nms_thresh, int human_idx, int max_human,
    int max_object, int64_t* __restrict__ out_index, int32_t* __restrict__ out_counts) {
  extern __shared__ __align__(16) unsigned char pre_smem[];
  float4* sbox = reinterpret_cast<float4*>(pre_smem);                  // [max_n] shifted boxes in score order
  float* sscore = reinterpret_cast<float*>(sbox + max_n);              // [max_n] raw scores (original order)
  int* slabel = reinterpret_cast<int*>(sscore + max_n);                // [max_n] labels (original order)
  int* sorder = slabel + max_n;                                        // [max_n] score rank -> original index
  unsigned char* sactive = reinterpret_cast<unsigned char*>(sorder + max_n);   // [max_n]
  unsigned char* ssupp = sactive + max_n;                              // [max_n] suppressed, by score rank
  __shared__ float red_max[kPreThreads / 32];
  __shared__ int s_nact;
  __shared__ float s_shift;

  const int b = blockIdx.x, tid = threadIdx.x;
  const int off = det_off[b], n = det_off[b + 1] - off;
  const int max_keep = max_human + max_object;
  if (tid == 0) s_nact = 0;
  // 1. load + score filter; max coordinate over the ACTIVE boxes (boxes[active].max())
  float mx = -INFINITY;
  for (int i = tid; i < n; i += kPreThreads) {
    const float s = scores[off + i];
    const bool act = s >= score_thresh;
    sscore[i] = s;
    slabel[i] = (int)labels[off + i];
    sactive[i] = act ? 1 : 0;
    if (act) {
      const float4 bx = boxes[off + i];
      mx = fmaxf(mx, fmaxf(fmaxf(bx.x, bx.y), fmaxf(bx.z, bx.w)));
    }
  }
  mx = warp_max(mx);
  if ((tid & 31) == 0) red_max[tid >> 5] = mx;
  __syncthreads();
  if (tid == 0) {
    float m = red_max[0];
    for (int w = 1; w < kPreThreads / 32; ++w) m = fmaxf(m, red_max[w]);
    s_shift = __fadd_rn(m, 1.0f);                                      // max_coordinate + 1
  }
  // 2. stable descending rank among the active boxes (ties keep the original order)
  for (int i = tid; i < n; i += kPreThreads) {
    if (!sactive[i]) continue;
    const float s = sscore[i];
    int r = 0;
    for (int j = 0; j < n; ++j) r += (sactive[j] && (sscore[j] > s || (sscore[j] == s && j < i))) ? 1 : 0;
    sorder[r] = i;
    atomicAdd(&s_nact, 1);
  }
  __syncthreads();
  const int nact = s_nact;
  const float shift = s_shift;
  // 3. shifted boxes in score order: box + float(label) * (max + 1)          (_batched_nms_coordinate_trick)
  for (int k = tid; k < nact; k += kPreThreads) {
    const int i = sorder[k];
    const float4 bx = boxes[off + i];
    const float o = __fmul_rn((float)slabel[i], shift);
    sbox[k] = make_float4(__fadd_rn(bx.x, o), __fadd_rn(bx.y, o), __fadd_rn(bx.z, o), __fadd_rn(bx.w, o));
    ssupp[k] = 0;
  }
  __syncthreads();
  for (int k = 0; k < nact; ++k) {                 // greedy: a surviving box removes every later box it overlaps
    if (!ssupp[k]) {                               // uniform across the CTA (read after the barrier below)
      const float4 a = sbox[k];
      for (int j = k + 1 + tid; j < nact; j += kPreThreads)
        if (!ssupp[j] && nms_iou_above(a, sbox[j], nms_thresh)) ssupp[j] = 1;
    }
    __syncthreads();
  }
  // 4. humans first, caps; global row indices
  if (tid == 0) {
    int nh = 0, no = 0;
    int64_t* dst = out_index + (int64_t)b * max_keep;
    for (int k = 0; k < nact; ++k) {
      if (ssupp[k]) continue;
      const int i = sorder[k];
      if (slabel[i] == human_idx) { if (nh < max_human) dst[nh++] = off + i; }
    }
    for (int k = 0; k < nact; ++k) {
      if (ssupp[k]) continue;
      const int i = sorder[k];
      if (slabel[i] != human_idx) { if (no < max_object) dst[nh + no++] = off + i; }
    }
    for (int z = nh + no; z < max_keep; ++z) dst[z] = -1;
    out_counts[2 * b] = nh + no;
    out_counts[2 * b + 1] = nh;
  }
}

}  // namespace skg

using namespace skg;

extern "C" int skg_preprocess_detections(const float* boxes, const int64_t* labels, const float* scores,
                                         const int32_t* det_off, int num_images, int max_dets_per_image,
                                         float box_score_thresh, float box_nms_thresh, int human_idx, int max_human,
                                         int max_object, int64_t* out_index, int32_t* out_counts, cudaStream_t stream) {
  if (num_images == 0) return SKG_OK;
  SKG_REQUIRE(boxes && labels && scores && det_off && out_index && out_counts, "null pointer argument");
  SKG_REQUIRE((reinterpret_cast<uintptr_t>(boxes) & 15u) == 0, "boxes must be 16-byte aligned (rows are read as float4)");
  SKG_REQUIRE(max_human >= 0 && max_object >= 0 && max_human + max_object > 0, "bad caps");
  const int max_n = max_dets_per_image < 1 ? 1 : max_dets_per_image;
  const size_t smem = (size_t)max_n * (sizeof(float4) + sizeof(float) + 2 * sizeof(int) + 2);
  SKG_REQUIRE(smem <= 200 * 1024, "too many detections per image for the shared-memory NMS (%d)", max_dets_per_image);
  static bool attr_set = false;
  if (!attr_set) {
    SKG_CUDA(cudaFuncSetAttribute(preprocess_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    attr_set = true;
  }
  preprocess_kernel<<<num_images, kPreThreads, smem, stream>>>((const float4*)boxes, labels, scores, det_off, max_n,
                                                              box_score_thresh, box_nms_thresh, human_idx, max_human,
                                                              max_object, out_index, out_counts);
  return check_launch("preprocess_kernel");
}
