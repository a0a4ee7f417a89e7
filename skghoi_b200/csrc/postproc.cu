// Prior scores, ground-truth association and the post-processing of the interaction head
// (reference head.py:703-767 associate_with_ground_truth / compute_prior_scores, head.py:291-337
// postprocess).  The reference does these with Python list comprehensions, `.item()` per pair and a
// `nonzero` per image; here every per-pair output position is a closed form of host-built per-box
// prefix sums, so the whole batch is one launch with no device->host sync.
// Compiled with -fmad=false (IoU threshold decisions and score products must round like the reference).
#include "common.cuh"
#include "skghoi_b200.h"

namespace skg {

struct PostTables {
  const int32_t* verb_off;   // [num_obj_cls + 1]  CSR of valid verbs per object class (ascending verb ids)
  const int32_t* verb_idx;   // [nnz]
  const int32_t* pair2cell;  // [P]
  const int32_t* cell_hrow;  // [cells] global human-node row
  const int32_t* cell_orow;  // [cells] global box row
  const int32_t* cell_img;   // [cells]
  const int32_t* hum_off;    // [B+1]
  const int32_t* box_off;    // [B+1]
  const int64_t* pair_off;   // [B+1]
  const int64_t* labels;     // [boxes] object class per box
  const float* scores;       // [boxes]
  const int32_t* box_prefix; // [boxes]  exclusive prefix (within image) of the per-box verb counts
  const int32_t* box_cnt;    // [boxes]  number of valid verbs of the box's class
  const int64_t* hum_base;   // [humans] output offset of the first entry of human row i (batch-global)
  const int32_t* logit_row;  // [P] row of pair p in the logits (NULL = p): the training layout keeps logits in cell rows
};

__device__ __forceinline__ float sigmoidf_ref(float x) { return 1.f / (1.f + expf(-x)); }

// One warp per valid pair.  Entry k of pair (i, j) is verb v = valid[label_j][k]:
//   score = sigmoid(logit[p, v]) * (s_h^pw * s_o^pw) * sigmoid(logit_s[p])           (head.py:316)
__global__ void postprocess_kernel(const float* __restrict__ logits, int ld_logits, int num_cls, PostTables t, int64_t pairs,
                                   float pw, const float* __restrict__ gt_labels /*[P, num_cls] or null*/,
                                   int64_t total_entries, int64_t* __restrict__ out_index, int64_t* __restrict__ out_pred,
                                   float* __restrict__ out_scores, float* __restrict__ out_prior /*[2, total]*/,
                                   float* __restrict__ out_weights /*[P]*/, float* __restrict__ out_labels /*[total]*/,
                                   float* __restrict__ out_unary /*[P]*/, int64_t* __restrict__ out_object /*[P]*/,
                                   const float4* __restrict__ boxes /*or null*/, float4* __restrict__ boxes_h,
                                   float4* __restrict__ boxes_o) {
  const int lane = threadIdx.x & 31;
  const int64_t p = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  if (p >= pairs) return;
  const int cell = __ldg(t.pair2cell + p);
  const int hr = __ldg(t.cell_hrow + cell), orow = __ldg(t.cell_orow + cell), b = __ldg(t.cell_img + cell);
  const int boff = __ldg(t.box_off + b);
  const int i = hr - __ldg(t.hum_off + b), j = orow - boff;
  const float s_h = __ldg(t.scores + boff + i), s_o = __ldg(t.scores + orow);
  const float ph = powf(s_h, pw), po = powf(s_o, pw);
  const float* __restrict__ lrow = logits + (t.logit_row ? (int64_t)__ldg(t.logit_row + p) : p) * ld_logits;
  const float w = sigmoidf_ref(__ldg(lrow + num_cls));
  const int64_t lab_o = __ldg(t.labels + orow);
  if (lane == 0) {
    out_weights[p] = w;
    if (out_object) out_object[p] = lab_o;
  }
  if (boxes && lane < 2) {           // boxes_h / boxes_o of the pair (head.py:975-976: coords[x_keep], coords[y_keep])
    if (lane == 0) boxes_h[p] = __ldg(boxes + boff + i); else boxes_o[p] = __ldg(boxes + orow);
  }
  if (gt_labels && out_unary) {     // unary_labels = clamp(sum_k labels[p, k], max=1)   (head.py:330)
    float s = 0.f;
    for (int v = lane; v < num_cls; v += 32) s += gt_labels[p * num_cls + v];
    s = warp_sum(s);
    if (lane == 0) out_unary[p] = fminf(s, 1.f);
  }
  if (!(s_h > 0.f)) return;          // prior_h row is all zero -> nonzero() yields nothing for this pair
  const int cnt = __ldg(t.box_cnt + orow);
  const int64_t base = __ldg(t.hum_base + hr) + __ldg(t.box_prefix + orow) - (j > i ? __ldg(t.box_cnt + boff + i) : 0);
  const int vo = __ldg(t.verb_off + lab_o);
  const int64_t pair_in_img = p - __ldg(t.pair_off + b);
  for (int k = lane; k < cnt; k += 32) {
    const int v = __ldg(t.verb_idx + vo + k);
    const float s = sigmoidf_ref(__ldg(lrow + v));
    const int64_t e = base + k;
    out_index[e] = pair_in_img;
    out_pred[e] = v;
    out_scores[e] = (s * (ph * po)) * w;
    out_prior[e] = ph;
    out_prior[total_entries + e] = po;
    if (out_labels) out_labels[e] = gt_labels[p * num_cls + v];
  }
}

// Dense prior [2, P, num_cls] (compute_prior_scores head.py:738-764) for the GraphHead API surface.
__global__ void prior_dense_kernel(PostTables t, int64_t pairs, int num_cls, float pw, float* __restrict__ prior) {
  const int lane = threadIdx.x & 31;
  const int64_t p = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  if (p >= pairs) return;
  const int cell = __ldg(t.pair2cell + p);
  const int hr = __ldg(t.cell_hrow + cell), orow = __ldg(t.cell_orow + cell), b = __ldg(t.cell_img + cell);
  const int boff = __ldg(t.box_off + b);
  const int i = hr - __ldg(t.hum_off + b);
  const float ph = powf(__ldg(t.scores + boff + i), pw), po = powf(__ldg(t.scores + orow), pw);
  for (int v = lane; v < num_cls; v += 32) { prior[p * num_cls + v] = 0.f; prior[(pairs + p) * num_cls + v] = 0.f; }
  __syncwarp();
  const int64_t lab_o = __ldg(t.labels + orow);
  const int vo = __ldg(t.verb_off + lab_o), cnt = __ldg(t.verb_off + lab_o + 1) - vo;
  for (int k = lane; k < cnt; k += 32) {
    const int v = __ldg(t.verb_idx + vo + k);
    prior[p * num_cls + v] = ph;
    prior[(pairs + p) * num_cls + v] = po;
  }
}

__device__ __forceinline__ float box_iou1(float4 a, float4 b) {   // torchvision box_iou element
  const float a1 = (a.z - a.x) * (a.w - a.y), a2 = (b.z - b.x) * (b.w - b.y);
  const float iw = fmaxf(fminf(a.z, b.z) - fmaxf(a.x, b.x), 0.f), ih = fmaxf(fminf(a.w, b.w) - fmaxf(a.y, b.y), 0.f);
  const float inter = iw * ih;
  return inter / ((a1 + a2) - inter);
}

// associate_with_ground_truth (head.py:703-719): labels[p, verb_g] = 1 where
// min(IoU(box_h, gt_h[g]), IoU(box_o, gt_o[g])) >= thresh.  One thread per pair, loop over the image's GT pairs.
__global__ void associate_gt_kernel(const float4* __restrict__ boxes, PostTables t, int64_t pairs, int num_cls,
                                    const int32_t* __restrict__ gt_off, const float4* __restrict__ gt_h,
                                    const float4* __restrict__ gt_o, const int64_t* __restrict__ gt_verb, float thresh,
                                    float* __restrict__ labels) {
  const int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (p >= pairs) return;
  const int cell = __ldg(t.pair2cell + p);
  const int hr = __ldg(t.cell_hrow + cell), orow = __ldg(t.cell_orow + cell), b = __ldg(t.cell_img + cell);
  const int boff = __ldg(t.box_off + b);
  const float4 bh = __ldg(boxes + boff + (hr - __ldg(t.hum_off + b))), bo = __ldg(boxes + orow);
  for (int g = __ldg(gt_off + b); g < __ldg(gt_off + b + 1); ++g) {
    const float m = fminf(box_iou1(bh, __ldg(gt_h + g)), box_iou1(bo, __ldg(gt_o + g)));
    int64_t v = __ldg(gt_verb + g);
    if (v < 0) v += num_cls;                                                      // tensor indexing wraps negative ids
    if (m >= thresh && v >= 0 && v < num_cls) labels[p * num_cls + v] = 1.f;   // range-checked on the host too (_gt_labels)
  }
}


// Output offsets of the post-processing, built on the device so the whole batch needs no host round trip:
//   box_cnt[k]    = number of valid verbs of box k's class
//   box_prefix[k] = exclusive prefix of box_cnt inside the image
//   row_tot[i]    = (score_i > 0) ? sum_j cnt_j - cnt_i : 0      entries emitted by human row i (head.py:315 nonzero)
//   img_entries[b] = sum_i row_tot[i]
// One CTA per image; the boxes of one image are scanned by warp 0 in chunks of 32.
__device__ __forceinline__ void post_counts_image(int b, int lane, const int32_t* __restrict__ verb_off,
                                                  const int64_t* __restrict__ labels, const float* __restrict__ scores,
                                                  const int32_t* __restrict__ box_off, const int32_t* __restrict__ hum_off,
                                                  int32_t* __restrict__ box_cnt, int32_t* __restrict__ box_prefix,
                                                  int64_t* __restrict__ hum_base, int64_t* __restrict__ img_entries) {
  const int lo = __ldg(box_off + b), n = __ldg(box_off + b + 1) - lo;
  const int h0 = __ldg(hum_off + b), nh = __ldg(hum_off + b + 1) - h0;
  int run = 0;
  for (int k0 = 0; k0 < n; k0 += 32) {
    const int k = k0 + lane;
    int c = 0;
    if (k < n) { const int64_t l = __ldg(labels + lo + k); c = __ldg(verb_off + l + 1) - __ldg(verb_off + l); }
    int inc = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += t; }
    if (k < n) { box_cnt[lo + k] = c; box_prefix[lo + k] = run + inc - c; }
    run += __shfl_sync(0xffffffffu, inc, 31);
  }
  __syncwarp();
  const int S = run;
  long long rrun = 0;
  for (int i0 = 0; i0 < nh; i0 += 32) {
    const int i = i0 + lane;
    long long t = 0;
    if (i < nh && __ldg(scores + lo + i) > 0.f) t = S - box_cnt[lo + i];
    long long inc = t;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const long long u = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += u; }
    if (i < nh) hum_base[h0 + i] = rrun + inc - t;       // image-local for now; post_bases_kernel adds the image base
    rrun += __shfl_sync(0xffffffffu, inc, 31);
  }
  if (lane == 0) img_entries[b] = rrun;
}

__global__ void post_counts_kernel(const int32_t* __restrict__ verb_off, const int64_t* __restrict__ labels,
                                   const float* __restrict__ scores, const int32_t* __restrict__ box_off,
                                   const int32_t* __restrict__ hum_off, int32_t* __restrict__ box_cnt,
                                   int32_t* __restrict__ box_prefix, int64_t* __restrict__ hum_base,
                                   int64_t* __restrict__ img_entries) {
  if (threadIdx.x >= 32) return;
  post_counts_image(blockIdx.x, threadIdx.x, verb_off, labels, scores, box_off, hum_off, box_cnt, box_prefix, hum_base, img_entries);
}

// The whole offset table in ONE single-CTA launch: a warp per image (round robin), then the image bases.
constexpr int kPostOffThreads = 512;
__global__ void __launch_bounds__(kPostOffThreads) post_offsets_kernel(
    const int32_t* __restrict__ verb_off, const int64_t* __restrict__ labels, const float* __restrict__ scores,
    const int32_t* __restrict__ box_off, const int32_t* __restrict__ hum_off, int B, int32_t* __restrict__ box_cnt,
    int32_t* __restrict__ box_prefix, int64_t* __restrict__ hum_base, int64_t* __restrict__ img_entries,
    int64_t* __restrict__ img_base) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int b = warp; b < B; b += kPostOffThreads / 32)
    post_counts_image(b, lane, verb_off, labels, scores, box_off, hum_off, box_cnt, box_prefix, hum_base, img_entries);
  __syncthreads();                                   // global writes of this CTA are visible to it after the barrier
  if (threadIdx.x == 0) {
    long long run = 0;
    for (int b = 0; b < B; ++b) { img_base[b] = run; run += img_entries[b]; }
    img_base[B] = run;
  }
  __syncthreads();
  const int total_h = __ldg(hum_off + B);
  for (int h = threadIdx.x; h < total_h; h += kPostOffThreads) {
    int lo = 0, hi = B;
    while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (__ldg(hum_off + mid) <= h) lo = mid; else hi = mid; }
    hum_base[h] += img_base[lo];
  }
}

__global__ void post_bases_kernel(const int32_t* __restrict__ hum_off, int B, const int64_t* __restrict__ img_entries,
                                  int64_t* __restrict__ img_base, int64_t* __restrict__ hum_base) {
  __shared__ long long carry;
  if (threadIdx.x == 0) {
    long long run = 0;
    for (int b = 0; b < B; ++b) { img_base[b] = run; run += img_entries[b]; }
    img_base[B] = run;
    carry = run;
  }
  __syncthreads();
  const int total_h = __ldg(hum_off + B);
  for (int h = threadIdx.x; h < total_h; h += blockDim.x) {
    int lo = 0, hi = B;
    while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (__ldg(hum_off + mid) <= h) lo = mid; else hi = mid; }
    hum_base[h] += img_base[lo];
  }
}

// boxes_h / boxes_o of every valid pair (head.py:975-976: coords[x_keep], coords[y_keep]).
__global__ void gather_pair_boxes_kernel(const float4* __restrict__ boxes, PostTables t, int64_t pairs,
                                         float4* __restrict__ boxes_h, float4* __restrict__ boxes_o) {
  const int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (p >= pairs) return;
  const int cell = __ldg(t.pair2cell + p);
  const int hr = __ldg(t.cell_hrow + cell), orow = __ldg(t.cell_orow + cell), b = __ldg(t.cell_img + cell);
  boxes_h[p] = __ldg(boxes + __ldg(t.box_off + b) + (hr - __ldg(t.hum_off + b)));
  boxes_o[p] = __ldg(boxes + orow);
}

}  // namespace skg

using namespace skg;

static int fill_tables(PostTables& t, const skg_post_tables* h) {
  SKG_REQUIRE(h && h->verb_off && h->verb_idx && h->pair2cell && h->cell_hrow && h->cell_orow && h->cell_img && h->hum_off &&
                  h->box_off && h->pair_off && h->labels && h->scores,
              "null pointer in skg_post_tables");
  t.verb_off = h->verb_off; t.verb_idx = h->verb_idx; t.pair2cell = h->pair2cell; t.cell_hrow = h->cell_hrow;
  t.cell_orow = h->cell_orow; t.cell_img = h->cell_img; t.hum_off = h->hum_off; t.box_off = h->box_off;
  t.pair_off = h->pair_off; t.labels = h->labels; t.scores = h->scores; t.box_prefix = h->box_prefix;
  t.box_cnt = h->box_cnt; t.hum_base = h->hum_base; t.logit_row = h->logit_row;
  return SKG_OK;
}

static int postprocess_launch(const float* logits, int ld_logits, int num_cls, const skg_post_tables* tables, int64_t pairs,
                              float prior_power, const float* gt_labels, int64_t total_entries, int64_t* out_index,
                              int64_t* out_pred, float* out_scores, float* out_prior, float* out_weights,
                              float* out_labels, float* out_unary, int64_t* out_object, const float* boxes, float* boxes_h,
                              float* boxes_o, cudaStream_t stream) {
  if (pairs == 0) return SKG_OK;
  PostTables t;
  int rc = fill_tables(t, tables);
  if (rc) return rc;
  SKG_REQUIRE(t.box_prefix && t.box_cnt && t.hum_base, "postprocess needs box_prefix / box_cnt / hum_base");
  SKG_REQUIRE(logits && out_weights && (total_entries == 0 || (out_index && out_pred && out_scores && out_prior)),
              "null output pointer");
  SKG_REQUIRE(!out_labels || gt_labels, "out_labels needs gt_labels");
  postprocess_kernel<<<(int)ceil_div64(pairs * 32, 256), 256, 0, stream>>>(logits, ld_logits, num_cls, t, pairs, prior_power,
                                                                          gt_labels, total_entries, out_index, out_pred,
                                                                          out_scores, out_prior, out_weights, out_labels,
                                                                          out_unary, out_object, (const float4*)boxes,
                                                                          (float4*)boxes_h, (float4*)boxes_o);
  return check_launch("postprocess_kernel");
}

extern "C" int skg_postprocess(const float* logits, int ld_logits, int num_cls, const skg_post_tables* tables, int64_t pairs,
                               float prior_power, const float* gt_labels, int64_t total_entries, int64_t* out_index,
                               int64_t* out_pred, float* out_scores, float* out_prior, float* out_weights,
                               float* out_labels, float* out_unary, int64_t* out_object, cudaStream_t stream) {
  return postprocess_launch(logits, ld_logits, num_cls, tables, pairs, prior_power, gt_labels, total_entries, out_index, out_pred,
                            out_scores, out_prior, out_weights, out_labels, out_unary, out_object, nullptr, nullptr, nullptr,
                            stream);
}

extern "C" int skg_prior_dense(const skg_post_tables* tables, int64_t pairs, int num_cls, float prior_power, float* prior,
                               cudaStream_t stream) {
  if (pairs == 0) return SKG_OK;
  PostTables t;
  int rc = fill_tables(t, tables);
  if (rc) return rc;
  SKG_REQUIRE(prior, "null output pointer");
  prior_dense_kernel<<<(int)ceil_div64(pairs * 32, 256), 256, 0, stream>>>(t, pairs, num_cls, prior_power, prior);
  return check_launch("prior_dense_kernel");
}

extern "C" int skg_associate_gt(const float* boxes, const skg_post_tables* tables, int64_t pairs, int num_cls,
                                const int32_t* gt_off, const float* gt_boxes_h, const float* gt_boxes_o,
                                const int64_t* gt_verb, float fg_iou_thresh, float* labels, cudaStream_t stream) {
  if (pairs == 0) return SKG_OK;
  PostTables t;
  int rc = fill_tables(t, tables);
  if (rc) return rc;
  SKG_REQUIRE(boxes && gt_off && labels, "null pointer argument");
  SKG_REQUIRE(((reinterpret_cast<uintptr_t>(boxes) | reinterpret_cast<uintptr_t>(gt_boxes_h) | reinterpret_cast<uintptr_t>(gt_boxes_o)) & 15u) == 0,
              "boxes and GT boxes must be 16-byte aligned (rows are read as float4)");
  SKG_CUDA(cudaMemsetAsync(labels, 0, sizeof(float) * pairs * num_cls, stream));
  associate_gt_kernel<<<(int)ceil_div64(pairs, 128), 128, 0, stream>>>((const float4*)boxes, t, pairs, num_cls, gt_off,
                                                                      (const float4*)gt_boxes_h, (const float4*)gt_boxes_o,
                                                                      gt_verb, fg_iou_thresh, labels);
  return check_launch("associate_gt_kernel");
}

extern "C" int skg_post_offsets(const int32_t* verb_off, const int64_t* labels, const float* scores, const int32_t* box_off,
                                const int32_t* hum_off, int num_images, int32_t* box_cnt, int32_t* box_prefix,
                                int64_t* hum_base, int64_t* img_entries, int64_t* img_base, cudaStream_t stream) {
  if (num_images == 0) return SKG_OK;
  SKG_REQUIRE(verb_off && labels && scores && box_off && hum_off && box_cnt && box_prefix && hum_base && img_entries && img_base,
              "null pointer argument");
  if (num_images <= 4 * (kPostOffThreads / 32)) {     // a batch of images: one single-CTA launch does both steps
    post_offsets_kernel<<<1, kPostOffThreads, 0, stream>>>(verb_off, labels, scores, box_off, hum_off, num_images, box_cnt,
                                                         box_prefix, hum_base, img_entries, img_base);
    return check_launch("post_offsets_kernel");
  }
  post_counts_kernel<<<num_images, 32, 0, stream>>>(verb_off, labels, scores, box_off, hum_off, box_cnt, box_prefix, hum_base,
                                                    img_entries);
  int rc = check_launch("post_counts_kernel");
  if (rc) return rc;
  post_bases_kernel<<<1, 256, 0, stream>>>(hum_off, num_images, img_entries, img_base, hum_base);
  return check_launch("post_bases_kernel");
}

extern "C" int skg_gather_pair_boxes(const float* boxes, const skg_post_tables* tables, int64_t pairs, float* boxes_h,
                                     float* boxes_o, cudaStream_t stream) {
  if (pairs == 0) return SKG_OK;
  PostTables t;
  int rc = fill_tables(t, tables);
  if (rc) return rc;
  SKG_REQUIRE(boxes && boxes_h && boxes_o, "null pointer argument");
  SKG_REQUIRE(((reinterpret_cast<uintptr_t>(boxes) | reinterpret_cast<uintptr_t>(boxes_h) | reinterpret_cast<uintptr_t>(boxes_o)) & 15u) == 0,
              "boxes, boxes_h and boxes_o must be 16-byte aligned (rows are read / written as float4)");
  gather_pair_boxes_kernel<<<(int)ceil_div64(pairs, 256), 256, 0, stream>>>((const float4*)boxes, t, pairs, (float4*)boxes_h,
                                                                           (float4*)boxes_o);
  return check_launch("gather_pair_boxes_kernel");
}

// The whole post-processing of a batch as one host call: output offsets (device), per-entry / per-pair outputs, pair boxes.
// scratch_i32 [2 * boxes], scratch_i64 [humans + 2 * B + 1] are caller-owned (the forward workspace has room for them).
extern "C" int skg_head_post(const float* logits, int ld_logits, int num_cls, const skg_post_tables* tables, const float* boxes,
                             int num_images, int total_boxes, int total_humans, int64_t pairs, float prior_power,
                             const float* gt_labels, int64_t total_entries, int32_t* scratch_i32, int64_t* scratch_i64,
                             int64_t* out_index, int64_t* out_pred, float* out_scores, float* out_prior, float* out_weights,
                             float* out_labels, float* out_unary, int64_t* out_object, float* boxes_h, float* boxes_o,
                             cudaStream_t stream) {
  if (pairs == 0 || num_images == 0) return SKG_OK;
  SKG_REQUIRE(tables && scratch_i32 && scratch_i64, "null pointer argument");
  skg_post_tables t = *tables;
  int32_t* box_cnt = scratch_i32; int32_t* box_prefix = scratch_i32 + total_boxes;
  int64_t* hum_base = scratch_i64; int64_t* img_entries = scratch_i64 + total_humans; int64_t* img_base = img_entries + num_images;
  int rc = skg_post_offsets(t.verb_off, t.labels, t.scores, t.box_off, t.hum_off, num_images, box_cnt, box_prefix, hum_base,
                            img_entries, img_base, stream);
  if (rc) return rc;
  t.box_cnt = box_cnt; t.box_prefix = box_prefix; t.hum_base = hum_base;
  const bool with_boxes = boxes && boxes_h && boxes_o;
  if (with_boxes)
    SKG_REQUIRE((((uintptr_t)boxes | (uintptr_t)boxes_h | (uintptr_t)boxes_o) & 15u) == 0, "boxes / boxes_h / boxes_o must be 16-byte aligned");
  // the pair boxes ride along in the per-pair warp of the post-processing kernel (one launch less on the critical path)
  return postprocess_launch(logits, ld_logits, num_cls, &t, pairs, prior_power, gt_labels, total_entries, out_index, out_pred,
                            out_scores, out_prior, out_weights, out_labels, out_unary, out_object, with_boxes ? boxes : nullptr,
                            boxes_h, boxes_o, stream);
}
