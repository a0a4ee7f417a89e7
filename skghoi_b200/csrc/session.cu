// Per-head serving session: the whole inference step of the interaction head behind TWO host calls ("an opaque
// handle owning pre-packed bf16 weights + TMA descriptors").  Reference: InteractionHead.forward from
// the pooled box features on (head.py:389-427) = GraphHead.forward (head.py:769-993) + predictor / suppressor (head.py:408-411)
// + postprocess (head.py:291-337).
//
//   skg_session_prepare  pure host work on the step's tiny detection lists (labels, scores, boxes per image): humans-first
//                        check (head.py:840-841), the packed segment table (written into a pinned staging slot), the capacity
//                        bucket and its workspace plan (cached per bucket), the closed-form result sizes (entries per image)
//   skg_session_run      ONE H2D copy of the segment table (+ the step's input addresses) into the workspace, the captured
//                        CUDA graph of the bucket (SKG_HEAD_GRAPH | SKG_HEAD_CAPACITY: whatever the batch geometry and the
//                        input addresses are, csrc/forward.cu), the two post-processing launches into the caller's flat outputs
//
// The session owns: a copy of the packed-weight descriptor, the verb-table pointers, the pinned staging ring (one event per
// slot), the plans of the buckets it has seen.  The graphs (with their tensor maps baked in) live in forward.cu's per-thread
// cache, keyed on (weights, bucket, workspace); plain launches take their tensor maps from gemm.cu's descriptor cache.
#include "common.cuh"
#include "skghoi_b200.h"
#include "head_plan.cuh"
#include <string.h>
#include <map>
#include <new>
#include <vector>

namespace skg {

struct SessionStep {
  bool valid = false;
  int slot = 0;
  int B = 0;
  skg_head_batch exact, cap;
  const skg_head_plan_t* plan = nullptr;   // of the capacity bucket
  int64_t offs[7];
  int64_t seg_bytes = 0;
  int64_t entries = 0;
};

struct BucketKey {
  int32_t B, boxes, humans; int64_t cells, seg_bytes;
  bool operator<(const BucketKey& o) const { return memcmp(this, &o, sizeof(*this)) < 0; }
};

}  // namespace skg

struct skg_session {
  skg_head_weights w;
  std::vector<int32_t> verb_len;
  const int32_t* verb_off_dev; const int32_t* verb_idx_dev;
  int num_obj_cls, max_images, flags;
  float prior_power;
  static constexpr int kSlots = 8;
  uint8_t* staging = nullptr;            // pinned, kSlots x slot_bytes
  int64_t slot_bytes = 0;
  cudaEvent_t slot_ev[kSlots] = {};
  bool slot_used[kSlots] = {};
  int next_slot = 0;
  std::vector<int64_t> n_h;              // scratch of prepare
  std::map<skg::BucketKey, skg_head_plan_t> plans;
  skg::SessionStep step;
};

using namespace skg;

extern "C" int skg_session_create(const skg_session_config* cfg, skg_session** out) {
  SKG_REQUIRE(cfg && out, "null pointer argument");
  SKG_REQUIRE(cfg->verb_len && cfg->verb_off_dev && cfg->verb_idx_dev && cfg->num_obj_cls > 0, "missing verb table");
  SKG_REQUIRE(cfg->max_images > 0, "max_images must be positive");
  skg_session* s = new (std::nothrow) skg_session();
  SKG_REQUIRE(s, "out of memory");
  s->w = cfg->weights;
  s->verb_len.assign(cfg->verb_len, cfg->verb_len + cfg->num_obj_cls);
  s->verb_off_dev = cfg->verb_off_dev; s->verb_idx_dev = cfg->verb_idx_dev;
  s->num_obj_cls = cfg->num_obj_cls; s->max_images = cfg->max_images;
  s->flags = cfg->flags & SKG_HEAD_PRECISE_LOGITS;
  s->prior_power = cfg->prior_power;
  s->slot_bytes = ((int64_t)40 * cfg->max_images + 24 + 15) / 16 * 16 + 32;
  s->slot_bytes = (s->slot_bytes + 255) / 256 * 256;
  if (cudaHostAlloc(reinterpret_cast<void**>(&s->staging), (size_t)s->slot_bytes * skg_session::kSlots, cudaHostAllocDefault) != cudaSuccess) {
    cudaGetLastError();
    delete s;
    set_error("cudaHostAlloc of the staging ring failed");
    return SKG_ERR_CUDA;
  }
  for (auto& e : s->slot_ev) {
    if (cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess) {
      cudaGetLastError();
      skg_session_destroy(s);
      set_error("cudaEventCreate failed");
      return SKG_ERR_CUDA;
    }
  }
  s->n_h.resize(cfg->max_images);
  *out = s;
  return SKG_OK;
}

extern "C" void skg_session_destroy(skg_session* s) {
  if (!s) return;
  for (auto& e : s->slot_ev) if (e) cudaEventDestroy(e);
  if (s->staging) cudaFreeHost(s->staging);
  delete s;
}

extern "C" int skg_session_prepare(skg_session* s, int num_images, const int64_t* n, const int64_t* labels, const float* scores,
                                   const double* hw, int feat_hw, skg_step_sizes* sizes, int64_t* pairs_per_image,
                                   int64_t* entries_per_image) {
  SKG_REQUIRE(s && sizes, "null pointer argument");
  SKG_REQUIRE(num_images >= 0 && num_images <= s->max_images, "num_images %d exceeds the session's max_images %d", num_images, s->max_images);
  SKG_REQUIRE(num_images == 0 || (n && labels && scores && hw), "null pointer argument");
  s->step.valid = false;
  memset(sizes, 0, sizeof(*sizes));
  const int B = num_images;
  // humans first (head.py:826-841): n_h = the leading run of human labels, no human label after it
  int64_t off = 0;
  for (int b = 0; b < B; ++b) {
    SKG_REQUIRE(n[b] >= 0, "negative box count");
    SKG_REQUIRE(n[b] <= s->w.ent_tot, "index out of range in self: image %d has %lld boxes, the entity table %d rows (head.py:572)", b,
                (long long)n[b], s->w.ent_tot);
    int64_t h = 0;
    while (h < n[b] && labels[off + h] == s->w.human_idx) ++h;
    for (int64_t k = h; k < n[b]; ++k)
      SKG_REQUIRE(labels[off + k] != s->w.human_idx, "Human detections are not permuted to the top");
    s->n_h[b] = h;
    off += n[b];
  }
  // staging slot: wait for the copy that last read it (normally long finished)
  const int slot = s->next_slot;
  s->next_slot = (slot + 1) % skg_session::kSlots;
  if (s->slot_used[slot]) SKG_CUDA(cudaEventSynchronize(s->slot_ev[slot]));
  uint8_t* host = s->staging + (int64_t)slot * s->slot_bytes;
  SessionStep& st = s->step;
  int64_t totals[5] = {0, 0, 0, 0, 0};
  const int64_t used = skg_segments_pack(n, s->n_h.data(), hw, B, host, s->slot_bytes, st.offs, totals);
  if (used < 0) return SKG_ERR_INVALID;
  memset(&st.exact, 0, sizeof(st.exact));
  st.exact.num_images = B; st.exact.total_boxes = (int32_t)totals[0]; st.exact.total_humans = (int32_t)totals[1];
  st.exact.total_cells = totals[2]; st.exact.total_pairs = totals[3]; st.exact.max_boxes_per_image = (int32_t)totals[4];
  st.exact.feat_hw = feat_hw; st.exact.seg_bytes = used;
  SKG_TRY(skg_head_capacity(&s->w, &st.exact, &st.cap));
  BucketKey key;
  memset(&key, 0, sizeof(key));
  key.B = B; key.boxes = st.cap.total_boxes; key.humans = st.cap.total_humans; key.cells = st.cap.total_cells; key.seg_bytes = used;
  auto it = s->plans.find(key);
  if (it == s->plans.end()) {
    skg_head_plan_t plan;
    SKG_TRY(make_plan(s->w, st.cap, s->flags | SKG_HEAD_CAPACITY, &plan));
    if (s->plans.size() > 512) s->plans.clear();
    it = s->plans.emplace(key, plan).first;
  }
  st.plan = &it->second;
  // result sizes: closed form on the host (the effective n_h of skipped images is what skg_segments_pack wrote)
  const int32_t* nh_eff = reinterpret_cast<const int32_t*>(host + st.offs[3]);
  int64_t entries = 0;
  off = 0;
  for (int b = 0; b < B; ++b) {
    int64_t S = 0, e = 0;
    for (int64_t k = 0; k < n[b]; ++k) {
      const int64_t l = labels[off + k];
      SKG_REQUIRE(l >= 0 && l < s->num_obj_cls, "object class %lld out of range", (long long)l);
      S += s->verb_len[l];
    }
    for (int64_t i = 0; i < nh_eff[b]; ++i)
      if (scores[off + i] > 0.f) e += S - s->verb_len[labels[off + i]];
    if (entries_per_image) entries_per_image[b] = e;
    if (pairs_per_image) pairs_per_image[b] = (int64_t)nh_eff[b] * (n[b] > 0 ? n[b] - 1 : 0);
    entries += e;
    off += n[b];
  }
  st.slot = slot; st.B = B; st.seg_bytes = used; st.entries = entries; st.valid = true;
  sizes->boxes = totals[0]; sizes->humans = totals[1]; sizes->cells = totals[2]; sizes->pairs = totals[3]; sizes->entries = entries;
  sizes->cap_boxes = st.cap.total_boxes; sizes->cap_humans = st.cap.total_humans; sizes->cap_cells = st.cap.total_cells;
  sizes->workspace_bytes = st.plan->total_bytes;
  sizes->out_i64 = 2 * entries + totals[3];
  sizes->out_f32 = 9 * totals[3] + 3 * entries;
  return SKG_OK;
}

extern "C" int skg_session_plans(const skg_session* s, skg_head_plan_t* capacity_plan, skg_head_plan_t* exact_plan) {
  SKG_REQUIRE(s && s->step.plan, "no prepared step");
  if (capacity_plan) *capacity_plan = *s->step.plan;
  if (exact_plan) SKG_TRY(make_plan(s->w, s->step.exact, s->flags, exact_plan));
  return SKG_OK;
}

extern "C" int skg_session_run(skg_session* s, const skg_step_io* io, void* workspace, int64_t workspace_bytes, cudaStream_t stream) {
  SKG_REQUIRE(s && io && workspace, "null pointer argument");
  SessionStep& st = s->step;
  SKG_REQUIRE(st.valid, "skg_session_run without skg_session_prepare");
  st.valid = false;
  const skg_head_plan_t& plan = *st.plan;
  SKG_REQUIRE(workspace_bytes >= plan.total_bytes, "workspace too small: %lld < %lld bytes", (long long)workspace_bytes, (long long)plan.total_bytes);
  const int64_t P = st.exact.total_pairs, L = st.entries;
  if (st.B == 0 || st.exact.total_cells == 0) return SKG_OK;
  SKG_REQUIRE(io->features3 && (io->box_features || io->box_features_bf16) && io->boxes && io->labels && io->scores && io->ent,
              "null input pointer");
  SKG_REQUIRE(io->out_i64 && io->out_f32, "null output pointer");
  uint8_t* base = reinterpret_cast<uint8_t*>(workspace);
  uint8_t* host = s->staging + (int64_t)st.slot * s->slot_bytes;
  unsigned long long* addr = reinterpret_cast<unsigned long long*>(host + st.offs[6]);
  addr[0] = reinterpret_cast<unsigned long long>(io->features3);
  addr[1] = reinterpret_cast<unsigned long long>(io->box_features);
  addr[2] = reinterpret_cast<unsigned long long>(io->boxes);
  addr[3] = 0;
  uint8_t* seg = base + plan.offset[B_SEG];
  SKG_CUDA(cudaMemcpyAsync(seg, host, (size_t)st.seg_bytes, cudaMemcpyHostToDevice, stream));
  SKG_CUDA(cudaEventRecord(s->slot_ev[st.slot], stream));
  s->slot_used[st.slot] = true;
  skg_head_batch b = st.cap;
  b.cell_off = reinterpret_cast<const int64_t*>(seg + st.offs[0]); b.pair_off = reinterpret_cast<const int64_t*>(seg + st.offs[1]);
  b.box_off = reinterpret_cast<const int32_t*>(seg + st.offs[2]); b.n_h = reinterpret_cast<const int32_t*>(seg + st.offs[3]);
  b.hum_off = reinterpret_cast<const int32_t*>(seg + st.offs[4]); b.hw = reinterpret_cast<const float*>(seg + st.offs[5]);
  b.box_features_bf16 = io->box_features_bf16;
  skg_head_weights w = s->w;
  w.ent = io->ent;
  SKG_TRY(skg_head_forward(&w, &b, s->flags | SKG_HEAD_GRAPH | SKG_HEAD_CAPACITY, workspace, workspace_bytes, stream));
  if (P == 0) return SKG_OK;
  skg_post_tables t;
  memset(&t, 0, sizeof(t));
  t.verb_off = s->verb_off_dev; t.verb_idx = s->verb_idx_dev;
  t.pair2cell = reinterpret_cast<const int32_t*>(base + plan.offset[B_PAIR2CELL]);
  t.cell_hrow = reinterpret_cast<const int32_t*>(base + plan.offset[B_CELL_HROW]);
  t.cell_orow = reinterpret_cast<const int32_t*>(base + plan.offset[B_CELL_OROW]);
  t.cell_img = reinterpret_cast<const int32_t*>(base + plan.offset[B_CELL_IMG]);
  t.hum_off = b.hum_off; t.box_off = b.box_off; t.pair_off = b.pair_off;
  t.labels = io->labels; t.scores = io->scores;
  // flat outputs (the layout engine._flat_f_views reads): int64 [index L | prediction L | object P],
  // fp32 [boxes_h 4P | boxes_o 4P | scores L | prior 2L | weights P]
  int64_t* o_index = io->out_i64; int64_t* o_pred = o_index + L; int64_t* o_obj = o_pred + L;
  float* o_bh = io->out_f32; float* o_bo = o_bh + 4 * P; float* o_scores = o_bo + 4 * P; float* o_prior = o_scores + L;
  float* o_weights = o_prior + 2 * L;
  return skg_head_post(reinterpret_cast<const float*>(base + plan.offset[B_LOGITS]), s->w.n_logit, s->w.num_cls, &t, io->boxes, st.B,
                       st.exact.total_boxes, st.exact.total_humans, P, s->prior_power, nullptr, L,
                       reinterpret_cast<int32_t*>(base + plan.offset[B_POST32]), reinterpret_cast<int64_t*>(base + plan.offset[B_POST64]),
                       o_index, o_pred, o_scores, o_prior, o_weights, nullptr, nullptr, o_obj, o_bh, o_bo, stream);
}
