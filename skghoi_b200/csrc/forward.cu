// The whole pair path of the interaction head as ONE host call (reference GraphHead.forward head.py:769-993 and
// the predictor / suppressor of InteractionHead.forward head.py:408-411), batched over every image of the step.
//
// Host-side orchestration only: it carves the caller's workspace, builds the epilogue descriptors and enqueues
// the kernels of spatial.cu / headops.cu / gemm.cu.  Independent chains run on a second (library-owned,
// per-thread) stream forked from and joined back into the caller's stream with events:
//     main : pair enumeration + spatial encoding -> spatial_head (3 layers)
//     side : box_head -> fc_head / fc_tail node rows -> per-node fc_1 projections ; global pooling -> fc_1 of attention_head_g
//     join : fused fc_2 GEMM of the four MBFs (the dominant launch)
//     main : adjacency row-dot GEMM -> softmax message sums -> message fc_3 + residual -> LayerNorm -> final attention head
//     side : attention_head_g fc_3 on the valid pairs (independent of the message passing)
//     join : predictor + suppressor logits
// so the few-CTA node-row GEMMs overlap the wide cell-row ones instead of leaving 140 SMs idle.
#include "common.cuh"
#include "skghoi_b200.h"
#include "head_plan.cuh"
#include <string.h>

namespace skg {

// library-owned side streams + an event pool for fork / join (per host thread and device)
struct SideStream {
  static constexpr int kStreams = 4, kEvents = 24;   // s[3] stands in for the legacy default stream (cannot be captured)
  int dev = -1;
  cudaStream_t s[kStreams] = {};
  cudaEvent_t ev[kEvents] = {};
  int next = 0;
};
static thread_local SideStream g_side;

static int get_side(SideStream** out) {
  int dev = 0;
  SKG_CUDA(cudaGetDevice(&dev));
  if (g_side.dev != dev) {
    if (g_side.dev >= 0) {   // device changed on this thread: drop the old objects (rare)
      for (auto& x : g_side.s) if (x) cudaStreamDestroy(x);
      for (auto& e : g_side.ev) if (e) cudaEventDestroy(e);
      g_side = SideStream();
    }
    for (auto& x : g_side.s) SKG_CUDA(cudaStreamCreateWithFlags(&x, cudaStreamNonBlocking));
    for (auto& e : g_side.ev) SKG_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    g_side.dev = dev;
  }
  g_side.next = 2;                                    // ev[0], ev[1] belong to the graph wrapper's fork / join
  *out = &g_side;
  return SKG_OK;
}

// `to` waits for everything enqueued on `from` so far
static int stream_dep(SideStream* sd, cudaStream_t from, cudaStream_t to) {
  SKG_REQUIRE(sd->next < SideStream::kEvents, "event pool exhausted");
  cudaEvent_t e = sd->ev[sd->next++];
  SKG_CUDA(cudaEventRecord(e, from));
  SKG_CUDA(cudaStreamWaitEvent(to, e, 0));
  return SKG_OK;
}

// optional per-launch timing (SKG_HEAD_PROFILE): one event after every launch, read back after a stream sync
struct Prof {
  static constexpr int kMax = 96;
  bool on = false;
  int n = 0;
  cudaEvent_t ev[kMax] = {};
  const char* name[kMax] = {};
  int sid[kMax] = {};
  float us[kMax] = {};
};
static thread_local Prof g_prof;

static void prof_mark(const char* name, int sid, cudaStream_t s) {
  Prof& p = g_prof;
  if (!p.on || p.n >= Prof::kMax) return;
  if (!p.ev[p.n]) cudaEventCreate(&p.ev[p.n]);
  cudaEventRecord(p.ev[p.n], s);
  p.name[p.n] = name; p.sid[p.n] = sid; p.n++;
}
#define MARK(name, s) prof_mark(name, (s) == st ? 0 : (s) == sa ? 1 : (s) == sb ? 2 : 3, (s))

}  // namespace skg

using namespace skg;

extern "C" const char* skg_head_buffer_name(int i) { return (i >= 0 && i < B_COUNT) ? kBufNames[i] : nullptr; }

extern "C" int skg_head_plan(const skg_head_weights* weights, const skg_head_batch* batch, int flags, skg_head_plan_t* plan) {
  SKG_REQUIRE(weights && batch && plan, "null pointer argument");
  return make_plan(*weights, *batch, flags, plan);
}

static int enqueue_forward(const skg_head_weights* wp, const skg_head_batch* bp, int flags, void* workspace,
                           int64_t workspace_bytes, cudaStream_t st);

// ---- CUDA-graph cache -------------------------------------------------------------------------------------------
// A step is ~35 small launches on four streams; replaying it as a graph removes the per-launch host cost and the
// inter-kernel launch gaps.  Key = every byte that reaches a launch (weights struct, batch struct incl. the input
// pointers, flags, workspace): the second call with an identical key captures the graph, later calls replay it.
// Any difference (new shapes, inputs at other addresses) falls back to plain stream launches.
struct GraphEntry {
  skg_head_weights w; skg_head_batch b; int flags; void* ws; int64_t ws_bytes; int dev;
  int seen = 0;
  cudaGraphExec_t exec = nullptr;
  long long launches = 0;
  unsigned long long stamp = 0;
};
static thread_local GraphEntry g_graphs[64];
static thread_local unsigned long long g_graph_clock = 0;
static thread_local long long g_graph_stats[3] = {0, 0, 0};   // replays, captures, plain launches

static bool same_key(const GraphEntry& e, const skg_head_weights& w, const skg_head_batch& b, int flags, void* ws, int64_t wsb, int dev) {
  return e.seen > 0 && e.flags == flags && e.ws == ws && e.ws_bytes == wsb && e.dev == dev && memcmp(&e.w, &w, sizeof(w)) == 0 &&
         memcmp(&e.b, &b, sizeof(b)) == 0;
}

extern "C" int skg_head_forward(const skg_head_weights* wp, const skg_head_batch* bp, int flags, void* workspace,
                                int64_t workspace_bytes, cudaStream_t st) {
  SKG_REQUIRE(wp && bp && workspace, "null pointer argument");
  if (!(flags & SKG_HEAD_GRAPH) || (flags & SKG_HEAD_PROFILE)) return enqueue_forward(wp, bp, flags, workspace, workspace_bytes, st);
  cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(st, &cs) != cudaSuccess || cs != cudaStreamCaptureStatusNone)
    return enqueue_forward(wp, bp, flags, workspace, workspace_bytes, st);   // the caller is capturing: just enqueue
  int dev = 0;
  SKG_CUDA(cudaGetDevice(&dev));
  if (st == nullptr || st == cudaStreamLegacy) {
    // the legacy default stream cannot be captured: run the step on a library stream forked from / joined into it
    SideStream* sd = nullptr;
    SKG_TRY(get_side(&sd));
    cudaStream_t run = sd->s[3];
    cudaEvent_t e0 = sd->ev[0], e1 = sd->ev[1];
    SKG_CUDA(cudaEventRecord(e0, st));
    SKG_CUDA(cudaStreamWaitEvent(run, e0, 0));
    const int rc = skg_head_forward(wp, bp, flags, workspace, workspace_bytes, run);
    SKG_CUDA(cudaEventRecord(e1, run));
    SKG_CUDA(cudaStreamWaitEvent(st, e1, 0));
    return rc;
  }
  GraphEntry* hit = nullptr; GraphEntry* victim = &g_graphs[0];
  for (auto& e : g_graphs) {
    if (same_key(e, *wp, *bp, flags, workspace, workspace_bytes, dev)) { hit = &e; break; }
    if (e.stamp < victim->stamp) victim = &e;
  }
  if (hit && hit->exec) {
    hit->stamp = ++g_graph_clock;
    SKG_CUDA(cudaGraphLaunch(hit->exec, st));
    add_launches(hit->launches);
    g_graph_stats[0]++;
    return SKG_OK;
  }
  if (!hit) {   // first sighting: plain launches (also sets the kernels' function attributes outside any capture)
    if (victim->exec) { cudaGraphExecDestroy(victim->exec); victim->exec = nullptr; }
    victim->w = *wp; victim->b = *bp; victim->flags = flags; victim->ws = workspace; victim->ws_bytes = workspace_bytes;
    victim->dev = dev; victim->seen = 1; victim->stamp = ++g_graph_clock;
    g_graph_stats[2]++;
    return enqueue_forward(wp, bp, flags, workspace, workspace_bytes, st);
  }
  // second sighting: capture, instantiate, launch
  hit->stamp = ++g_graph_clock;
  const long long l0 = skg_launch_count();
  if (cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal) != cudaSuccess) {
    cudaGetLastError();
    return enqueue_forward(wp, bp, flags, workspace, workspace_bytes, st);
  }
  const int rc = enqueue_forward(wp, bp, flags, workspace, workspace_bytes, st);
  cudaGraph_t graph = nullptr;
  const cudaError_t ce = cudaStreamEndCapture(st, &graph);
  const long long captured = skg_launch_count() - l0;
  add_launches(-captured);                       // nothing ran yet: count launches when the graph is launched
  if (rc != SKG_OK || ce != cudaSuccess || !graph) {
    cudaGetLastError();
    if (graph) cudaGraphDestroy(graph);
    hit->seen = 0;
    if (rc != SKG_OK) return rc;
    return enqueue_forward(wp, bp, flags, workspace, workspace_bytes, st);
  }
  cudaGraphExec_t exec = nullptr;
  const cudaError_t ie = cudaGraphInstantiate(&exec, graph, 0);
  cudaGraphDestroy(graph);
  if (ie != cudaSuccess || !exec) {
    cudaGetLastError();
    hit->seen = 0;
    return enqueue_forward(wp, bp, flags, workspace, workspace_bytes, st);
  }
  hit->exec = exec; hit->launches = captured;
  g_graph_stats[1]++;
  SKG_CUDA(cudaGraphLaunch(exec, st));
  add_launches(captured);
  return SKG_OK;
}

// byte offset of the input-address block inside the packed segment table (skg_segments_pack layout)
static inline int64_t seg_ptr_offset(int64_t B) {
  const int64_t o_hw = 8 * (B + 1) * 2 + 4 * (B + 1) + 4 * B + 4 * (B + 1);
  return (o_hw + 12 * B + 15) / 16 * 16;
}

extern "C" int skg_head_capacity(const skg_head_weights* wp, const skg_head_batch* exact, skg_head_batch* cap) {
  SKG_REQUIRE(wp && exact && cap, "null pointer argument");
  *cap = *exact;
  auto up = [](int64_t v, int64_t g) { return v <= 0 ? g : (v + g - 1) / g * g; };
  // cells: one GEMM M tile (128 rows) up to 2048, then <= 6.25 % slack; node rows: one M tile
  const int64_t C = exact->total_cells;
  int64_t g = 128;
  while (g * 16 < C) g *= 2;
  cap->total_cells = up(C, g);
  cap->total_pairs = cap->total_cells;
  cap->total_boxes = (int32_t)up(exact->total_boxes, 128);
  cap->total_humans = (int32_t)up(exact->total_humans, 128);
  cap->max_boxes_per_image = wp->ent_tot;
  cap->features3 = nullptr; cap->box_features = nullptr; cap->boxes = nullptr;   // read from the segment table's address block
  return SKG_OK;
}

static int enqueue_forward(const skg_head_weights* wp, const skg_head_batch* bp, int flags, void* workspace,
                           int64_t workspace_bytes, cudaStream_t st) {
  const skg_head_weights& w = *wp;
  const skg_head_batch& b = *bp;
  skg_head_plan_t plan;
  SKG_TRY(make_plan(w, b, flags, &plan));
  SKG_REQUIRE(workspace_bytes >= plan.total_bytes, "workspace too small: %lld < %lld bytes", (long long)workspace_bytes,
              (long long)plan.total_bytes);
  SKG_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 255u) == 0, "workspace must be 256-byte aligned");
  const int64_t C = b.total_cells, P = b.total_pairs;
  const int NB = b.total_boxes, NH = b.total_humans, B = b.num_images, R = w.R, E = w.E;
  if (C == 0 || B == 0) return SKG_OK;
  SKG_REQUIRE(C < (int64_t)1 << 31, "too many grid cells in one batch");
  const bool cap = (flags & SKG_HEAD_CAPACITY) != 0;
  SKG_REQUIRE(b.box_off && b.n_h && b.hum_off && b.cell_off && b.pair_off && b.hw, "null batch pointer");
  SKG_REQUIRE(cap || (b.features3 && (b.box_features || b.box_features_bf16) && b.boxes), "null batch pointer");
  SKG_REQUIRE(!cap || !(flags & (SKG_HEAD_TRAIN | SKG_HEAD_WANT_TAPS)), "SKG_HEAD_CAPACITY is an inference mode (no TRAIN / WANT_TAPS)");
  SKG_REQUIRE(!cap || b.seg_bytes >= seg_ptr_offset(B) + 32, "SKG_HEAD_CAPACITY needs the packed segment table inside the workspace");
  SKG_REQUIRE(b.max_boxes_per_image <= w.ent_tot, "index out of range in self: a box index addresses the %d-row entity table (head.py:572)",
              w.ent_tot);
  const bool train = (flags & SKG_HEAD_TRAIN) != 0;
  const bool precise = !train && (flags & SKG_HEAD_PRECISE_LOGITS) != 0 && w.Wpred_hh && w.Wpred_lo;
  SKG_REQUIRE(precise || w.Wpred, "missing predictor weights");
  uint8_t* base = reinterpret_cast<uint8_t*>(workspace);
  // capacity mode: the batch sizes above are CAPACITIES; the true counts (box_off[B], cell_off[B], ...) and the three input
  // addresses live in the segment table the caller uploaded into buffer "seg" for THIS step
  const unsigned long long* slots = cap ? reinterpret_cast<const unsigned long long*>(base + plan.offset[B_SEG] + seg_ptr_offset(B)) : nullptr;
  if (cap) {
    SKG_REQUIRE(reinterpret_cast<const uint8_t*>(b.cell_off) == base + plan.offset[B_SEG],
                "SKG_HEAD_CAPACITY: the segment table must sit in workspace buffer \"seg\"");
  }
  auto F = [&](int i) { return reinterpret_cast<float*>(base + plan.offset[i]); };
  auto H = [&](int i) { return reinterpret_cast<void*>(base + plan.offset[i]); };
  auto I32 = [&](int i) { return reinterpret_cast<int32_t*>(base + plan.offset[i]); };
  auto I64 = [&](int i) { return reinterpret_cast<int64_t*>(base + plan.offset[i]); };
  auto HR = [&](int i, int64_t row) { return reinterpret_cast<void*>(base + plan.offset[i] + row * plan.cols[i] * 2); };
  auto FR = [&](int i, int64_t row) { return reinterpret_cast<float*>(base + plan.offset[i] + row * plan.cols[i] * 4); };
  const int64_t sk_bytes = plan.cols[B_SPLITK0] * 4;
  SideStream* side = nullptr;
  SKG_TRY(get_side(&side));
  cudaStream_t sa = side->s[0], sb = side->s[1], sc = side->s[2];
  skg_epilogue_group g;

  g_prof.on = (flags & SKG_HEAD_PROFILE) != 0; g_prof.n = 0;
  MARK("start", st);
  // ================= fork: node-row chain on sa, global-feature chain on sb
  SKG_TRY(stream_dep(side, st, sa));
  SKG_TRY(stream_dep(side, st, sb));
  SKG_TRY(stream_dep(side, st, sc));
  MARK("start", sa); MARK("start", sb); MARK("start", sc);
  // ---- sc: the two small preparations of the fc_head | fc_tail GEMM run beside the box head instead of in front of it
  //      (they were ~15 us at the head of the step's critical path)
  SKG_TRY(skg_node_prep(w.ent, w.ent_dim, w.ent_tot, b.box_off, b.n_h, b.hum_off, B, NB, H(B_OIN), w.fc_k, E, I32(B_BOX2HUM), sc));
  MARK("node_prep", sc);
  if (w.Wfh_ent) {   // constant human-side entity term of fc_head, re-folded from the table's CURRENT contents
    SKG_TRY(skg_fc_bias_fold(w.bfh, w.Wfh_ent, w.ent, w.ent_dim, w.human_idx, R, const_cast<float*>(w.bfc), sc));
    MARK("fc_bias_fold", sc);
  }
  // ---- sa: box head (head.py:635-641, 812); its second layer writes straight into the fc input rows
  const void* xb = b.box_features_bf16;            // pooled features already in operand layout (native RoIAlign)
  if (!xb) {
    if (cap) SKG_TRY(cast_pad_bf16_indirect(slots + 1, w.pool_k, b.box_off + B, NB, w.pool_k, H(B_XB), w.K0, w.K0, sa));
    else SKG_TRY(skg_cast_pad_bf16(b.box_features, w.pool_k, nullptr, 0, NB, w.pool_k, H(B_XB), w.K0, w.K0, sa));
    MARK("cast box_features", sa);
    xb = H(B_XB);
  }
  g = group0(); g.flags = SKG_EPI_RELU; g.bias = w.bb1; g.out_bf16 = H(B_X1); g.ld_ob = E;
  SKG_TRY(skg_gemm_bf16_splitk(xb, w.K0, w.Wb1, w.K0, NB, E, w.K0, &g, 1, 0, H(B_SPLITK1), sk_bytes, sa));
  MARK("gemm box_head.1", sa);
  g = group0(); g.flags = SKG_EPI_RELU; g.bias = w.bb2; g.out_bf16 = H(B_OIN); g.ld_ob = w.fc_k;
  SKG_TRY(skg_gemm_bf16(H(B_X1), E, w.Wb2, E, NB, E, E, &g, 1, 0, sa));
  MARK("gemm box_head.3", sa);
  // ---- sa: fc_head and fc_tail as ONE GEMM over all boxes with stacked weights (head.py:875-885): only n_h (resp. n)
  //      rows of the reference's [N2, 1074] inputs are distinct; human rows are compacted through box2hum
  SKG_TRY(stream_dep(side, sc, sa));   // entity columns, box -> human map and the folded bias are in place
  {
    skg_epilogue_group gg[2];
    gg[0] = group0(); gg[0].flags = SKG_EPI_RELU; gg[0].bias = w.bfc; gg[0].out_f32 = F(B_HN32); gg[0].ld_of = R;
    gg[0].out_bf16 = H(B_HN16); gg[0].ld_ob = R; gg[0].out_row = I32(B_BOX2HUM);
    gg[1] = group0(); gg[1].n_begin = R; gg[1].flags = SKG_EPI_RELU; gg[1].bias = w.bfc; gg[1].out_f32 = F(B_ON32); gg[1].ld_of = R;
    gg[1].out_bf16 = H(B_ON16); gg[1].ld_ob = R;
    SKG_TRY(skg_gemm_bf16(H(B_OIN), w.fc_k, w.Wfc, w.fc_k, NB, 2 * R, w.fc_k, gg, 2, 0, sa));
    MARK("gemm fc_head|fc_tail", sa);
  }
  // ---- per-node halves of the MBF appearance terms: humans -> [U_h | V_so] (sc), boxes -> [U_o | V_os] (sa)
  SKG_TRY(stream_dep(side, sa, sc));
  MARK("start", sc);
  g = group0(); g.bias = w.b_hproj; g.out_f32 = F(B_HPROJ); g.ld_of = 2 * R;
  SKG_TRY(skg_gemm_bf16(H(B_HN16), R, w.W_hproj, R, NH, 2 * R, R, &g, 1, 0, sc));
  MARK("gemm hproj", sc);
  g = group0(); g.bias = w.b_oproj; g.out_f32 = F(B_OPROJ); g.ld_of = 2 * R;
  SKG_TRY(skg_gemm_bf16(H(B_ON16), R, w.W_oproj, R, NB, 2 * R, R, &g, 1, 0, sa));
  MARK("gemm oproj", sa);
  // ---- sb: global feature term of attention_head_g (head.py:811, 971)
  if (w.Kg == w.Cg) {
    if (cap) SKG_TRY(global_pool_indirect(slots, (int64_t)B * w.Cg, b.feat_hw, H(B_GFEAT), nullptr, sb));
    else SKG_TRY(skg_global_pool(b.features3, (int64_t)B * w.Cg, b.feat_hw, H(B_GFEAT), nullptr, sb));
  } else {
    if (cap) SKG_TRY(global_pool_indirect(slots, (int64_t)B * w.Cg, b.feat_hw, nullptr, F(B_G32), sb));
    else SKG_TRY(skg_global_pool(b.features3, (int64_t)B * w.Cg, b.feat_hw, nullptr, F(B_G32), sb));
    SKG_TRY(skg_cast_pad_bf16(F(B_G32), w.Cg, nullptr, 0, B, w.Cg, H(B_GFEAT), w.Kg, w.Kg, sb));
  }
  MARK("global_pool", sb);
  g = group0(); g.bias = w.bg1; g.out_f32 = F(B_GT); g.ld_of = R;
  SKG_TRY(skg_gemm_bf16(H(B_GFEAT), w.Kg, w.Wg1, w.Kg, B, R, w.Kg, &g, 1, 0, sb));
  MARK("gemm attention_head_g.fc_1", sb);

  // ================= main: pairs + spatial head (head.py:847-868, 662-669, 888)
  SKG_TRY(pairs_spatial_encode_impl(cap ? nullptr : b.boxes, cap ? slots + 2 : nullptr, b.box_off, b.n_h, b.cell_off, b.pair_off, b.hum_off,
                                    b.hw, B, C, 23, 1e-10f, 1, F(B_ENC), I64(B_X), I64(B_Y), I64(B_XKEEP), I64(B_YKEEP), I32(B_CELL2PAIR),
                                    I32(B_CELL_HROW), I32(B_CELL_OROW), I32(B_CELL_IMG), I32(B_PAIR2CELL), I32(B_NANFLAG), st));
  MARK("pairs_spatial_encode", st);
  if (b.maps_ready_event) {   // geometry-only work of the caller (labels, normalisers, TransH selection) may start on another stream
    cudaStreamCaptureStatus cst = cudaStreamCaptureStatusNone;
    SKG_CUDA(cudaStreamIsCapturing(st, &cst));
    if (cst != cudaStreamCaptureStatusNone)
      SKG_CUDA(cudaEventRecordWithFlags(reinterpret_cast<cudaEvent_t>(b.maps_ready_event), st, cudaEventRecordExternal));
    else
      SKG_CUDA(cudaEventRecord(reinterpret_cast<cudaEvent_t>(b.maps_ready_event), st));
  }
  SKG_TRY(skg_spatial_fc1(F(B_ENC), w.Wsp0, w.bsp0, C, 46, 128, H(B_S1), st));
  MARK("spatial_fc1", st);
  g = group0(); g.flags = SKG_EPI_RELU; g.bias = w.bsp1; g.out_bf16 = H(B_S2); g.ld_ob = 256;
  SKG_TRY(skg_gemm_bf16(H(B_S1), 128, w.Wsp1, 128, (int)C, 256, 128, &g, 1, 0, st));
  MARK("gemm spatial_head.2", st);
  g = group0(); g.flags = SKG_EPI_RELU; g.bias = w.bsp2; g.out_bf16 = H(B_SP); g.ld_ob = R;
  SKG_TRY(skg_gemm_bf16(H(B_S2), 256, w.Wsp2, 256, (int)C, R, 256, &g, 1, 0, st));
  MARK("gemm spatial_head.4", st);
  // ================= join, then the fused fc_2 GEMM of the four MBFs that read the spatial features (N = 4R)
  SKG_TRY(stream_dep(side, sa, st));
  SKG_TRY(stream_dep(side, sb, st));
  SKG_TRY(stream_dep(side, sc, st));
  {
    skg_epilogue_group gq[4];
    for (int i = 0; i < 4; ++i) { gq[i] = group0(); gq[i].n_begin = i * R; gq[i].flags = SKG_EPI_RELU; gq[i].bias = w.b2cat; gq[i].ld_ob = R; }
    gq[0].mul_p = F(B_HPROJ); gq[0].row_p = I32(B_CELL_HROW); gq[0].ld_p = 2 * R;
    gq[0].mul_q = F(B_OPROJ); gq[0].row_q = I32(B_CELL_OROW); gq[0].ld_q = 2 * R;
    const int ld_raw = (int)plan.cols[B_SRAW];     // R, or 4R in training: the raw fc_2 outputs of all four MBFs are kept
    gq[0].out_bf16 = H(B_HATT); gq[0].raw_bf16 = H(B_SRAW); gq[0].ld_raw = ld_raw;
    if (train)
      for (int i = 1; i < 4; ++i) { gq[i].raw_bf16 = reinterpret_cast<__nv_bfloat16*>(H(B_SRAW)) + i * R; gq[i].ld_raw = ld_raw; }
    gq[1].mul_p = F(B_OPROJ) + R; gq[1].row_p = I32(B_CELL_OROW); gq[1].ld_p = 2 * R; gq[1].out_bf16 = H(B_HOS);
    gq[2].mul_p = F(B_HPROJ) + R; gq[2].row_p = I32(B_CELL_HROW); gq[2].ld_p = 2 * R; gq[2].out_bf16 = H(B_HSO);
    gq[3].mul_p = F(B_GT); gq[3].row_p = I32(B_CELL_IMG); gq[3].ld_p = R; gq[3].out_bf16 = H(B_HG);
    SKG_TRY(skg_gemm_bf16(H(B_SP), R, w.W2cat, R, (int)C, 4 * R, R, gq, 4, 0, st));
    MARK("gemm fused fc_2 x4", st);
  }
  // ================= fork sb: attention_head_g fc_3 on the valid pairs (head.py:971) does not depend on the messages
  const int feat_ld = (int)plan.cols[B_FEAT];
  SKG_TRY(stream_dep(side, st, sb));
  MARK("start", sb);
  g = group0(); g.flags = SKG_EPI_RELU; g.bias = w.b3g; g.out_bf16 = reinterpret_cast<__nv_bfloat16*>(H(B_FEAT)) + R; g.ld_ob = feat_ld;
  g.out_row = train ? nullptr : I32(B_CELL2PAIR);
  if (precise) { g.lo_bf16 = reinterpret_cast<__nv_bfloat16*>(H(B_FEAT)) + 3 * R; g.ld_lo = feat_ld; }
  SKG_TRY(skg_gemm_bf16(H(B_HG), R, w.W3g, R, (int)C, R, R, &g, 1, 0, sb));
  MARK("gemm attention_head_g.fc_3", sb);
  // ---- attention weights -> adjacency (head.py:896-897); the two iterations of head.py:892 are identical
  const int nt = 4 * (R / 256);
  g = group0(); g.flags = SKG_EPI_RELU; g.bias = w.b3att; g.dot_w = w.w_adj; g.dot_out = F(B_DOTPART); g.ld_dot = nt;
  if (train) { g.out_bf16 = H(B_WTS); g.ld_ob = R; }      // the backward needs the ReLU mask and the values (dw_adj)
  SKG_TRY(skg_gemm_bf16(H(B_HATT), R, w.W3att, R, (int)C, R, R, &g, 1, 256, st));
  MARK("gemm attention_head.fc_3 + adjacency dot", st);
  // ---- messages (head.py:907-925)
  SKG_TRY(skg_message_aggregate(F(B_DOTPART), nt, w.b_adj, H(B_HOS), H(B_HSO), R, b.box_off, b.hum_off, b.cell_off, B, NB, NH,
                                b.max_boxes_per_image, R, HR(B_Z, 0), HR(B_Z, NH), F(B_ADJ), st));
  MARK("message_aggregate", st);
  SKG_TRY(stream_dep(side, st, sc));
  MARK("start", sc);
  g = group0(); g.flags = SKG_EPI_RELU; g.bias = w.b3os; g.residual = F(B_HN32); g.ld_r = R; g.out_f32 = FR(B_PRE, 0); g.ld_of = R;
  SKG_TRY(skg_gemm_bf16(HR(B_Z, 0), R, w.W3os, R, NH, R, R, &g, 1, 0, sc));
  MARK("gemm obj_to_sub.fc_3", sc);
  g = group0(); g.flags = SKG_EPI_RELU; g.bias = w.b3so; g.residual = F(B_ON32); g.ld_r = R; g.out_f32 = FR(B_PRE, NH); g.ld_of = R;
  SKG_TRY(skg_gemm_bf16(HR(B_Z, NH), R, w.W3so, R, NB, R, R, &g, 1, 0, st));
  MARK("gemm sub_to_obj.fc_3", st);
  SKG_TRY(stream_dep(side, sc, st));
  SKG_TRY(skg_layernorm(F(B_PRE), NH + NB, NH, R, w.ln_h_w, w.ln_h_b, w.ln_o_w, w.ln_o_b, w.ln_eps,
                        (flags & SKG_HEAD_WANT_TAPS) ? F(B_ENC32) : nullptr, H(B_ENC16), st));
  MARK("layernorm", st);
  // ---- final attention head on the valid pairs (head.py:966-970)
  SKG_TRY(stream_dep(side, st, sc));
  MARK("start", sc);
  g = group0(); g.out_f32 = F(B_U2H); g.ld_of = R;
  SKG_TRY(skg_gemm_bf16(HR(B_ENC16, 0), R, w.W_hproj, R, NH, R, R, &g, 1, 0, sc));
  MARK("gemm U2h", sc);
  g = group0(); g.bias = w.b_oproj; g.out_f32 = F(B_U2O); g.ld_of = R;
  SKG_TRY(skg_gemm_bf16(HR(B_ENC16, NH), R, w.W_oproj, R, NB, R, R, &g, 1, 0, st));
  MARK("gemm U2o", st);
  SKG_TRY(stream_dep(side, sc, st));
  SKG_TRY(skg_hadamard_cells(H(B_SRAW), plan.cols[B_SRAW], F(B_U2H), R, F(B_U2O), R, I32(B_CELL_HROW), I32(B_CELL_OROW), C, R, H(B_H2), R, st));
  MARK("hadamard_cells", st);
  g = group0(); g.flags = SKG_EPI_RELU; g.bias = w.b3att; g.out_bf16 = H(B_FEAT); g.ld_ob = feat_ld; g.out_row = train ? nullptr : I32(B_CELL2PAIR);
  if (precise) { g.lo_bf16 = reinterpret_cast<__nv_bfloat16*>(H(B_FEAT)) + 2 * R; g.ld_lo = feat_ld; }
  SKG_TRY(skg_gemm_bf16(H(B_H2), R, w.W3att, R, (int)C, R, R, &g, 1, 0, st));
  MARK("gemm attention_head.fc_3 (pairs)", st);
  SKG_TRY(stream_dep(side, sb, st));
  // ================= predictor + suppressor (head.py:410-411) as one N = num_cls + 1 GEMM
  const int64_t LR = train ? C : P;      // logit rows: pair-compacted, or cell rows in training
  if (LR > 0) {
    g = group0();
    if (precise && w.Wpred_hhl) {
      // split bf16 as ONE launch: logits = [hi | lo | hi] x [W_hi | W_hi | W_lo]^T (3 bf16 passes ~ fp32 product), A = the
      // [hi | lo] feature rows read cyclically along K = 6R; the skinny N = 128 problem fills the SMs through split-K
      g.bias = w.bpred; g.out_f32 = F(B_LOGITS); g.ld_of = w.n_logit;
      SKG_TRY(skg_gemm_bf16_wrapk(H(B_FEAT), feat_ld, 4 * R, w.Wpred_hhl, 6 * R, (int)LR, w.n_logit, 6 * R, &g, 1, 0, H(B_SPLITK0), sk_bytes, st));
      MARK("gemm predictor [hi|lo|hi]x[Whi|Whi|Wlo]", st);
    } else if (precise) {   // two launches: logits = [hi | lo] x [W_hi | W_hi]^T + hi x W_lo^T
      g.bias = w.bpred; g.out_f32 = F(B_TMP); g.ld_of = w.n_logit_pad;
      SKG_TRY(skg_gemm_bf16(H(B_FEAT), feat_ld, w.Wpred_hh, 4 * R, (int)LR, w.n_logit, 4 * R, &g, 1, 0, st));
      MARK("gemm predictor [hi|lo]x[Whi|Whi]", st);
      g = group0(); g.residual = F(B_TMP); g.ld_r = w.n_logit_pad; g.out_f32 = F(B_LOGITS); g.ld_of = w.n_logit;
      SKG_TRY(skg_gemm_bf16(H(B_FEAT), feat_ld, w.Wpred_lo, 2 * R, (int)LR, w.n_logit, 2 * R, &g, 1, 0, st));
      MARK("gemm predictor hi x Wlo", st);
    } else {
      g.bias = w.bpred; g.out_f32 = F(B_LOGITS); g.ld_of = w.n_logit;
      SKG_TRY(skg_gemm_bf16(H(B_FEAT), feat_ld, w.Wpred, 2 * R, (int)LR, w.n_logit, 2 * R, &g, 1, 0, st));
      MARK("gemm predictor", st);
    }
  }
  if (g_prof.on) {   // profiling mode synchronises: per-launch time = gap to the previous event on the same stream
    SKG_CUDA(cudaStreamSynchronize(st));
    SKG_CUDA(cudaStreamSynchronize(sa)); SKG_CUDA(cudaStreamSynchronize(sb)); SKG_CUDA(cudaStreamSynchronize(sc));
    Prof& p = g_prof;
    int last[4] = {-1, -1, -1, -1};
    for (int i = 0; i < p.n; ++i) {
      float ms = 0.f;
      if (last[p.sid[i]] >= 0) cudaEventElapsedTime(&ms, p.ev[last[p.sid[i]]], p.ev[i]);
      p.us[i] = ms * 1e3f;
      last[p.sid[i]] = i;
    }
    p.on = false;
  }
  return SKG_OK;
}

extern "C" void skg_head_graph_stats(long long* replays, long long* captures, long long* plain) {
  if (replays) *replays = g_graph_stats[0];
  if (captures) *captures = g_graph_stats[1];
  if (plain) *plain = g_graph_stats[2];
}

extern "C" int skg_head_profile_get(int i, const char** name, int* stream_id, float* us) {
  const Prof& p = g_prof;
  if (i < 0 || i >= p.n) return -1;
  if (name) *name = p.name[i];
  if (stream_id) *stream_id = p.sid[i];
  if (us) *us = p.us[i];
  return 0;
}

extern "C" int64_t skg_post_count_host(const int32_t* verb_len, int num_obj_cls, const int64_t* labels, const float* scores,
                                       const int64_t* box_off, const int64_t* n_h, int num_images, int64_t* img_entries) {
  if (!verb_len || !labels || !scores || !box_off || !n_h) { set_error("null pointer argument"); return -1; }
  int64_t total = 0;
  for (int b = 0; b < num_images; ++b) {
    const int64_t lo = box_off[b], hi = box_off[b + 1];
    int64_t S = 0;
    for (int64_t k = lo; k < hi; ++k) {
      if (labels[k] < 0 || labels[k] >= num_obj_cls) { set_error("object class %lld out of range", (long long)labels[k]); return -1; }
      S += verb_len[labels[k]];
    }
    int64_t e = 0;
    for (int64_t i = 0; i < n_h[b]; ++i)
      if (scores[lo + i] > 0.f) e += S - verb_len[labels[lo + i]];
    if (img_entries) img_entries[b] = e;
    total += e;
  }
  return total;
}

// Host-side packing of the per-batch segment table (engine.Segments): one buffer, one H2D copy.
//   layout: cell_off i64[B+1] | pair_off i64[B+1] | box_off i32[B+1] | n_h i32[B] | hum_off i32[B+1] | hw f32[3B] | pad to 16
//           | ptrs u64[4] = addresses of features3, box_features (fp32), boxes, 0 -- filled by the caller for SKG_HEAD_CAPACITY
// n_h_eff[b] = 0 for images the reference skips (n_h == 0 or n <= 1, head.py:829-839).  offsets[7] receive the byte offsets
// of the seven arrays, totals[5] = boxes, humans, cells, pairs, max boxes per image.  Returns the bytes used or -1.
extern "C" int64_t skg_segments_pack(const int64_t* n, const int64_t* n_h, const double* hw, int num_images, void* out,
                                     int64_t out_bytes, int64_t* offsets, int64_t* totals) {
  const int64_t B = num_images;
  if (B < 0 || !out || !offsets || !totals || (B > 0 && (!n || !n_h || !hw))) { set_error("null pointer argument"); return -1; }
  const int64_t o_cell = 0, o_pair = o_cell + 8 * (B + 1), o_box = o_pair + 8 * (B + 1), o_nh = o_box + 4 * (B + 1),
                o_hum = o_nh + 4 * B, o_hw = o_hum + 4 * (B + 1);
  const int64_t o_ptr = (o_hw + 12 * B + 15) / 16 * 16;   // input-address block of the capacity mode (4 x u64, zero here)
  const int64_t used = o_ptr + 32;
  if (out_bytes < used) { set_error("segment buffer too small: %lld < %lld", (long long)out_bytes, (long long)used); return -1; }
  uint8_t* base = reinterpret_cast<uint8_t*>(out);
  int64_t* cell_off = reinterpret_cast<int64_t*>(base + o_cell);
  int64_t* pair_off = reinterpret_cast<int64_t*>(base + o_pair);
  int32_t* box_off = reinterpret_cast<int32_t*>(base + o_box);
  int32_t* nh = reinterpret_cast<int32_t*>(base + o_nh);
  int32_t* hum_off = reinterpret_cast<int32_t*>(base + o_hum);
  float* hwf = reinterpret_cast<float*>(base + o_hw);
  int64_t boxes = 0, humans = 0, cells = 0, pairs = 0, max_n = 0;
  for (int64_t b = 0; b < B; ++b) {
    if (n_h[b] < 0 || n_h[b] > n[b]) { set_error("n_h must satisfy 0 <= n_h <= n"); return -1; }
    cell_off[b] = cells; pair_off[b] = pairs; box_off[b] = (int32_t)boxes; hum_off[b] = (int32_t)humans;
    const int64_t h = (n_h[b] == 0 || n[b] <= 1) ? 0 : n_h[b];
    nh[b] = (int32_t)h;
    boxes += n[b]; humans += h; cells += h * n[b]; pairs += h * (n[b] > 0 ? n[b] - 1 : 0);
    if (n[b] > max_n) max_n = n[b];
    hwf[3 * b] = (float)hw[2 * b]; hwf[3 * b + 1] = (float)hw[2 * b + 1]; hwf[3 * b + 2] = (float)(hw[2 * b] * hw[2 * b + 1]);
  }
  if (boxes >= ((int64_t)1 << 31) || cells >= ((int64_t)1 << 31)) { set_error("too many boxes / grid cells in one batch"); return -1; }
  cell_off[B] = cells; pair_off[B] = pairs; box_off[B] = (int32_t)boxes; hum_off[B] = (int32_t)humans;
  for (int64_t i = o_hw + 12 * B; i < used; ++i) base[i] = 0;
  offsets[0] = o_cell; offsets[1] = o_pair; offsets[2] = o_box; offsets[3] = o_nh; offsets[4] = o_hum; offsets[5] = o_hw;
  offsets[6] = o_ptr;
  totals[0] = boxes; totals[1] = humans; totals[2] = cells; totals[3] = pairs; totals[4] = max_n;
  return used;
}
