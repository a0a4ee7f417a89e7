// Backward of the whole pair path as ONE host call (the training step of BASELINE configs[2]; reference: autograd through
// GraphHead.forward head.py:769-993, the predictor / suppressor head.py:408-411 and the two focal losses head.py:153-205).
//
// Host-side orchestration only, on the buffers skg_head_forward(SKG_HEAD_TRAIN) left in the workspace.  With G the gradient
// of a layer's pre-activation (ReLU gate already applied), X its input and W [N, K] its weight:
//     dX = gate_X(G W)    one skg_gemm_bf16 launch: A = G, B = W^T (skg_head_weights_t), the ReLU mask of X in the epilogue
//     dW = G^T X          one launch on M-major copies made by skg_transpose_bf16 (split-K over the row dimension)
//     db = colsum(G)      fused into the transposition of G
// Everything between the GEMMs is one of the kernels of trainops.cu.  Order (dependencies top to bottom):
//   losses -> predictor -> final attention heads (pairs) -> [node side: U2 tables -> LayerNorm -> message fc_3 -> aggregation
//   -> adjacency -> attention weights] -> Hadamard backward of all four MBFs -> fused fc_2 -> spatial head
//   -> node projections -> fc_head / fc_tail -> box_head
#include "common.cuh"
#include "skghoi_b200.h"
#include "head_plan.cuh"
#include "graph_cache.cuh"

namespace skg {

static thread_local GraphCache g_bwd_graphs;

struct BwdCtx {
  uint8_t* base;
  skg_head_plan_t plan;
  void* sk; int64_t sk_bytes;       // split-K scratch (zero on entry, left zero)
  void* tra; void* trb;             // M-major operand scratch
  cudaStream_t st;
};

// dW[N, K] (+)= G[M, N]^T X[x_rows ? x_rows[m] : m, K];  db[N] += colsum(G)
static int wgrad(BwdCtx& c, const void* G, int64_t ldg, int N, const void* X, int64_t ldx, int K, const int32_t* x_rows, int64_t M,
                 float* dW, int64_t ld_dw, float* db, bool accumulate) {
  if (M == 0) return SKG_OK;
  const int64_t ld_t = align_up(M, 64);
  SKG_REQUIRE((int64_t)N * ld_t <= c.plan.cols[B_TRA] && (int64_t)K * ld_t <= c.plan.cols[B_TRB], "transpose scratch too small (N=%d K=%d M=%lld)",
              N, K, (long long)M);
  SKG_TRY(skg_transpose_bf16(G, ldg, M, N, nullptr, c.tra, ld_t, db, c.st));
  SKG_TRY(skg_transpose_bf16(X, ldx, M, K, x_rows, c.trb, ld_t, nullptr, c.st));
  skg_epilogue_group g = group0();
  g.out_f32 = dW; g.ld_of = (int)ld_dw;
  if (accumulate) { g.residual = dW; g.ld_r = (int)ld_dw; }
  return skg_gemm_bf16_splitk(c.tra, ld_t, c.trb, ld_t, N, K, (int)M, &g, 1, 0, c.sk, c.sk_bytes, c.st);
}

// dX[M, K] = gate(G[M, N] W[N, K] (+ residual)):  B operand = W^T stored [K, ld_wt] (N contiguous)
static int dgrad(BwdCtx& c, const void* G, int64_t ldg, int64_t M, int N, const void* WT, int64_t ld_wt, int K, const void* mask,
                 int64_t ld_mask, const float* residual, int64_t ld_r, void* out_bf16, int64_t ld_ob, float* out_f32, int64_t ld_of,
                 const int32_t* out_row) {
  if (M == 0) return SKG_OK;
  skg_epilogue_group g = group0();
  g.mask_bf16 = mask; g.ld_mask = (int)ld_mask;
  g.residual = residual; g.ld_r = (int)ld_r;
  g.out_bf16 = out_bf16; g.ld_ob = (int)ld_ob;
  g.out_f32 = out_f32; g.ld_of = (int)ld_of;
  g.out_row = out_row;
  return skg_gemm_bf16(G, ldg, WT, ld_wt, (int)M, K, N, &g, 1, 0, c.st);
}

}  // namespace skg

using namespace skg;

extern "C" int skg_head_backward(const skg_head_weights* wp, const skg_head_weights_t* wtp, const skg_head_batch* bp,
                                 const skg_head_loss_args* lp, const skg_head_grads* gp, int flags, void* workspace,
                                 int64_t workspace_bytes, cudaStream_t st) {
  return skg_head_backward_staged(wp, wtp, bp, lp, gp, flags, workspace, workspace_bytes, st, nullptr, 0);
}

// Stage events (data-parallel training: the gradient all-reduce overlaps the rest of the backward).  stage_events[k] is
// recorded on `st` when the gradients of group k are FINAL:
//   0: predictor | suppressor, attention_head_g.fc_3 / fc_1, LayerNorms, message fc_3, adjacency, attention_head.fc_3
//   1: the node-side chain -- per-node projections (attention_head.fc_1, sub_to_obj.fc_1, obj_to_sub.fc_1), fc_head | fc_tail,
//      box_head (69 % of the bytes: it is scheduled BEFORE the cell-side chain so that its all-reduce has cover)
//   2: fused fc_2 of the four MBFs (16.8 MB), as soon as its weight gradient is done
//   3: spatial head (1.2 MB: the only group whose all-reduce has no cover -- the end of the backward)
static int backward_enqueue(const skg_head_weights* wp, const skg_head_weights_t* wtp, const skg_head_batch* bp,
                            const skg_head_loss_args* lp, const skg_head_grads* gp, int flags, void* workspace,
                            int64_t workspace_bytes, cudaStream_t st, void* const* stage_events, int num_stage_events, bool capturing);

extern "C" void skg_head_backward_graph_stats(long long* replays, long long* captures, long long* plain) {
  if (replays) *replays = g_bwd_graphs.stats[0];
  if (captures) *captures = g_bwd_graphs.stats[1];
  if (plain) *plain = g_bwd_graphs.stats[2];
}

extern "C" int skg_head_backward_staged(const skg_head_weights* wp, const skg_head_weights_t* wtp, const skg_head_batch* bp,
                                        const skg_head_loss_args* lp, const skg_head_grads* gp, int flags, void* workspace,
                                        int64_t workspace_bytes, cudaStream_t st, void* const* stage_events, int num_stage_events) {
  SKG_REQUIRE(wp && wtp && bp && lp && gp && workspace, "null pointer argument");
  if (!(flags & SKG_HEAD_GRAPH))
    return backward_enqueue(wp, wtp, bp, lp, gp, flags, workspace, workspace_bytes, st, stage_events, num_stage_events, false);
  // SKG_HEAD_GRAPH: the ~130 launches replayed as one graph when every byte that reaches a launch repeats (weights, batch
  // geometry, loss-argument and gradient ADDRESSES, workspace, stage events) -- a training loop over a fixed batch geometry
  // with stable buffers; anything else takes the plain launches
  SKG_REQUIRE(lp->tables, "incomplete loss arguments");
  KeyBuilder kb;
  int dev = 0;
  SKG_CUDA(cudaGetDevice(&dev));
  kb.add(dev); kb.add(*wp); kb.add(*wtp); kb.add(*bp); kb.add(*lp->tables); kb.add(lp->gt_labels); kb.add(lp->n_p); kb.add(lp->g_hoi);
  kb.add(lp->g_int); kb.add(lp->prior_power); kb.add(*gp); kb.add(flags); kb.add(workspace); kb.add(workspace_bytes);
  for (int k = 0; k < num_stage_events; ++k) kb.add(stage_events ? stage_events[k] : nullptr);
  return graph_cached_run(g_bwd_graphs, kb.bytes, st, [&](cudaStream_t s, bool capturing) {
    return backward_enqueue(wp, wtp, bp, lp, gp, flags, workspace, workspace_bytes, s, stage_events, num_stage_events, capturing);
  });
}

static int backward_enqueue(const skg_head_weights* wp, const skg_head_weights_t* wtp, const skg_head_batch* bp,
                            const skg_head_loss_args* lp, const skg_head_grads* gp, int flags, void* workspace,
                            int64_t workspace_bytes, cudaStream_t st, void* const* stage_events, int num_stage_events, bool capturing) {
  auto stage_done = [&](int k) -> int {
    if (stage_events && k < num_stage_events && stage_events[k]) {
      // inside a capture the record must become an EXTERNAL event-record node: it then fires at every replay and streams
      // outside the graph can wait on it
      if (capturing) SKG_CUDA(cudaEventRecordWithFlags(reinterpret_cast<cudaEvent_t>(stage_events[k]), st, cudaEventRecordExternal));
      else SKG_CUDA(cudaEventRecord(reinterpret_cast<cudaEvent_t>(stage_events[k]), st));
    }
    return SKG_OK;
  };
  SKG_REQUIRE(wp && wtp && bp && lp && gp && workspace, "null pointer argument");
  SKG_REQUIRE(flags & SKG_HEAD_TRAIN, "skg_head_backward needs the SKG_HEAD_TRAIN workspace layout");
  const skg_head_weights& w = *wp;
  const skg_head_weights_t& wt = *wtp;
  const skg_head_batch& b = *bp;
  const skg_head_grads& d = *gp;
  BwdCtx c;
  SKG_TRY(make_plan(w, b, flags, &c.plan));
  const skg_head_plan_t& plan = c.plan;
  SKG_REQUIRE(workspace_bytes >= plan.total_bytes, "workspace too small: %lld < %lld bytes", (long long)workspace_bytes, (long long)plan.total_bytes);
  const int64_t C = b.total_cells, P = b.total_pairs;
  const int NB = b.total_boxes, NH = b.total_humans, B = b.num_images, R = w.R, E = w.E;
  if (C == 0 || B == 0) return SKG_OK;
  SKG_REQUIRE(lp->tables && lp->gt_labels && lp->n_p, "incomplete loss arguments");
  SKG_REQUIRE(wt.Wb2T && wt.Wsp1T && wt.Wsp2T && wt.WfcT && wt.W_hprojT && wt.W_oprojT && wt.W2catT && wt.W3attT && wt.W3osT && wt.W3soT &&
                  wt.W3gT && wt.WpredT, "missing transposed weights");
  c.base = reinterpret_cast<uint8_t*>(workspace);
  c.st = st;
  uint8_t* base = c.base;
  auto F = [&](int i) { return reinterpret_cast<float*>(base + plan.offset[i]); };
  auto H = [&](int i) { return reinterpret_cast<__nv_bfloat16*>(base + plan.offset[i]); };
  auto I32 = [&](int i) { return reinterpret_cast<int32_t*>(base + plan.offset[i]); };
  auto bytes = [&](int i) { return (size_t)(plan.rows[i] * plan.cols[i] * plan.elem_size[i]); };
  c.sk = base + plan.offset[B_SPLITK0]; c.sk_bytes = plan.cols[B_SPLITK0] * 4;
  c.tra = H(B_TRA); c.trb = H(B_TRB);
  const int NLP = w.n_logit_pad;
  const int feat_ld = (int)plan.cols[B_FEAT];      // 2R
  const int ld_s = (int)plan.cols[B_SRAW];         // 4R
  const __nv_bfloat16* xb = b.box_features_bf16 ? reinterpret_cast<const __nv_bfloat16*>(b.box_features_bf16) : H(B_XB);

  // ---- zero what is accumulated or only partially written
  SKG_CUDA(cudaMemsetAsync(H(B_DLOGITS), 0, bytes(B_DLOGITS), st));
  SKG_CUDA(cudaMemsetAsync(F(B_DOPROJ), 0, bytes(B_DOPROJ), st));
  SKG_CUDA(cudaMemsetAsync(F(B_DU2O), 0, bytes(B_DU2O), st));
  SKG_CUDA(cudaMemsetAsync(F(B_DGT), 0, bytes(B_DGT), st));
  SKG_CUDA(cudaMemsetAsync(H(B_GFC), 0, bytes(B_GFC), st));
  SKG_TRY(skg_hum2box(I32(B_BOX2HUM), NB, I32(B_HUM2BOX), st));

  // ---- losses -> d logits (cell rows)
  SKG_TRY(skg_head_loss_backward(F(B_LOGITS), w.n_logit, w.num_cls, lp->tables, I32(B_PAIR2CELL), P, lp->prior_power, lp->gt_labels,
                                 lp->n_p, lp->g_hoi, lp->g_int, H(B_DLOGITS), NLP, st));
  // ---- predictor | suppressor (head.py:410-411)
  SKG_TRY(wgrad(c, H(B_DLOGITS), NLP, NLP, H(B_FEAT), feat_ld, 2 * R, nullptr, C, d.Wpred, 2 * R, d.bpred, false));
  SKG_TRY(dgrad(c, H(B_DLOGITS), NLP, C, NLP, wt.WpredT, NLP, 2 * R, H(B_FEAT), feat_ld, nullptr, 0, H(B_GF), 2 * R, nullptr, 0, nullptr));
  // ---- final attention heads on the pairs (head.py:966-973): fc_3 of attention_head (first of two uses) and attention_head_g
  SKG_TRY(wgrad(c, H(B_GF), 2 * R, R, H(B_H2), R, R, nullptr, C, d.W3att, R, d.b3att, false));
  SKG_TRY(wgrad(c, H(B_GF) + R, 2 * R, R, H(B_HG), R, R, nullptr, C, d.W3g, R, d.b3g, false));
  SKG_TRY(dgrad(c, H(B_GF), 2 * R, C, R, wt.W3attT, R, R, H(B_H2), R, nullptr, 0, H(B_DP2), R, nullptr, 0, nullptr));
  SKG_TRY(dgrad(c, H(B_GF) + R, 2 * R, C, R, wt.W3gT, R, R, H(B_HG), R, nullptr, 0, H(B_DPG), R, nullptr, 0, nullptr));
  // ---- branch products of the final pass: node-side sums dU2h / dU2o / dGt and the g column block of dS
  skg_hadamard_bwd hb;
  memset(&hb, 0, sizeof(hb));
  hb.S = H(B_SRAW); hb.ld_s = ld_s;
  hb.dP_att = H(B_DPATT); hb.dP_fin = H(B_DP2); hb.dP_os = H(B_DPOS); hb.dP_so = H(B_DPSO); hb.dP_g = H(B_DPG); hb.ld_dp = R;
  hb.hproj = F(B_HPROJ); hb.oproj = F(B_OPROJ); hb.U2h = F(B_U2H); hb.U2o = F(B_U2O); hb.GT = F(B_GT);
  hb.dS = H(B_DS); hb.ld_ds = 4 * R;
  hb.dhproj = F(B_DHPROJ); hb.doproj = F(B_DOPROJ); hb.dU2h = F(B_DU2H); hb.dU2o = F(B_DU2O); hb.dGT = F(B_DGT);
  hb.box_off = b.box_off; hb.hum_off = b.hum_off; hb.cell_off = b.cell_off; hb.num_images = B; hb.R = R;
  SKG_TRY(skg_hadamard_backward(&hb, NH, 1, st));
  // ---- U2h = enc_h W1h^T, U2o = enc_o W1o^T + b1 (attention_head.fc_1 split per node), attention_head_g.fc_1
  __nv_bfloat16* n16 = H(B_DNODE16);               // [NH + NB + B, 2R] bf16 operand copies of fp32 node-row gradients
  __nv_bfloat16* n16_o = n16 + (int64_t)NH * 2 * R;
  __nv_bfloat16* n16_g = n16 + (int64_t)(NH + NB) * 2 * R;
  SKG_TRY(skg_cast_pad_bf16(F(B_DU2H), R, nullptr, 0, NH, R, n16, 2 * R, R, st));
  SKG_TRY(skg_cast_pad_bf16(F(B_DU2O), R, nullptr, 0, NB, R, n16_o, 2 * R, R, st));
  SKG_TRY(skg_cast_pad_bf16(F(B_DGT), R, nullptr, 0, B, R, n16_g, 2 * R, R, st));
  SKG_TRY(wgrad(c, n16, 2 * R, R, H(B_ENC16), R, R, nullptr, NH, d.W_hproj, R, nullptr, false));
  SKG_TRY(wgrad(c, n16_o, 2 * R, R, H(B_ENC16) + (int64_t)NH * R, R, R, nullptr, NB, d.W_oproj, R, d.b_oproj, false));
  SKG_TRY(wgrad(c, n16_g, 2 * R, R, H(B_GFEAT), w.Kg, w.Kg, nullptr, B, d.Wg1, w.Kg, d.bg1, false));
  SKG_TRY(dgrad(c, n16, 2 * R, NH, R, wt.W_hprojT, 2 * R, R, nullptr, 0, nullptr, 0, nullptr, 0, F(B_DENC), R, nullptr));
  SKG_TRY(dgrad(c, n16_o, 2 * R, NB, R, wt.W_oprojT, 2 * R, R, nullptr, 0, nullptr, 0, nullptr, 0, F(B_DENC) + (int64_t)NH * R, R, nullptr));
  // ---- LayerNorm + residual + message ReLU (head.py:909-914, 921-925)
  SKG_TRY(skg_layernorm_backward(F(B_PRE), F(B_DENC), F(B_HN32), F(B_ON32), NH + NB, NH, R, w.ln_h_w, w.ln_o_w, w.ln_eps, F(B_DPRE),
                                 H(B_DM16), d.ln_h_w, d.ln_h_b, d.ln_o_w, d.ln_o_b, st));
  // ---- message fc_3 (applied after the weighted sum: rows = nodes)
  __nv_bfloat16* dm_o = H(B_DM16) + (int64_t)NH * R;
  __nv_bfloat16* z_o = H(B_Z) + (int64_t)NH * R;
  __nv_bfloat16* dz_o = H(B_DZ) + (int64_t)NH * R;
  SKG_TRY(wgrad(c, H(B_DM16), R, R, H(B_Z), R, R, nullptr, NH, d.W3os, R, d.b3os, false));
  SKG_TRY(wgrad(c, dm_o, R, R, z_o, R, R, nullptr, NB, d.W3so, R, d.b3so, false));
  SKG_TRY(dgrad(c, H(B_DM16), R, NH, R, wt.W3osT, R, R, nullptr, 0, nullptr, 0, H(B_DZ), R, nullptr, 0, nullptr));
  SKG_TRY(dgrad(c, dm_o, R, NB, R, wt.W3soT, R, R, nullptr, 0, nullptr, 0, dz_o, R, nullptr, 0, nullptr));
  // ---- softmax-weighted message sums (head.py:907-921) and adjacency (head.py:896-897)
  SKG_TRY(skg_message_aggregate_backward(F(B_ADJ), H(B_DZ), dz_o, H(B_HOS), H(B_HSO), R, b.box_off, b.hum_off, b.cell_off, B, NB, NH,
                                         b.max_boxes_per_image, R, H(B_DPOS), H(B_DPSO), R, F(B_DADJH), F(B_DADJO), st));
  SKG_TRY(skg_adjacency_backward(F(B_DADJH), F(B_DADJO), H(B_WTS), R, w.w_adj, C, R, H(B_DWTS), R, d.w_adj, d.b_adj, st));
  SKG_TRY(wgrad(c, H(B_DWTS), R, R, H(B_HATT), R, R, nullptr, C, d.W3att, R, d.b3att, true));     // second use of attention_head.fc_3
  SKG_TRY(stage_done(0));
  SKG_TRY(dgrad(c, H(B_DWTS), R, C, R, wt.W3attT, R, R, H(B_HATT), R, nullptr, 0, H(B_DPATT), R, nullptr, 0, nullptr));
  // ---- branch products of the first pass: dS (att | os | so) and the node-side sums dhproj / doproj
  SKG_TRY(skg_hadamard_backward(&hb, NH, 2, st));
  // ---- node projections [U_h | V_so] = hN W_hproj^T, [U_o | V_os] = oN W_oproj^T
  SKG_TRY(skg_cast_pad_bf16(F(B_DHPROJ), 2 * R, nullptr, 0, NH, 2 * R, n16, 2 * R, 2 * R, st));
  SKG_TRY(skg_cast_pad_bf16(F(B_DOPROJ), 2 * R, nullptr, 0, NB, 2 * R, n16_o, 2 * R, 2 * R, st));
  SKG_TRY(wgrad(c, n16, 2 * R, 2 * R, H(B_HN16), R, R, nullptr, NH, d.W_hproj, R, d.b_hproj, true));
  SKG_TRY(wgrad(c, n16_o, 2 * R, 2 * R, H(B_ON16), R, R, nullptr, NB, d.W_oproj, R, d.b_oproj, true));
  // d(fc_head | fc_tail output) = gate(projection path + residual path of the LayerNorm input); rows scattered to boxes
  SKG_TRY(dgrad(c, n16, 2 * R, NH, 2 * R, wt.W_hprojT, 2 * R, R, H(B_HN16), R, F(B_DPRE), R, H(B_GFC), 2 * R, nullptr, 0, I32(B_HUM2BOX)));
  SKG_TRY(dgrad(c, n16_o, 2 * R, NB, 2 * R, wt.W_oprojT, 2 * R, R, H(B_ON16), R, F(B_DPRE) + (int64_t)NH * R, R, H(B_GFC) + R, 2 * R, nullptr, 0,
                nullptr));
  // ---- fc_head | fc_tail (head.py:694-701, 884-885) as the stacked [2R, fc_k] GEMM of the forward, then box_head
  SKG_TRY(wgrad(c, H(B_GFC), 2 * R, 2 * R, H(B_OIN), w.fc_k, w.fc_k, nullptr, NB, d.Wfc, w.fc_k, d.bfc, false));
  SKG_TRY(dgrad(c, H(B_GFC), 2 * R, NB, 2 * R, wt.WfcT, 2 * R, E, H(B_OIN), w.fc_k, nullptr, 0, H(B_DX2), E, nullptr, 0, nullptr));
  SKG_TRY(wgrad(c, H(B_DX2), E, E, H(B_X1), E, E, nullptr, NB, d.Wb2, E, d.bb2, false));
  SKG_TRY(dgrad(c, H(B_DX2), E, NB, E, wt.Wb2T, E, E, H(B_X1), E, nullptr, 0, H(B_DX1), E, nullptr, 0, nullptr));
  SKG_TRY(wgrad(c, H(B_DX1), E, E, xb, w.K0, w.K0, nullptr, NB, d.Wb1, w.K0, d.bb1, false));
  SKG_TRY(stage_done(1));
  // ---- fused fc_2 of the four MBFs, then the spatial head (head.py:662-669): the cell-side chain, independent of the
  //      node-side chain above once the Hadamard backward has split dS from dhproj / doproj
  SKG_TRY(wgrad(c, H(B_DS), 4 * R, 4 * R, H(B_SP), R, R, nullptr, C, d.W2cat, R, d.b2cat, false));
  SKG_TRY(stage_done(2));
  SKG_TRY(dgrad(c, H(B_DS), 4 * R, C, 4 * R, wt.W2catT, 4 * R, R, H(B_SP), R, nullptr, 0, H(B_DSP), R, nullptr, 0, nullptr));
  SKG_TRY(wgrad(c, H(B_DSP), R, R, H(B_S2), 256, 256, nullptr, C, d.Wsp2, 256, d.bsp2, false));
  SKG_TRY(dgrad(c, H(B_DSP), R, C, R, wt.Wsp2T, R, 256, H(B_S2), 256, nullptr, 0, H(B_DS2), 256, nullptr, 0, nullptr));
  SKG_TRY(wgrad(c, H(B_DS2), 256, 256, H(B_S1), 128, 128, nullptr, C, d.Wsp1, 128, d.bsp1, false));
  SKG_TRY(dgrad(c, H(B_DS2), 256, C, 256, wt.Wsp1T, 256, 128, H(B_S1), 128, nullptr, 0, H(B_DS1), 128, nullptr, 0, nullptr));
  SKG_TRY(skg_spatial_fc1_backward(F(B_ENC), H(B_DS1), C, 46, 128, d.Wsp0, d.bsp0, st));
  SKG_TRY(stage_done(3));
  return SKG_OK;
}
