// Workspace layout of the whole-path forward / backward (csrc/forward.cu, csrc/backward.cu): one caller-owned buffer
// carved into named sub-buffers; the plan is a pure function of the sizes in skg_head_weights / skg_head_batch and the flags.
#pragma once
#include "common.cuh"
#include "skghoi_b200.h"
#include <string.h>

namespace skg {

enum Buf : int {
  B_SPLITK0 = 0, B_SPLITK1, B_SEG, B_ENC, B_X, B_Y, B_XKEEP, B_YKEEP, B_CELL2PAIR, B_CELL_HROW, B_CELL_OROW, B_CELL_IMG, B_PAIR2CELL, B_NANFLAG,
  B_S1, B_S2, B_SP, B_XB, B_X1, B_BOX2HUM, B_OIN, B_HN32, B_HN16, B_ON32, B_ON16, B_HPROJ, B_OPROJ, B_GFEAT, B_G32,
  B_GT, B_HATT, B_SRAW, B_HOS, B_HSO, B_HG, B_DOTPART, B_Z, B_ADJ, B_PRE, B_ENC16, B_ENC32, B_U2H, B_U2O, B_H2, B_FEAT,
  B_TMP, B_LOGITS, B_POST32, B_POST64,
  // training only (SKG_HEAD_TRAIN): what the backward keeps / produces (csrc/backward.cu)
  B_WTS, B_HUM2BOX, B_DLOGITS, B_GF, B_DP2, B_DPG, B_DPATT, B_DPOS, B_DPSO, B_DWTS, B_DS, B_DSP, B_DS2, B_DS1, B_DADJH, B_DADJO,
  B_DHPROJ, B_DOPROJ, B_DU2H, B_DU2O, B_DGT, B_DNODE16, B_DENC, B_DPRE, B_DM16, B_DZ, B_GFC, B_DX2, B_DX1, B_TRA, B_TRB,
  B_COUNT
};
static_assert(B_COUNT <= SKG_HEAD_NUM_BUFFERS, "raise SKG_HEAD_NUM_BUFFERS");

static const char* kBufNames[B_COUNT] = {
  "splitk_main", "splitk_side", "seg", "enc", "x", "y", "x_keep", "y_keep", "cell2pair", "cell_hrow", "cell_orow", "cell_img", "pair2cell", "nan_flag",
  "s1", "s2", "sp", "xb", "x1", "box2hum", "o_in", "hN32", "hN16", "oN32", "oN16", "hproj", "oproj", "gfeat", "g32",
  "Gt", "H_att", "S_raw", "H_os", "H_so", "H_g", "dot_part", "Z", "adjacency", "pre", "enc16", "enc32", "U2h", "U2o", "H2", "feat",
  "tmp", "logits", "post_i32", "post_i64",
  "wts", "hum2box", "dlogits", "dfeat", "dP_fin", "dP_g", "dP_att", "dP_os", "dP_so", "dwts", "dS", "dsp", "ds2", "ds1", "dadj_h", "dadj_o",
  "dhproj", "doproj", "dU2h", "dU2o", "dGt", "dnode16", "denc", "dpre", "dmsg16", "dZ", "dfc", "dx2", "dx1", "tr_a", "tr_b"};

static inline int64_t align_up(int64_t v, int64_t a) { return (v + a - 1) / a * a; }

static int make_plan(const skg_head_weights& w, const skg_head_batch& b, int flags, skg_head_plan_t* p) {
  SKG_REQUIRE(w.R > 0 && w.R % 256 == 0, "representation_size must be a multiple of 256 for the tensor-core path, got %d", w.R);
  SKG_REQUIRE(w.E > 0 && w.E % 128 == 0, "node_encoding_size must be a multiple of 128, got %d", w.E);
  SKG_REQUIRE(w.K0 % 8 == 0 && w.fc_k % 8 == 0 && w.Kg % 8 == 0 && w.K0 >= w.pool_k && w.Kg >= w.Cg, "padded K sizes must be multiples of 8");
  SKG_REQUIRE(w.fc_k >= w.E + w.ent_dim, "fc_k must cover node encoding + TransH dim");
  SKG_REQUIRE(w.sp_in == 46, "spatial_head input must be 46 (2 x 23 ratio features, head.py:663)");
  SKG_REQUIRE(w.n_logit == w.num_cls + 1 && w.n_logit_pad >= w.n_logit && w.n_logit_pad % 128 == 0, "bad logit sizes");
  SKG_REQUIRE(b.num_images >= 0 && b.total_boxes >= 0 && b.total_humans >= 0 && b.total_cells >= 0 && b.total_pairs >= 0, "negative sizes");
  const int64_t C = b.total_cells, P = b.total_pairs, NB = b.total_boxes, NH = b.total_humans, B = b.num_images;
  const int R = w.R;
  const bool train = (flags & SKG_HEAD_TRAIN) != 0;
  const bool precise = (flags & SKG_HEAD_PRECISE_LOGITS) != 0 && !train;
  memset(p, 0, sizeof(*p));
  p->num_buffers = B_COUNT;
  auto set = [&](int i, int64_t rows, int64_t cols, int es) { p->rows[i] = rows; p->cols[i] = cols; p->elem_size[i] = es; };
  // split-K scratch of the two streams: FIRST in the layout so their position does not depend on the batch shape;
  // must be zero before the first call on a workspace and are left zero by every call (skg_gemm_bf16_splitk)
  const int64_t sk_elems = skg_gemm_splitk_scratch_bytes(0, 0) / 4;
  set(B_SPLITK0, 1, sk_elems, 4); set(B_SPLITK1, 1, sk_elems, 4);
  set(B_SEG, 1, b.seg_bytes > 0 ? b.seg_bytes : 0, 1);
  set(B_ENC, C, 46, 4); set(B_X, C, 1, 8); set(B_Y, C, 1, 8); set(B_XKEEP, P, 1, 8); set(B_YKEEP, P, 1, 8);
  set(B_CELL2PAIR, C, 1, 4); set(B_CELL_HROW, C, 1, 4); set(B_CELL_OROW, C, 1, 4); set(B_CELL_IMG, C, 1, 4);
  set(B_PAIR2CELL, P, 1, 4); set(B_NANFLAG, B > 0 ? B : 1, 1, 4);
  set(B_S1, C, 128, 2); set(B_S2, C, 256, 2); set(B_SP, C, R, 2);
  set(B_XB, NB, w.K0, 2); set(B_X1, NB, w.E, 2); set(B_BOX2HUM, NB, 1, 4); set(B_OIN, NB, w.fc_k, 2);
  set(B_HN32, NH, R, 4); set(B_HN16, NH, R, 2); set(B_ON32, NB, R, 4); set(B_ON16, NB, R, 2);
  set(B_HPROJ, NH, 2 * R, 4); set(B_OPROJ, NB, 2 * R, 4);
  set(B_GFEAT, B, w.Kg, 2); set(B_G32, B, w.Cg, 4); set(B_GT, B, R, 4);
  set(B_HATT, C, R, 2); set(B_SRAW, C, train ? 4 * R : R, 2); set(B_HOS, C, R, 2); set(B_HSO, C, R, 2); set(B_HG, C, R, 2);
  set(B_DOTPART, C, 4 * (R / 256), 4); set(B_Z, NH + NB, R, 2); set(B_ADJ, C, 1, 4); set(B_PRE, NH + NB, R, 4);
  set(B_ENC16, NH + NB, R, 2); set(B_ENC32, (flags & SKG_HEAD_WANT_TAPS) ? NH + NB : 0, R, 4);
  set(B_U2H, NH, R, 4); set(B_U2O, NB, R, 4); set(B_H2, C, R, 2);
  // training keeps the pair features and the logits in CELL rows (the diagonal cells i == j ride along and are masked by
  // the loss): no scatter / gather between the pair-compacted and the cell layouts anywhere in the backward
  const int64_t FR = train ? C : P;
  set(B_FEAT, FR, precise ? 4 * R : 2 * R, 2); set(B_TMP, precise ? P : 0, w.n_logit_pad, 4); set(B_LOGITS, FR, w.n_logit, 4);
  set(B_POST32, 2 * NB, 1, 4); set(B_POST64, NH + 2 * B + 1, 1, 8);   // scratch of skg_head_post
  if (train) {
    const int64_t NN = NH + NB;
    const int64_t Cp = align_up(C, 64), NBp = align_up(NB, 64);
    set(B_WTS, C, R, 2); set(B_HUM2BOX, NH, 1, 4);
    set(B_DLOGITS, C, w.n_logit_pad, 2); set(B_GF, C, 2 * R, 2);
    set(B_DP2, C, R, 2); set(B_DPG, C, R, 2); set(B_DPATT, C, R, 2); set(B_DPOS, C, R, 2); set(B_DPSO, C, R, 2); set(B_DWTS, C, R, 2);
    set(B_DS, C, 4 * R, 2); set(B_DSP, C, R, 2); set(B_DS2, C, 256, 2); set(B_DS1, C, 128, 2);
    set(B_DADJH, C, 1, 4); set(B_DADJO, C, 1, 4);
    set(B_DHPROJ, NH, 2 * R, 4); set(B_DOPROJ, NB, 2 * R, 4); set(B_DU2H, NH, R, 4); set(B_DU2O, NB, R, 4); set(B_DGT, B, R, 4);
    set(B_DNODE16, NN + B, 2 * R, 2);      // bf16 operand copies of the node-row gradients (reused stage by stage)
    set(B_DENC, NN, R, 4); set(B_DPRE, NN, R, 4); set(B_DM16, NN, R, 2); set(B_DZ, NN, R, 2);
    set(B_GFC, NB, 2 * R, 2); set(B_DX2, NB, w.E, 2); set(B_DX1, NB, w.E, 2);
    // M-major (transposed) operand scratch of the weight-gradient GEMMs: tr_a holds G^T, tr_b holds X^T
    int64_t ta = 4 * R * Cp, tb = 2 * R * Cp;
    const int64_t kmax = w.K0 > w.fc_k ? w.K0 : w.fc_k;
    if (kmax * NBp > tb) tb = kmax * NBp;
    if (2 * R * NBp > ta) ta = 2 * R * NBp;
    if ((int64_t)R * 64 > ta) ta = (int64_t)R * 64;
    if ((int64_t)w.Kg * 64 > tb) tb = (int64_t)w.Kg * 64;
    set(B_TRA, 1, ta, 2); set(B_TRB, 1, tb, 2);
  }
  int64_t off = 0;
  for (int i = 0; i < B_COUNT; ++i) {
    p->offset[i] = off;
    off = align_up(off + p->rows[i] * p->cols[i] * p->elem_size[i], 256);
  }
  p->total_bytes = off > 0 ? off : 256;
  return SKG_OK;
}

static inline skg_epilogue_group group0() {
  skg_epilogue_group g;
  memset(&g, 0, sizeof(g));
  return g;
}

#define SKG_TRY(call)        \
  do {                       \
    int rc__ = (call);       \
    if (rc__) return rc__;   \
  } while (0)

}  // namespace skg
