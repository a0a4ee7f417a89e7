/* skghoi_b200 -- C ABI of the B200-native SKGHOI interaction-head library (libskghoi_b200.so).
 *
 * This is the drop-in boundary: plain pointers and sizes, no torch types.  Every entry point
 * replaces one piece of the reference's Python operator surface and cites it (paths relative to
 * the reference checkout; "head.py" = heads/adamixer_transH_spatial_r50_head.py).
 *
 * Conventions
 *   - all pointers are DEVICE pointers on the caller's current device unless marked "host";
 *   - caller allocates every buffer; the library allocates nothing in steady state;
 *   - work is enqueued on `stream` (cudaStream_t passed as void*-compatible handle) and the call
 *     returns without synchronising;
 *   - return value: 0 = ok, SKG_ERR_INVALID (-1) bad argument, SKG_ERR_CUDA (-2) CUDA failure,
 *     SKG_ERR_UNSUPPORTED (-3); skg_last_error() gives the text (thread-local);
 *   - fp32 tensors are row-major contiguous; indices follow the reference (int64 for the
 *     torch-visible index tensors, int32 for internal row maps).
 */
#ifndef SKGHOI_B200_H_
#define SKGHOI_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SKG_ABI_VERSION 1

#ifndef __CUDA_RUNTIME_H__
typedef struct CUstream_st* cudaStream_t;
#endif

const char* skg_last_error(void);
int skg_abi_version(void);
int skg_num_sms(void);
/* number of kernel launches this library has enqueued in the process so far (bench.py "gpu_launches") */
long long skg_launch_count(void);

/* ---- pair enumeration + spatial encoding --------------------------------------------------
 * Replaces, for a whole batch in one launch: the meshgrid / nonzero(x != y) pair construction of
 * GraphHead.forward (head.py:847-860), ops.compute_spatial_ratio_encodings (ops.py:85-157,
 * num_features = 23 -> 46 columns) or ops.compute_spatial_encodings (ops.py:20-83, 18 -> 36), and
 * the NaN guard (head.py:866-868, nan_guard != 0).
 * Segment table (device): box_off[B+1] prefix of n, n_h[B], cell_off[B+1] prefix of n_h*n,
 * pair_off[B+1] prefix of n_h*(n-1), hum_off[B+1] prefix of n_h, hw[B,3] = (h, w, h*w) as fp32.
 * Outputs: enc[cells, 2F] (required); all index outputs optional (NULL to skip):
 *   x, y [cells] int64 (flattened meshgrid, head.py:860); x_keep, y_keep [pairs] int64 (head.py:852);
 *   cell2pair [cells] (pair rank or -1), cell_hrow / cell_orow / cell_img [cells], pair2cell [pairs];
 *   nan_flag [B] (1 when the image's encoding held a NaN before the guard). */
int skg_pairs_spatial_encode(const float* boxes, const int32_t* box_off, const int32_t* n_h,
                             const int64_t* cell_off, const int64_t* pair_off, const int32_t* hum_off,
                             const float* hw, int num_images, int64_t total_cells, int num_features,
                             float eps, int nan_guard, float* enc, int64_t* x, int64_t* y,
                             int64_t* x_keep, int64_t* y_keep, int32_t* cell2pair, int32_t* cell_hrow,
                             int32_t* cell_orow, int32_t* cell_img, int32_t* pair2cell, int32_t* nan_flag,
                             cudaStream_t stream);

/* ops.compute_spatial_encodings / compute_spatial_ratio_encodings on explicit box rows
 * (ops.py:20-23, 85-88): boxes_1/boxes_2 [rows,4] are the concatenated lists, row_off[num_lists+1]
 * their prefix offsets, hw[num_lists,3] = (h, w, h*w). */
int skg_spatial_encode_rows(const float* boxes_1, const float* boxes_2, const int64_t* row_off,
                            const float* hw, int num_lists, int64_t rows, int num_features, float eps,
                            float* enc, int32_t* nan_flag, cudaStream_t stream);

/* ---- TransH ------------------------------------------------------------------------------
 * TransH.forward (heads/TransH/TransH.py:87-106, 'normal' mode): score[t] =
 * || norm(h - (h.w^)w^) + norm(r) - norm(t - (t.w^)w^) ||_p  with w^ = w/max(||w||,1e-12);
 * norm() applied when norm_flag.  h_out/r_out/w_out/t_out [T,dim] receive the RAW gathered rows
 * the fork returns (all four or none). */
int skg_transh_forward(const float* ent, const float* rel, const float* nrm, int ent_tot, int rel_tot,
                       int dim, const int64_t* head, const int64_t* relation, const int64_t* tail,
                       int64_t num_triples, int p_norm, int norm_flag, float* h_out, float* r_out,
                       float* w_out, float* t_out, float* score, cudaStream_t stream);

/* transH_head.forward (head.py:558-582) without materialising the index vectors: triple
 * t = cell*rel_tot + r has head = head_const, relation = r, tail = cell_tail[cell]
 * (cell_tail NULL -> tail = cell). */
int skg_transh_forward_grid(const float* ent, const float* rel, const float* nrm, int ent_tot, int rel_tot,
                            int dim, int64_t head_const, const int64_t* cell_tail, int64_t num_cells,
                            int p_norm, int norm_flag, float* h_out, float* r_out, float* w_out,
                            float* t_out, float* score, cudaStream_t stream);

/* Backward of skg_transh_forward wrt the three tables: d_* += dscore[t] * d score[t] / d table
 * (fp32 atomics; caller zeroes d_ent/d_rel/d_nrm). */
int skg_transh_backward(const float* ent, const float* rel, const float* nrm, int ent_tot, int rel_tot,
                        int dim, const int64_t* head, const int64_t* relation, const int64_t* tail,
                        int64_t num_triples, int p_norm, int norm_flag, const float* dscore, float* d_ent,
                        float* d_rel, float* d_nrm, cudaStream_t stream);

/* MarginLoss.forward (heads/MarginLoss.py:28-36) for p_score[rows] broadcast against
 * n_score[rows, cols]; use_adv selects the self-adversarial weighting (MarginLoss.py:25-30).
 * loss[0] receives the scalar.  d_p[rows] / d_n[rows, cols] (both or neither) receive
 * grad_scale * d loss / d score. */
int skg_margin_loss(const float* p_score, const float* n_score, int64_t rows, int64_t cols, float margin,
                    float adv_temperature, int use_adv, float* loss, float* d_p, float* d_n,
                    float grad_scale, cudaStream_t stream);

/* ---- losses --------------------------------------------------------------------------------
 * ops.binary_focal_loss (ops.py:159-211), element-wise: loss[i] (optional) and, when dx is set,
 * dx[i] = grad_out[i] * d loss[i] / d x[i].  Reductions (sum / mean) are done by the caller. */
int skg_focal_loss(const float* x, const float* y, int64_t n, float alpha, float gamma, float eps,
                   float* loss, const float* grad_out, float* dx, cudaStream_t stream);

/* ---- bf16 tensor-core GEMM with fused epilogues ---------------------------------------------
 * C[M,N] = epilogue(A[M,K] x B[N,K]^T), A and B bf16 with K contiguous (B in nn.Linear weight
 * layout), fp32 accumulation in TMEM (tcgen05.mma).  Carries every dense layer of the head:
 * MultiBranchFusion / MessageMBF (head.py:431-530, concatenated-branch form), box_head,
 * spatial_head, fc_head / fc_tail (head.py:635-701), predictor / suppressor (models.py:176-177).
 *
 * The output columns are split into up to 4 groups (sorted, n_begin multiples of block_n); for a
 * row r and a column c of group g (lc = c - n_begin):
 *     v = acc + bias[c]                                   (bias NULL -> 0; indexed by GLOBAL column)
 *     raw_bf16[out_row[r]][lc] = v                        (optional copy before the product)
 *     v *= mul_p[row_p[r]][lc] (+ mul_q[row_q[r]][lc])    (optional Hadamard with gathered fp32 tables)
 *     v = max(v, 0)                                       (flags & SKG_EPI_RELU)
 *     v += residual[row_r[r]][lc]                         (optional)
 *     dot_out[out_row[r]][4*tile + h] = sum_lc v * dot_w[lc]  (optional fused row-dot; four partials per N tile,
 *                                                          one per column quarter h, summed by the consumer)
 *     v = mask_bf16[r][lc] > 0 ? v : 0                     (optional: the ReLU mask of a backward (dX) GEMM)
 *     out_bf16 / out_f32 [out_row[r]][lc] = v             (optional; out_row NULL = r, < 0 = skip row)
 * row_* NULL means the identity map.  bias, dot_w and the gathered tables are read in 32-column
 * chunks, so they (and out_bf16 / raw_bf16 rows) must be padded to a multiple of 32 columns;
 * out_f32 is masked at N exactly. */
#define SKG_EPI_RELU 1
typedef struct skg_epilogue_group {
  int32_t n_begin;
  int32_t flags;
  const float* bias;
  const float* mul_p; const int32_t* row_p; int32_t ld_p;
  const float* mul_q; const int32_t* row_q; int32_t ld_q;
  const float* residual; const int32_t* row_r; int32_t ld_r;
  const float* dot_w; float* dot_out; int32_t ld_dot;
  void* out_bf16; int32_t ld_ob;
  float* out_f32; int32_t ld_of;
  void* raw_bf16; int32_t ld_raw;
  const int32_t* out_row;
  void* lo_bf16; int32_t ld_lo;   /* optional: bf16(v - float(bf16(v))), the rounding residual of out_bf16 */
  const void* mask_bf16; int32_t ld_mask;   /* optional ReLU-backward gate, applied last before the stores:
                                               v = mask_bf16[r][lc] > 0 ? v : 0  (r = GEMM row, before out_row) */
} skg_epilogue_group;

/* groups is a HOST array of num_groups descriptors (copied into the launch).  block_n = 128, 256
 * or 0 (auto: 256 when N % 256 == 0).  lda / ldb are row pitches in elements (multiples of 8). */
int skg_gemm_bf16(const void* A, int64_t lda, const void* B, int64_t ldb, int M, int N, int K,
                  const skg_epilogue_group* groups, int num_groups, int block_n, cudaStream_t stream);

/* Same GEMM with split-K for shapes whose tiles alone leave most SMs idle (M = node rows of a batch):
 * `scratch` is a caller-owned region of at least skg_gemm_splitk_scratch_bytes() bytes that must be ZERO on
 * entry and is left zero on exit (partial tiles are reduced with fp32 red.global.add; the last split to arrive
 * runs the epilogue and re-zeroes its tile).  One scratch region may serve successive launches of ONE stream.
 * scratch NULL (or too small) -> no split, identical to skg_gemm_bf16. */
int skg_gemm_bf16_splitk(const void* A, int64_t lda, const void* B, int64_t ldb, int M, int N, int K,
                         const skg_epilogue_group* groups, int num_groups, int block_n, void* scratch,
                         int64_t scratch_bytes, cudaStream_t stream);
/* The same with A read CYCLICALLY along K: A has a_k_wrap columns (a multiple of 64, <= K; 0 = plain) and k-block kb reads
 * columns (64 kb) mod a_k_wrap -- the split-bf16 predictor [hi | lo | hi] x [W_hi | W_hi | W_lo]^T as ONE K = 6R launch over
 * the [hi | lo] feature rows (head.py:410-411 in ~fp32 precision, 3 bf16 passes). */
int skg_gemm_bf16_wrapk(const void* A, int64_t lda, int a_k_wrap, const void* B, int64_t ldb, int M, int N, int K,
                        const skg_epilogue_group* groups, int num_groups, int block_n, void* scratch,
                        int64_t scratch_bytes, cudaStream_t stream);
int64_t skg_gemm_splitk_scratch_bytes(int max_tiles_m, int max_n);

/* Allocation-free skg_gemm_bf16_splitk (csrc/noalloc.cu).  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  Same GEMM, same arguments minus the scratch: the split-K scratch is a file-scope device array inside the library
 * of SKG_GEMM_STATIC_SCRATCH_BYTES (every GPU that loads the library holds it until exit), and the call makes no
 * cudaMalloc / cudaFree / allocating library call.
 *   limit   ceil(M / 128) * 128 x ceil(N / 256) * 256 fp32 + 4096 bytes <= SKG_GEMM_STATIC_SCRATCH_BYTES (M x N up to
 *           1024 x 1024 after padding; the head's box_head fc6 at up to 1024 boxes); larger shapes fail with
 *           SKG_ERR_INVALID and a message naming the bytes needed -- nothing is launched
 *   sharing every call on a device uses the ONE array; the part a call uses is cleared on `stream` in front of its launch
 *           and left zero behind it.  Calls that overlap on the GPU (two streams, two threads) RACE: serialise them, or use
 *           skg_gemm_bf16_splitk with one caller-owned region per stream.
 * block_n = 0 (auto), 128 or 256.  Results equal skg_gemm_bf16_splitk's up to fp32 summation order. */
#define SKG_GEMM_STATIC_SCRATCH_BYTES (4 * 1024 * 1024 + 4096)
int skg_gemm_bf16_splitk_static(const void* A, int64_t lda, const void* B, int64_t ldb, int M, int N, int K,
                                const skg_epilogue_group* groups, int num_groups, int block_n, cudaStream_t stream);
int64_t skg_gemm_static_scratch_bytes(void);

/* ---- glue kernels between the GEMMs (csrc/headops.cu) ---------------------------------------- */

/* dst[r, 0:cols_pad] = bf16(relu?(src[r, 0:cols] + bias)), zero padded (operand packing for the GEMMs). */
int skg_cast_pad_bf16(const float* src, int64_t ld_src, const float* bias, int relu, int64_t rows, int cols,
                      void* dst, int64_t ld_dst, int cols_pad, cudaStream_t stream);

/* spatial_head layer 0 (head.py:663-664), fp32 math: out_bf16[cells,128] = relu(enc[cells,46] W^T + b). */
int skg_spatial_fc1(const float* enc, const float* weight, const float* bias, int64_t cells, int in_features,
                    int out_features, void* out_bf16, cudaStream_t stream);

/* Inputs of fc_head / fc_tail (head.py:875-885): h_in[humans, ld] = cat(node_enc[human box], E[human_idx]),
 * o_in[boxes, ld] = cat(node_enc[box], E[box index within image]) as bf16, zero padded to ld. */
int skg_node_inputs(const void* node_enc_bf16, int enc_dim, const float* ent, int ent_dim, int ent_tot,
                    int human_idx, const int32_t* box_off, const int32_t* hum_off, int num_images,
                    int total_boxes, int total_humans, void* h_in, void* o_in, int ld, cudaStream_t stream);

/* Fused-fc variant of skg_node_inputs: fills only the TransH entity columns o_in[box, enc_dim:ld] = bf16(E[index of the
 * box inside its image]) (tail_stack == E[y], head.py:572; the node encoding columns are written by the box_head GEMM)
 * and box2hum[boxes] = human row of the box (first n_h boxes of every image) or -1. */
int skg_node_prep(const float* ent, int ent_dim, int ent_tot, const int32_t* box_off, const int32_t* n_h,
                  const int32_t* hum_off, int num_images, int total_boxes, void* o_in, int ld, int enc_dim,
                  int32_t* box2hum, cudaStream_t stream);

/* bias_out[r] = bias_in[r] + sum_k W_ent[r, k] * ent[entity, k]  (r < rows): the constant contribution of the human side's
 * entity columns of fc_head (head.py:571, 877-884: head_stack == E[human_idx] for every row). */
int skg_fc_bias_fold(const float* bias_in, const float* W_ent, const float* ent, int ent_dim, int entity, int rows,
                     float* bias_out, cudaStream_t stream);

/* avg_pool(features['3']).flatten(1) (head.py:811): planes = B*C, hw = H*W. */
int skg_global_pool(const float* features, int64_t planes, int hw, void* out_bf16, float* out_f32,
                    cudaStream_t stream);

/* Message aggregation (head.py:897, 909-910, 916-921): adjacency = sum(dot_part) + adj_bias,
 * Z_h[i] = sum_j softmax_j(adj[i,:]) H_os[i,j], Z_o[j] = sum_i softmax_i(adj[:,j]) H_so[i,j] (bf16). */
int skg_message_aggregate(const float* dot_part, int num_part, float adj_bias, const void* H_os, const void* H_so,
                          int64_t ld_h, const int32_t* box_off, const int32_t* hum_off, const int64_t* cell_off,
                          int num_images, int total_boxes, int total_humans, int max_per_image, int width,
                          void* Z_h, void* Z_o, float* adjacency, cudaStream_t stream);

/* LayerNorm (head.py:912-914, 923-925): rows [0, rows_a) use (w_a, b_a), the rest (w_b, b_b). */
int skg_layernorm(const float* x, int64_t rows, int64_t rows_a, int width, const float* w_a, const float* b_a,
                  const float* w_b, const float* b_b, float eps, float* out_f32, void* out_bf16,
                  cudaStream_t stream);

/* Branch product of the final attention head (head.py:966-970):
 * H[cell] = bf16(relu(S_raw[cell] * (U_h[hrow[cell]] + U_o[orow[cell]]))). */
int skg_hadamard_cells(const void* S_raw, int64_t ld_s, const float* U_h, int64_t ld_uh, const float* U_o,
                       int64_t ld_uo, const int32_t* hrow, const int32_t* orow, int64_t cells, int width,
                       void* H, int64_t ld_hh, cudaStream_t stream);

/* ---- priors, ground-truth association, post-processing (csrc/postproc.cu) ---------------------- */
typedef struct skg_post_tables {
  const int32_t* verb_off;    /* [num_obj_cls+1] CSR of valid verbs per object class, ascending (object_class_to_target_class) */
  const int32_t* verb_idx;    /* [nnz] */
  const int32_t* pair2cell;   /* [P]     from skg_pairs_spatial_encode */
  const int32_t* cell_hrow;   /* [cells] */
  const int32_t* cell_orow;   /* [cells] */
  const int32_t* cell_img;    /* [cells] */
  const int32_t* hum_off;     /* [B+1] */
  const int32_t* box_off;     /* [B+1] */
  const int64_t* pair_off;    /* [B+1] */
  const int64_t* labels;      /* [boxes] object class per box */
  const float* scores;        /* [boxes] detection scores */
  const int32_t* box_prefix;  /* [boxes]  exclusive within-image prefix of box_cnt (postprocess only) */
  const int32_t* box_cnt;     /* [boxes]  number of valid verbs of the box's class (postprocess only) */
  const int64_t* hum_base;    /* [humans] batch-global output offset of each human row (postprocess only) */
  const int32_t* logit_row;   /* [P] row of pair p inside `logits` (NULL = p).  The training layout (SKG_HEAD_TRAIN) keeps the
                                 logits in cell rows: logit_row = pair2cell */
} skg_post_tables;

/* InteractionHead.postprocess (head.py:291-337) for the whole batch: logits [P, ld] holds num_cls verb
 * logits followed by the suppressor logit.  For every (pair, valid verb) entry e (row-major nonzero order
 * of prior[0], head.py:315): index, prediction, scores = sigmoid(logit) * prior_h * prior_o * weights,
 * prior [2, total_entries]; per pair: weights = sigmoid(suppressor logit), object class; with gt_labels
 * [P, num_cls]: labels per entry and unary_labels per pair (head.py:325-330).  tables is a HOST struct
 * of DEVICE pointers. */
int skg_postprocess(const float* logits, int ld_logits, int num_cls, const skg_post_tables* tables, int64_t pairs,
                    float prior_power, const float* gt_labels, int64_t total_entries, int64_t* out_index,
                    int64_t* out_pred, float* out_scores, float* out_prior, float* out_weights,
                    float* out_labels, float* out_unary, int64_t* out_object, cudaStream_t stream);

/* Output offsets of skg_postprocess built on the device (no host round trip): box_cnt / box_prefix [boxes],
 * hum_base [humans] (batch-global), img_entries [B] and img_base [B+1] (entries per image and their prefix) --
 * the closed form of the per-image nonzero(prior[0]) ordering of head.py:315.  verb_off as in skg_post_tables. */
int skg_post_offsets(const int32_t* verb_off, const int64_t* labels, const float* scores, const int32_t* box_off,
                     const int32_t* hum_off, int num_images, int32_t* box_cnt, int32_t* box_prefix,
                     int64_t* hum_base, int64_t* img_entries, int64_t* img_base, cudaStream_t stream);

/* boxes_h / boxes_o [P, 4] of every valid pair (head.py:975-976: coords[x_keep], coords[y_keep]). */
int skg_gather_pair_boxes(const float* boxes, const skg_post_tables* tables, int64_t pairs, float* boxes_h,
                          float* boxes_o, cudaStream_t stream);

/* skg_post_offsets + skg_postprocess + skg_gather_pair_boxes as ONE host call (InteractionHead.postprocess for a batch,
 * head.py:291-337, 975-976).  tables->box_cnt / box_prefix / hum_base are ignored and built into the caller's scratch:
 * scratch_i32 [2 * total_boxes], scratch_i64 [total_humans + 2 * num_images + 1]. */
int skg_head_post(const float* logits, int ld_logits, int num_cls, const skg_post_tables* tables, const float* boxes,
                  int num_images, int total_boxes, int total_humans, int64_t pairs, float prior_power,
                  const float* gt_labels, int64_t total_entries, int32_t* scratch_i32, int64_t* scratch_i64,
                  int64_t* out_index, int64_t* out_pred, float* out_scores, float* out_prior, float* out_weights,
                  float* out_labels, float* out_unary, int64_t* out_object, float* boxes_h, float* boxes_o,
                  cudaStream_t stream);

/* compute_prior_scores (head.py:721-767): dense prior [2, P, num_cls]. */
int skg_prior_dense(const skg_post_tables* tables, int64_t pairs, int num_cls, float prior_power, float* prior,
                    cudaStream_t stream);

/* associate_with_ground_truth (head.py:703-719): labels [P, num_cls] (zeroed here), GT pairs in CSR by image. */
int skg_associate_gt(const float* boxes, const skg_post_tables* tables, int64_t pairs, int num_cls,
                     const int32_t* gt_off, const float* gt_boxes_h, const float* gt_boxes_o,
                     const int64_t* gt_verb, float fg_iou_thresh, float* labels, cudaStream_t stream);


/* ---- backward-pass kernels of the training step (csrc/trainops.cu; BASELINE configs[2]) -----------------------------
 * The reference trains the head with autograd (utils.py:213-229: forward, loss.backward(), optimizer step); these are the
 * backward counterparts of the fused forward kernels above, so that the whole step runs on this library. */

/* out[c, r] = in[row_map ? row_map[r] : r, c] for r < rows, c < cols (bf16; rows with a negative map entry read as zero):
 * the M-major copies the weight-gradient GEMMs dW = G^T X consume.  colsum (optional, fp32 [cols], caller-zeroed)
 * += column sums of `in` -- the bias gradient of the layer comes for free with the transposition of G. */
int skg_transpose_bf16(const void* in, int64_t ld_in, int64_t rows, int cols, const int32_t* row_map, void* out,
                       int64_t ld_out, float* colsum, cudaStream_t stream);

/* d(g_hoi * hoi_loss + g_int * interactiveness_loss) / d logits (compute_interaction_classification_loss head.py:153-177,
 * compute_interactiveness_loss head.py:180-205 through postprocess head.py:308-316, binary_focal_loss ops.py:159-211):
 * logits [*, ld_logits] fp32 (num_cls verb logits + the suppressor logit; the row of pair p is logit_row[p], or p when
 * NULL), gt_labels [P, num_cls], n_p[2] the two normalisers, g_hoi / g_int device scalars (NULL = 1).  dlogits
 * [*, ld_d] bf16 must be ZERO on entry; entries of valid (pair, verb) combinations and the suppressor column are written. */
int skg_head_loss_backward(const float* logits, int ld_logits, int num_cls, const skg_post_tables* tables,
                           const int32_t* logit_row, int64_t pairs, float prior_power, const float* gt_labels,
                           const float* n_p, const float* g_hoi, const float* g_int, void* dlogits_bf16, int ld_d,
                           cudaStream_t stream);

/* Backward of the branch products fc_1(app) * fc_2(sp) of the four MBFs (MultiBranchFusion head.py:469-474, MessageMBF
 * head.py:509-527), all five products of a cell in one pass.  S [cells, 4R] raw fc_2 outputs (attention_head | obj_to_sub |
 * sub_to_obj | attention_head_g); dP_* [cells, R] gradients of the pre-activations (att = attention weights pass,
 * fin = final attention pass on the refined nodes); hproj [NH, 2R] = (U_h | V_so), oproj [NB, 2R] = (U_o | V_os), U2h,
 * U2o, GT the appearance-side tables.  Writes dS [cells, 4R] (bf16) and the per-node sums dhproj, dU2h (stored) and
 * doproj, dU2o, dGT (ACCUMULATED with fp32 atomics: caller zeroes them). */
typedef struct skg_hadamard_bwd {
  const void* S; int64_t ld_s;
  const void* dP_att; const void* dP_fin; const void* dP_os; const void* dP_so; const void* dP_g; int64_t ld_dp;
  const float* hproj; const float* oproj; const float* U2h; const float* U2o; const float* GT;
  void* dS; int64_t ld_ds;
  float* dhproj; float* doproj; float* dU2h; float* dU2o; float* dGT;
  const int32_t* box_off; const int32_t* hum_off; const int64_t* cell_off;
  int32_t num_images, R;
} skg_hadamard_bwd;
/* phase 1 (needs dP_fin, dP_g): dS[:, 3R:4R], dU2h, dU2o, dGT.  phase 2 (needs all five): dS[:, 0:3R], dhproj, doproj. */
int skg_hadamard_backward(const skg_hadamard_bwd* args, int total_humans, int phase, cudaStream_t stream);

/* Backward of skg_message_aggregate (head.py:907-921): given dZ_h [NH, width] / dZ_o [NB, width] (bf16) writes the gated
 * gradients of the message pre-activations dP_os / dP_so [cells, ld_dp] (bf16) and the two halves of the adjacency
 * gradient, dadj_h / dadj_o [cells] (softmax over an image's boxes resp. humans); adjacency [cells] as the forward stored it. */
int skg_message_aggregate_backward(const float* adjacency, const void* dZ_h, const void* dZ_o, const void* H_os,
                                   const void* H_so, int64_t ld_h, const int32_t* box_off, const int32_t* hum_off,
                                   const int64_t* cell_off, int num_images, int total_boxes, int total_humans,
                                   int max_per_image, int width, void* dP_os, void* dP_so, int64_t ld_dp, float* dadj_h,
                                   float* dadj_o, cudaStream_t stream);

/* Backward of the adjacency Linear(R -> 1) on wts = relu(fc_3(...)) kept in bf16 (head.py:896-897): dwts [cells, ld_dw]
 * (bf16, ReLU gated); dw_adj [width] and db_adj [1] are ACCUMULATED. */
int skg_adjacency_backward(const float* dadj_h, const float* dadj_o, const void* wts, int64_t ld_w, const float* w_adj,
                           int64_t cells, int width, void* dwts, int64_t ld_dw, float* dw_adj, float* db_adj,
                           cudaStream_t stream);

/* Backward of skg_layernorm on x = base + relu(msg) (head.py:909-914, 921-925): dx [rows, width] fp32 (the gradient that
 * also flows down the residual path), dmsg_bf16 = dx gated by (x - base > 0); dw_* / db_* [width] are ACCUMULATED.
 * Rows [0, rows_a) use (base_a, w_a), the rest (base_b [rows - rows_a, width], w_b). */
int skg_layernorm_backward(const float* x, const float* dy, const float* base_a, const float* base_b, int64_t rows,
                           int64_t rows_a, int width, const float* w_a, const float* w_b, float eps, float* dx,
                           void* dmsg_bf16, float* dw_a, float* db_a, float* dw_b, float* db_b, cudaStream_t stream);

/* Weight / bias gradient of skg_spatial_fc1 (head.py:663): dW [out, in] and db [out] are ACCUMULATED. */
int skg_spatial_fc1_backward(const float* enc, const void* g_bf16, int64_t cells, int in_features, int out_features,
                             float* dW, float* db, cudaStream_t stream);

/* hum2box[box2hum[k]] = k: inverse of the box -> human-row map of skg_node_prep (human rows are the first n_h boxes). */
int skg_hum2box(const int32_t* box2hum, int total_boxes, int32_t* hum2box, cudaStream_t stream);

/* Descriptor-driven block copy: weight packing (the reference's per-branch fp32 parameters, state_dict layout of SURVEY
 * 8(a), -> the concatenated bf16 operands of skg_head_weights, plus their transposes for the dX GEMMs) and gradient
 * unpacking (packed fp32 gradients -> the per-branch .grad tensors) as ONE launch each.  descs / tiles live in DEVICE
 * memory; tiles holds (descriptor index, first row, first column) triples of 32 x 32 tiles. */
#define SKG_BLOCK_SRC_BF16 1
#define SKG_BLOCK_DST_BF16 2
#define SKG_BLOCK_ACCUMULATE 4   /* fp32 dst += (atomic) instead of = */
#define SKG_BLOCK_SCALE 8        /* multiply by `scale` */
typedef struct skg_block_desc {
  const void* src; void* dst; void* dst_t;      /* dst_t (optional): bf16 [cols, ld_dst_t] transposed copy */
  int32_t rows, cols;
  int64_t ld_src, ld_dst, ld_dst_t;
  int32_t flags; float scale;
} skg_block_desc;
int skg_block_copy(const skg_block_desc* descs_dev, const int32_t* tiles_dev, int num_tiles, cudaStream_t stream);

/* ---- the whole pair path in one call (csrc/forward.cu) ------------------------------------------
 * GraphHead.forward + predictor / suppressor for a batch (head.py:769-993, 408-411): pair
 * enumeration + spatial encoding, spatial_head, box_head, fc_head / fc_tail on the distinct node rows
 * with the TransH entity rows, the fused fc_2 GEMM of the four MBFs, adjacency, message aggregation,
 * LayerNorm, the final attention heads on the valid pairs and the logits.  The host mirror packs
 * the weights once (engine.PackedHead) and hands raw DEVICE pointers; the library owns no weights.
 * All intermediates live in ONE caller-allocated workspace whose layout skg_head_plan reports
 * (so the caller can read any tap: enc, x_keep, adjacency, ...); no allocation, no sync.
 * The first two buffers ("splitk_main", "splitk_side", fixed size, at the start of the workspace) are the
 * split-K scratch of the two streams: they must be ZERO before the first call on a workspace and every call
 * leaves them zero (see skg_gemm_bf16_splitk). */
typedef struct skg_head_weights {
  int32_t R, E, K0, pool_k, fc_k, Kg, Cg, num_cls, n_logit, n_logit_pad, human_idx, ent_tot, ent_dim, sp_in;
  const void* Wb1; const float* bb1;          /* box_head.1  [E, K0] bf16 (K padded)   head.py:637 */
  const void* Wb2; const float* bb2;          /* box_head.3  [E, E]                    head.py:639 */
  const float* Wsp0; const float* bsp0;       /* spatial_head.0 [128, 46] fp32         head.py:663 */
  const void* Wsp1; const float* bsp1;        /* spatial_head.2 [256, 128]             head.py:665 */
  const void* Wsp2; const float* bsp2;        /* spatial_head.4 [R, 256]               head.py:667 */
  const void* Wfh; const float* bfh;          /* fc_head.0 [R, fc_k]                   head.py:694 */
  const void* Wft; const float* bft;          /* fc_tail.0 [R, fc_k]                   head.py:698 */
  const void* Wfc; const float* bfc;          /* fused [2R, fc_k]: fc_head rows (entity columns zeroed, their constant
                                                 E[human_idx] term folded into bfc) then fc_tail rows */
  const void* W_hproj; const float* b_hproj;  /* [2R, R]: attention_head.fc_1[:, :E] | sub_to_obj.fc_1 */
  const void* W_oproj; const float* b_oproj;  /* [2R, R]: attention_head.fc_1[:, E:] | obj_to_sub.fc_1 */
  const void* Wg1; const float* bg1;          /* attention_head_g.fc_1 [R, Kg] */
  const void* W2cat; const float* b2cat;      /* fc_2 of attention_head | obj_to_sub | sub_to_obj | attention_head_g  [4R, R] */
  const void* W3att; const float* b3att;      /* fc_3 (branch-concatenated) [R, R] */
  const void* W3os; const float* b3os;
  const void* W3so; const float* b3so;
  const void* W3g; const float* b3g;
  const float* w_adj; float b_adj;            /* adjacency Linear(R -> 1) head.py:644 */
  const float* ln_h_w; const float* ln_h_b; const float* ln_o_w; const float* ln_o_b; float ln_eps;
  const void* Wpred; const float* bpred;      /* [n_logit_pad, 2R] predictor rows then the suppressor row (models.py:176-177) */
  const void* Wpred_hh; const void* Wpred_lo; /* split-bf16 predictor operands ([hi | hi] and lo), NULL -> single bf16 pass */
  const float* ent;                           /* TransH entity table [ent_tot, ent_dim] (head.py:574-578) */
  const void* Wpred_hhl;                      /* [n_logit_pad, 6R] = [W_hi | W_hi | W_lo]: the split-bf16 predictor as one launch
                                                 against the cyclically read [hi | lo] features; NULL -> the two-launch form */
  const float* Wfh_ent;                       /* fc_head.0.weight[:, E:] fp32 [R, ent_dim]: when set, every forward re-folds
                                                 bfc[:R] = bfh[:R] + Wfh_ent @ ent[human_idx] ON THE DEVICE (bfc is then
                                                 written by the library), so the constant human-side entity term follows the
                                                 table's current contents -- also under CUDA-graph replay */
} skg_head_weights;

typedef struct skg_head_batch {
  int32_t num_images, total_boxes, total_humans, max_boxes_per_image, feat_hw;
  int64_t total_cells, total_pairs;
  int64_t seg_bytes;          /* room to reserve in the workspace for the packed segment table (buffer "seg"; may be 0) */
  const float* features3;     /* [B, Cg, feat_hw]  features['3'] (head.py:811) */
  const float* box_features;  /* [boxes, pool_k]   pooled box features (head.py:387) */
  const float* boxes;         /* [boxes, 4] */
  const int32_t* box_off; const int32_t* n_h; const int32_t* hum_off;   /* segment table, as skg_pairs_spatial_encode */
  const int64_t* cell_off; const int64_t* pair_off; const float* hw;
  const void* box_features_bf16;  /* optional [boxes, K0] bf16 rows (K0 = pool_k rounded up to 8, zero padded): the pooled
                                     features already in the first GEMM's operand layout, e.g. written by
                                     skg_multiscale_roi_align; when set, box_features is ignored and may be NULL */
  void* maps_ready_event;         /* optional cudaEvent_t recorded right after the pair enumeration (the FIRST kernel of the
                                     step): x / y / x_keep / y_keep / cell and pair maps are then final, so work that needs only
                                     the batch geometry (GT association, loss normalisers, TransH selection: head.py:926-963)
                                     can run on another stream under the GEMM chain; NULL = none */
} skg_head_batch;

#define SKG_HEAD_NUM_BUFFERS 96
typedef struct skg_head_plan_t {
  int64_t total_bytes;
  int32_t num_buffers;
  int64_t offset[SKG_HEAD_NUM_BUFFERS];     /* byte offset inside the workspace (256-byte aligned) */
  int64_t rows[SKG_HEAD_NUM_BUFFERS];
  int64_t cols[SKG_HEAD_NUM_BUFFERS];       /* row pitch in elements */
  int32_t elem_size[SKG_HEAD_NUM_BUFFERS];  /* 2 = bf16, 4 = fp32 / int32, 8 = int64 */
} skg_head_plan_t;

/* name of workspace buffer i ("enc", "x_keep", "logits", ...) or NULL past the end */
const char* skg_head_buffer_name(int i);
/* sizes only (pointers of `batch` are not read) */
int skg_head_plan(const skg_head_weights* weights, const skg_head_batch* batch, int flags, skg_head_plan_t* plan);
#define SKG_HEAD_PRECISE_LOGITS 1   /* split-bf16 predictor (3 bf16 passes, ~fp32 logits) */
#define SKG_HEAD_WANT_TAPS 2        /* also keep fp32 copies of the refined node encodings */
#define SKG_HEAD_GRAPH 8            /* replay the step as a CUDA graph when the same (weights, batch incl. input pointers,
                                       workspace) key repeats; any change falls back to plain launches */
#define SKG_HEAD_TRAIN 16           /* training forward (skg_head_backward follows): keeps the raw fc_2 outputs of all four MBFs
                                       and the attention weights, the pair features and logits stay in CELL rows (buffers
                                       "feat" [cells, 2R], "logits" [cells, n_logit]; the diagonal cells i == j are masked by
                                       the loss), single-pass predictor */
#define SKG_HEAD_PROFILE 4          /* record an event after every launch and SYNCHRONISE at the end (diagnostics only) */
#define SKG_HEAD_CAPACITY 32        /* the sizes in skg_head_batch are CAPACITIES (skg_head_capacity): every launch is sized by
                                       them, the true counts (box_off[B], hum_off[B], cell_off[B]) and the addresses of
                                       features3 / box_features / boxes are read on the device from the packed segment table
                                       (skg_segments_pack layout incl. its address block) that the caller uploads into
                                       workspace buffer "seg" before every step.  Rows beyond the true counts are padding
                                       (zero encodings, cell2pair = -1, never read back).  With SKG_HEAD_GRAPH one captured
                                       graph then serves EVERY batch that fits the capacity bucket, whatever its per-image
                                       detection counts and input addresses (real detection streams change them every
                                       batch, head.py:826-841).  Inference only */
/* capacity bucket of a batch: cells rounded up to a GEMM row tile (128 up to 2048 cells, then <= 6.25 % slack), node rows to
 * 128, pairs = cells; the input addresses of `cap` are cleared (they travel in the segment table) */
int skg_head_capacity(const skg_head_weights* weights, const skg_head_batch* exact, skg_head_batch* cap);
/* launch i of the last profiled skg_head_forward on this thread: label, stream (0 caller's, 1 side), microseconds since
 * the previous event on that stream; returns -1 past the end */
int skg_head_profile_get(int i, const char** name, int* stream_id, float* us);
/* graph-cache counters of this thread: replays, captures, first-sighting plain launches */
void skg_head_graph_stats(long long* replays, long long* captures, long long* plain);
int skg_head_forward(const skg_head_weights* weights, const skg_head_batch* batch, int flags, void* workspace,
                     int64_t workspace_bytes, cudaStream_t stream);

/* n_p of the two focal losses from the GT labels alone (head.py:167-205: number of positive entry labels, number of pairs
 * with any positive label) -- what `(labels != 0).sum()` / `(unary_labels != 0).sum()` give after skg_head_post, computable
 * before the GEMM chain has produced a single logit.  gt_labels [pairs, num_cls], out [2] fp32 (zeroed here). */
int skg_loss_normalisers(const float* gt_labels, const skg_post_tables* tables, int64_t pairs, int num_cls, float* out,
                         cudaStream_t stream);

/* TransH positive / negative selection of the training loss on the device (head.py:926-963): per image b -- the label rows
 * [pair_off[b], pair_off[b+1]) of gt_labels [P, num_cls] -- skg_select_count writes the number of positive labels to
 * n_pos[b] and their total to n_pos[num_images]; after the caller has sized the outputs from that total, skg_select_fill
 * writes the positives' flat indices (p * num_cls + verb, flat order, image by image) and, aligned with them, as many
 * zero-label flat indices of the SAME image drawn uniformly without replacement (the distribution of
 * randperm(#zeros)[:#pos], head.py:939; counter-hash generator seeded by `seed`). */
int skg_select_count(const float* gt_labels, const int64_t* pair_off, int num_images, int num_cls, int32_t* n_pos,
                     cudaStream_t stream);
int skg_select_fill(const float* gt_labels, const int64_t* pair_off, int num_images, int num_cls, const int32_t* n_pos,
                    unsigned long long seed, int64_t* pos_flat, int64_t* neg_flat, cudaStream_t stream);

/* ---- per-head serving session (csrc/session.cu) ----------------------------------------------------------------------------
 * The opaque handle of SURVEY 8(b): it owns a copy of the packed-weight descriptor (the bf16 operands stay owned by the
 * caller and must outlive it), the verb table, a pinned staging ring for the per-step segment table and the workspace plans
 * of the capacity buckets it has seen.  One inference step of InteractionHead.forward from the pooled box features on
 * (head.py:389-427) is
 *     skg_session_prepare   host only: humans-first check (head.py:840-841), segment table, capacity bucket, result sizes
 *     (the caller sizes the workspace and the two flat outputs from `sizes`)
 *     skg_session_run       1 H2D copy (segment table + input addresses), the bucket's CUDA graph, 2 post-processing launches
 * whatever the batch geometry and the input addresses are.  One session per host thread and stream. */
typedef struct skg_session skg_session;
typedef struct skg_session_config {
  skg_head_weights weights;
  const int32_t* verb_len;        /* HOST   [num_obj_cls] valid verbs per object class (object_class_to_target_class) */
  const int32_t* verb_off_dev;    /* DEVICE [num_obj_cls + 1] CSR of the same table */
  const int32_t* verb_idx_dev;    /* DEVICE [nnz] */
  int32_t num_obj_cls;
  int32_t max_images;             /* largest batch a step may carry */
  int32_t flags;                  /* SKG_HEAD_PRECISE_LOGITS or 0 */
  float prior_power;              /* 2.8 in inference (head.py:742) */
} skg_session_config;
typedef struct skg_step_sizes {
  int64_t boxes, humans, cells, pairs, entries;   /* true counts of the step; entries = sum over images of nonzero(prior[0]) */
  int64_t cap_boxes, cap_humans, cap_cells;       /* its capacity bucket */
  int64_t workspace_bytes;                        /* the bucket's workspace (256-byte aligned base, zero-filled on first use) */
  int64_t out_i64, out_f32;                       /* elements of the two flat outputs of skg_session_run */
} skg_step_sizes;
typedef struct skg_step_io {
  const float* features3;          /* [B, Cg, feat_hw] */
  const float* box_features;       /* [boxes, pool_k] fp32, or NULL with box_features_bf16 */
  const void* box_features_bf16;   /* [cap_boxes, K0] bf16 operand rows (skg_multiscale_roi_align), or NULL */
  const float* boxes;              /* [boxes, 4] */
  const int64_t* labels;           /* [boxes] */
  const float* scores;             /* [boxes] */
  const float* ent;                /* TransH entity table [ent_tot, ent_dim] */
  int64_t* out_i64;                /* [index entries | prediction entries | object pairs] */
  float* out_f32;                  /* [boxes_h 4*pairs | boxes_o 4*pairs | scores entries | prior 2*entries | weights pairs] */
} skg_step_io;
int skg_session_create(const skg_session_config* config, skg_session** session);
void skg_session_destroy(skg_session* session);
/* all HOST pointers: n [B] boxes per image, labels / scores [sum n] (humans first in every image), hw [B, 2] image sizes as
 * doubles; pairs_per_image / entries_per_image [B] may be NULL.  Errors: "index out of range in self" when an image has more
 * boxes than the entity table has rows (head.py:572), "Human detections are not permuted to the top" (head.py:840-841) */
int skg_session_prepare(skg_session* session, int num_images, const int64_t* n, const int64_t* labels, const float* scores,
                        const double* hw, int feat_hw, skg_step_sizes* sizes, int64_t* pairs_per_image, int64_t* entries_per_image);
int skg_session_run(skg_session* session, const skg_step_io* io, void* workspace, int64_t workspace_bytes, cudaStream_t stream);
/* workspace plans of the prepared step: of its capacity bucket (offsets) and of its exact sizes (row counts) */
int skg_session_plans(const skg_session* session, skg_head_plan_t* capacity_plan, skg_head_plan_t* exact_plan);
/* tensor-map descriptor cache of this thread (csrc/gemm.cu): lookups served from the cache, cuTensorMapEncodeTiled calls */
void skg_tensor_map_cache_stats(long long* hits, long long* encodes);

/* ---- backward of the whole pair path (csrc/backward.cu): the training step of BASELINE configs[2] ------------------------
 * Reference: loss.backward() through GraphHead.forward / the predictor and the two focal losses (utils.py:213-229,
 * head.py:153-205, 769-993).  Call order per step: skg_head_forward(flags | SKG_HEAD_TRAIN) -> skg_head_post on the cell-row
 * logits (logit rows = pair2cell) -> skg_head_backward with the SAME weights / batch / workspace.  Every dX and dW
 * product is a launch of skg_gemm_bf16 (dX: the transposed weight copies of skg_head_weights_t as B operand, the ReLU gate
 * of the layer's input fused into the epilogue; dW = G^T X on M-major copies made by skg_transpose_bf16, whose column sums
 * are the bias gradients), the element-wise and per-node reductions are the kernels of csrc/trainops.cu.
 * `grads` mirrors the packed operand layout of skg_head_weights in fp32; the caller ZEROES it before the call (bias and
 * per-node gradients are accumulated with atomics, weight gradients are stored).  skg_block_copy maps it back onto the
 * reference's per-branch parameter layout. */
typedef struct skg_head_weights_t {     /* transposed bf16 copies: B operands of the dX GEMMs (K-major = N of the forward) */
  const void* Wb2T;      /* [E, E] */
  const void* Wsp1T;     /* [128, 256] */
  const void* Wsp2T;     /* [256, R] */
  const void* WfcT;      /* [fc_k, 2R] */
  const void* W_hprojT;  /* [R, 2R] */
  const void* W_oprojT;  /* [R, 2R] */
  const void* W2catT;    /* [R, 4R] */
  const void* W3attT; const void* W3osT; const void* W3soT; const void* W3gT;   /* [R, R] */
  const void* WpredT;    /* [2R, n_logit_pad] */
} skg_head_weights_t;

typedef struct skg_head_grads {         /* fp32, shapes of the packed operands in skg_head_weights */
  float* Wb1; float* bb1; float* Wb2; float* bb2; float* Wsp0; float* bsp0; float* Wsp1; float* bsp1; float* Wsp2; float* bsp2;
  float* Wfc; float* bfc; float* W_hproj; float* b_hproj; float* W_oproj; float* b_oproj; float* Wg1; float* bg1;
  float* W2cat; float* b2cat; float* W3att; float* b3att; float* W3os; float* b3os; float* W3so; float* b3so; float* W3g; float* b3g;
  float* w_adj; float* b_adj; float* ln_h_w; float* ln_h_b; float* ln_o_w; float* ln_o_b; float* Wpred; float* bpred;
} skg_head_grads;

typedef struct skg_head_loss_args {
  const skg_post_tables* tables;   /* HOST struct of DEVICE pointers, as skg_head_post */
  const float* gt_labels;          /* [P, num_cls] from skg_associate_gt */
  const float* n_p;                /* [2] normalisers of hoi_loss / interactiveness_loss (rank-averaged when distributed) */
  const float* g_hoi; const float* g_int;   /* upstream gradients of the two losses (device scalars, NULL = 1) */
  float prior_power;               /* 1.0 in training (head.py:742) */
} skg_head_loss_args;

int skg_head_backward(const skg_head_weights* weights, const skg_head_weights_t* weights_t, const skg_head_batch* batch,
                      const skg_head_loss_args* loss, const skg_head_grads* grads, int flags, void* workspace,
                      int64_t workspace_bytes, cudaStream_t stream);
/* The same with stage events for data-parallel training (reference: DDP's bucketed all-reduce overlapping backward,
 * utils.py:200-229): stage_events[k] (cudaEvent_t handles, NULL entries skipped) is recorded on `stream` when the gradients of
 * group k are final, so the caller can start that group's all-reduce on another stream while the backward continues:
 *   0: Wpred bpred W3g b3g Wg1 bg1 ln_* W3os b3os W3so b3so w_adj b_adj W3att b3att
 *   1: W_hproj b_hproj W_oproj b_oproj Wfc bfc Wb2 bb2 Wb1 bb1   (the node-side chain incl. box_head: 69 % of the bytes; it
 *      runs BEFORE the cell-side chain so that its all-reduce is covered)
 *   2: W2cat b2cat                                               (fused fc_2 of the four MBFs, 16.8 MB)
 *   3: Wsp2 bsp2 Wsp1 bsp1 Wsp0 bsp0                             (spatial head, 1.2 MB: the end of the backward) */
/* flags | SKG_HEAD_GRAPH: the launch sequence is replayed as a CUDA graph when every byte that reaches a launch repeats
 * (weights, batch geometry, loss-argument / gradient addresses, workspace, stage events): first sighting plain launches,
 * second capture, then replays; skg_head_backward_graph_stats counts them for this thread */
void skg_head_backward_graph_stats(long long* replays, long long* captures, long long* plain);
#define SKG_HEAD_BACKWARD_STAGES 4
int skg_head_backward_staged(const skg_head_weights* weights, const skg_head_weights_t* weights_t, const skg_head_batch* batch,
                             const skg_head_loss_args* loss, const skg_head_grads* grads, int flags, void* workspace,
                             int64_t workspace_bytes, cudaStream_t stream, void* const* stage_events, int num_stage_events);

/* Host-side packing of the per-batch segment table every batched kernel consumes (all HOST pointers): n / n_h [B]
 * boxes and humans per image, hw [B,2] = (h, w) as doubles.  `out` receives
 *   cell_off i64[B+1] | pair_off i64[B+1] | box_off i32[B+1] | n_h i32[B] | hum_off i32[B+1] | hw f32[3B] (h, w, h*w)
 *   | pad to 16 | addr u64[4]  (device addresses of features3, fp32 box_features, boxes, 0: zeroed here, filled in by the
 *   caller for SKG_HEAD_CAPACITY)
 * (n_h forced to 0 for images the reference skips, head.py:829-839); offsets[7] = byte offsets of the seven arrays,
 * totals[5] = boxes, humans, cells, pairs, max boxes per image.  Returns the bytes used (multiple of 16) or -1. */
int64_t skg_segments_pack(const int64_t* n, const int64_t* n_h, const double* hw, int num_images, void* out,
                          int64_t out_bytes, int64_t* offsets, int64_t* totals);

/* Host-side closed form of the number of (pair, verb) entries postprocess emits per image
 * (head.py:315 nonzero(prior[0])): all arrays are HOST pointers; returns the total, or -1. */
int64_t skg_post_count_host(const int32_t* verb_len, int num_obj_cls, const int64_t* labels, const float* scores,
                            const int64_t* box_off, const int64_t* n_h, int num_images, int64_t* img_entries);

/* box_roi_pool of the interaction head (torchvision MultiScaleRoIAlign(['0','1','2','3'], 7, 2), built at models.py:158-162
 * and called at head.py:387) for a whole batch: FPN level per box = clamp(floor(canonical_level + log2(sqrt(area) /
 * canonical_scale) + 1e-6), k_min, k_max) - k_min (torchvision poolers.LevelMapper), then roi_align (aligned = False,
 * sampling_ratio^2 bilinear samples per bin) on that level's map.
 * feats / feat_h / feat_w / scales are HOST arrays of num_levels entries: device pointers of the [B, C, H_l, W_l] fp32
 * maps, their sizes and spatial scales.  boxes [sum n, 4] fp32 xyxy in image coordinates (device, 16-byte aligned),
 * box_off [B+1] int32 (device) the per-image prefix sum.  Outputs (either may be NULL): out_f32 [sum n, C, pooled,
 * pooled] fp32, the reference layout; out_bf16 [sum n, ld_bf16] bf16, the same block rounded to bf16 in the K-major
 * layout box_head's first GEMM reads (skips the fp32 tensor and the cast). */
int skg_multiscale_roi_align(const float* const* feats, const int32_t* feat_h, const int32_t* feat_w, const float* scales,
                             int num_levels, int channels, int num_images, const float* boxes, const int32_t* box_off,
                             int64_t num_boxes, int pooled, int sampling_ratio, int k_min, int k_max, float canonical_scale,
                             int canonical_level, float* out_f32, void* out_bf16, int64_t ld_bf16, cudaStream_t stream);

/* InteractionHead.preprocess (head.py:92-151) for a whole batch, on the device: score >= box_score_thresh filter,
 * class-wise NMS exactly as torchvision.ops.batched_nms computes it for small inputs (boxes shifted by
 * label * (max active coordinate + 1), IoU = inter / (Sa + Sb - inter) > thresh, stable score order), survivors in
 * descending score order with the human_idx detections first, at most max_human humans + max_object others.
 * boxes [sum N, 4] fp32 xyxy (16-byte aligned), labels [sum N] int64, scores [sum N] fp32 are the images' detections
 * concatenated (ground-truth boxes already prepended in training, head.py:105-116); det_off [B+1] int32 their prefix
 * sum; max_dets_per_image sizes the shared-memory NMS.  out_index [B, max_human + max_object] int64: GLOBAL rows of the
 * kept detections of image b in final order, padded with -1; out_counts [B, 2] int32 = (n, n_h). */
int skg_preprocess_detections(const float* boxes, const int64_t* labels, const float* scores, const int32_t* det_off,
                              int num_images, int max_dets_per_image, float box_score_thresh, float box_nms_thresh,
                              int human_idx, int max_human, int max_object, int64_t* out_index, int32_t* out_counts,
                              cudaStream_t stream);

/* Negative sampler of the KG trainer (OpenKE sampling() / getBatch(), OpenKE/openke/base/Base.cpp:78-197, with the
 * filtered corrupt_head / corrupt_tail of Corrupt.h:7-83 and the per-thread LCG of Random.h:18-28), bit-exact with the
 * reference for any `work_threads`.  The index arrays are what importTrainFiles builds (Reader.h:52-160), on the device:
 * head_hrt / tail_hrt [3, T] int64 = rows h | r | t of the de-duplicated training triples sorted by (h, r, t) resp.
 * (t, r, h); lef_* / rig_* [E] the entity ranges in those orderings; left_mean / right_mean [R] fp32 for the "bern"
 * head/tail coin (both NULL: fair coin, prob = 500/1000).  seeds_in [work_threads] = next_random[] before the call,
 * seeds_out (distinct buffer) after it.  Outputs batch_h / batch_t / batch_r int64 and batch_y fp32 of length
 * batch_size * (1 + neg_rate): positives first, then neg_rate blocks of negatives (batch_y = 1 / -1).
 * mode 0: corrupt head or tail per the coin; -1: always the head; +1: always the tail.  negRelRate is not supported.
 * Where the reference divides by zero -- (entity, relation) linked to EVERY entity -- the positive is kept. */
int skg_kg_sample(const int64_t* head_hrt, const int64_t* tail_hrt, const int64_t* lef_head, const int64_t* rig_head,
                  const int64_t* lef_tail, const int64_t* rig_tail, const float* left_mean, const float* right_mean,
                  int64_t train_total, int64_t ent_total, const uint64_t* seeds_in, uint64_t* seeds_out, int work_threads,
                  int64_t batch_size, int neg_rate, int mode, int64_t* batch_h, int64_t* batch_t, int64_t* batch_r,
                  float* batch_y, cudaStream_t stream);

/* TransH knowledge-graph training as ONE persistent cooperative kernel (csrc/kgtrain.cu): `num_steps` optimiser steps of
 * the reference loop OpenKE/openke/config/Trainer.py:43-99 -- negative sampling (the sampler above: same per-worker LCG
 * streams, same batches), TransH scores (heads/TransH/TransH.py:56-106), MarginLoss without the adversarial weighting
 * (heads/MarginLoss.py:28-36), NegativeSampling strategy with regul_rate = 0, backward, plain SGD -- per launch, the phases of
 * a step separated by grid-wide barriers; the three tables and their gradients live in shared memory (graphs whose tables
 * exceed 100 KB take the eager Trainer).  seeds: [2, work_threads], the seeds of the first step in row 0; after the call the
 * current seeds sit in row num_steps & 1.  ent / rel / nrm are updated in place; loss_out[s] = the loss of step s. */
typedef struct skg_kg_train_args {
  const int64_t* head_hrt; const int64_t* tail_hrt;       /* [3, T] rows h | r | t in head order / tail order (skg_kg_sample) */
  const int64_t* lef_head; const int64_t* rig_head; const int64_t* lef_tail; const int64_t* rig_tail;   /* [E] */
  const float* left_mean; const float* right_mean;         /* [R] bern statistics, or both NULL */
  int64_t train_total, ent_total, rel_total;
  uint64_t* seeds;
  int32_t work_threads; int32_t neg_rate; int32_t mode; int32_t dim; int32_t p_norm; int32_t norm_flag; int32_t num_steps; int32_t pad_;
  int64_t batch_size;
  float* ent; float* rel; float* nrm;                       /* [E, dim], [R, dim], [R, dim] fp32 */
  float margin, lr;
  float* loss_out;                                          /* [num_steps] */
  void* scratch; int64_t scratch_bytes;                     /* >= skg_kg_train_scratch_bytes, 256-byte aligned */
} skg_kg_train_args;
int64_t skg_kg_train_scratch_bytes(int64_t batch_size, int neg_rate, int64_t ent_total, int64_t rel_total, int dim);
int skg_kg_train_steps(const skg_kg_train_args* args, cudaStream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* SKGHOI_B200_H_ */
