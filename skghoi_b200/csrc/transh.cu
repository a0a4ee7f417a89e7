// TransH knowledge-graph scoring, its backward, and the margin loss
// (reference heads/TransH/TransH.py:56-106, heads/MarginLoss.py:28-36, heads/NegativeSampling.py:30-63).
//
// Bandwidth-bound gather kernel: the three embedding tables (80x50 + 2x117x50 fp32 = 62.8 KB for
// the HOI label graph) are staged once per CTA in shared memory; one warp handles one triple at a
// time with the embedding dimension spread over the lanes, so every global access (index reads,
// the optional materialised h_, r, w, t_ rows, the score) is a coalesced warp-wide transaction and
// the per-triple reductions are register shuffles.  Tables that do not fit in shared memory
// (large public KGs) are read through L2 instead.
#include "common.cuh"
#include "skghoi_b200.h"
#include "transh.cuh"

namespace skg {

constexpr int kTranshThreads = 256;
constexpr int kTranshWarps = kTranshThreads / 32;

// d(tables) += d(score) * d score / d tables.  Gradients accumulate into fp32 tables with atomics;
// when the tables fit they are privatised per CTA in shared memory and flushed once.
template <int EPL, bool SMEM_TABLES>
__global__ void __launch_bounds__(kTranshThreads) transh_backward_kernel(
    const float* __restrict__ ent, const float* __restrict__ rel, const float* __restrict__ nrm, int ent_tot, int rel_tot,
    int dim, TripleIndex idx, int64_t T, int p_norm, int norm_flag, const float* __restrict__ dscore,
    float* __restrict__ d_ent, float* __restrict__ d_rel, float* __restrict__ d_nrm) {
  extern __shared__ __align__(16) float smem[];
  float* gE = d_ent; float* gR = d_rel; float* gW = d_nrm;
  const int nE = ent_tot * dim, nR = rel_tot * dim;
  if (SMEM_TABLES) {
    gE = smem; gR = gE + nE; gW = gR + nR;
    for (int e = threadIdx.x; e < nE + 2 * nR; e += blockDim.x) smem[e] = 0.f;
    __syncthreads();
  }
  const int lane = threadIdx.x & 31;
  const int64_t warp = (int64_t)blockIdx.x * kTranshWarps + (threadIdx.x >> 5);
  const int64_t nwarps = (int64_t)gridDim.x * kTranshWarps;
  for (int64_t t = warp; t < T; t += nwarps) {
    const float g = __ldg(dscore + t);
    if (g == 0.f) continue;    // warp-uniform
    int64_t hi, ri, ti;
    idx.get(t, hi, ri, ti);
    float h[EPL], tl[EPL], r[EPL], w[EPL], s[EPL];
#pragma unroll
    for (int e = 0; e < EPL; ++e) {
      const int d = lane + 32 * e;
      const bool in = d < dim;
      h[e] = in ? __ldg(ent + hi * dim + d) : 0.f; tl[e] = in ? __ldg(ent + ti * dim + d) : 0.f;
      r[e] = in ? __ldg(rel + ri * dim + d) : 0.f; w[e] = in ? __ldg(nrm + ri * dim + d) : 0.f;
    }
    TranshTriple st;
    transh_triple_fwd<EPL>(h, tl, r, w, p_norm, norm_flag, s, st);
    float g_h[EPL], g_t[EPL], g_r[EPL], g_w[EPL];
    transh_triple_bwd<EPL>(h, tl, r, w, s, st, g, p_norm, norm_flag, g_h, g_t, g_r, g_w);
#pragma unroll
    for (int e = 0; e < EPL; ++e) {
      const int d = lane + 32 * e;
      if (d < dim) {
        atomicAdd(gE + hi * dim + d, g_h[e]);
        atomicAdd(gE + ti * dim + d, g_t[e]);
        atomicAdd(gR + ri * dim + d, g_r[e]);
        atomicAdd(gW + ri * dim + d, g_w[e]);
      }
    }
  }
  if (SMEM_TABLES) {
    __syncthreads();
    for (int e = threadIdx.x; e < nE; e += blockDim.x) { const float v = smem[e]; if (v != 0.f) atomicAdd(d_ent + e, v); }
    for (int e = threadIdx.x; e < nR; e += blockDim.x) {
      const float v = smem[nE + e]; if (v != 0.f) atomicAdd(d_rel + e, v);
      const float u = smem[nE + nR + e]; if (u != 0.f) atomicAdd(d_nrm + e, u);
    }
  }
}

// MarginLoss.forward (MarginLoss.py:28-36): p [rows] broadcast against n [rows, cols].
//   plain: mean_{rows,cols}(max(p - n, -margin)) + margin
//   adv  : mean_rows( sum_cols softmax_cols(-n*T) * max(p - n, -margin) ) + margin  (weights detached)
// One CTA per row chunk; partial sums are reduced with one atomicAdd per CTA into loss[0], which the
// host wrapper zeroes first and then finalises (+margin) with a tiny second kernel.
__global__ void margin_loss_kernel(const float* __restrict__ p, const float* __restrict__ n, int64_t rows, int64_t cols,
                                   float margin, float adv_t, int use_adv, float* __restrict__ loss_acc,
                                   float* __restrict__ d_p, float* __restrict__ d_n, float gscale) {
  __shared__ float red[32];
  float local = 0.f;
  for (int64_t row = blockIdx.x; row < rows; row += gridDim.x) {
    const float pv = __ldg(p + row);
    const float* nr = n + row * cols;
    float mx = -INFINITY, den = 1.f;
    if (use_adv) {   // softmax over the row
      for (int64_t c = threadIdx.x; c < cols; c += blockDim.x) mx = fmaxf(mx, -nr[c] * adv_t);
      mx = warp_max(mx);
      if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = mx;
      __syncthreads();
      mx = -INFINITY;
      for (int i = 0; i < (blockDim.x >> 5); ++i) mx = fmaxf(mx, red[i]);
      __syncthreads();
      float s = 0.f;
      for (int64_t c = threadIdx.x; c < cols; c += blockDim.x) s += expf(-nr[c] * adv_t - mx);
      s = warp_sum(s);
      if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
      __syncthreads();
      den = 0.f;
      for (int i = 0; i < (blockDim.x >> 5); ++i) den += red[i];
      __syncthreads();
    }
    float dp = 0.f;
    for (int64_t c = threadIdx.x; c < cols; c += blockDim.x) {
      const float nv = nr[c];
      const float diff = pv - nv;
      const bool active = diff > -margin;          // torch.max(a, b): gradient to a where a > b
      const float wgt = use_adv ? expf(-nv * adv_t - mx) / den : 1.f;
      local += wgt * fmaxf(diff, -margin);
      if (d_n) {
        const float gsc = active ? gscale * wgt : 0.f;
        d_n[row * cols + c] = -gsc;
        dp += gsc;
      }
    }
    if (d_p) {
      dp = warp_sum(dp);
      if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = dp;
      __syncthreads();
      if (threadIdx.x == 0) { float a = 0.f; for (int i = 0; i < (blockDim.x >> 5); ++i) a += red[i]; d_p[row] = a; }
      __syncthreads();
    }
  }
  local = warp_sum(local);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = local;
  __syncthreads();
  if (threadIdx.x == 0) {
    float a = 0.f;
    for (int i = 0; i < (blockDim.x >> 5); ++i) a += red[i];
    atomicAdd(loss_acc, a);
  }
}

__global__ void margin_loss_finalize(float* loss, float inv_count, float margin) { loss[0] = loss[0] * inv_count + margin; }

template <int EPL>
static int launch_transh(bool backward, const float* ent, const float* rel, const float* nrm, int ent_tot, int rel_tot,
                         int dim, const TripleIndex& idx, int64_t T, int p_norm, int norm_flag, float* h_out,
                         float* r_out, float* w_out, float* t_out, float* score, const float* dscore, float* d_ent,
                         float* d_rel, float* d_nrm, cudaStream_t stream) {
  const size_t table_bytes = ((size_t)ent_tot + 2 * (size_t)rel_tot) * dim * sizeof(float);
  const bool fits = table_bytes <= 96 * 1024;
  const int64_t want = ceil_div64(T, kTranshWarps * 4);
  const int per_sm = fits ? (table_bytes > 48 * 1024 ? 2 : 4) : 8;
  int grid = (int)(want < (int64_t)num_sms() * per_sm ? want : (int64_t)num_sms() * per_sm);
  if (grid < 1) grid = 1;
  if (fits) {
    auto k = transh_backward_kernel<EPL, true>;
    SKG_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)table_bytes));
    k<<<grid, kTranshThreads, table_bytes, stream>>>(ent, rel, nrm, ent_tot, rel_tot, dim, idx, T, p_norm, norm_flag,
                                                     dscore, d_ent, d_rel, d_nrm);
  } else {
    transh_backward_kernel<EPL, false><<<grid, kTranshThreads, 0, stream>>>(ent, rel, nrm, ent_tot, rel_tot, dim, idx, T,
                                                                            p_norm, norm_flag, dscore, d_ent, d_rel, d_nrm);
  }
  return check_launch("transh_backward_kernel");
}

static int dispatch_transh(bool backward, const float* ent, const float* rel, const float* nrm, int ent_tot, int rel_tot,
                           int dim, const TripleIndex& idx, int64_t T, int p_norm, int norm_flag, float* h_out,
                           float* r_out, float* w_out, float* t_out, float* score, const float* dscore, float* d_ent,
                           float* d_rel, float* d_nrm, cudaStream_t stream) {
  SKG_REQUIRE(dim >= 1 && dim <= 256, "TransH dim must be in [1, 256], got %d", dim);
  SKG_REQUIRE(p_norm == 1 || p_norm == 2, "p_norm must be 1 or 2, got %d", p_norm);
  SKG_REQUIRE(ent_tot > 0 && rel_tot > 0 && T >= 0, "bad table sizes");
  if (T == 0) return SKG_OK;
  SKG_REQUIRE(ent && rel && nrm, "null table pointer");
#define SKG_TRANSH_CASE(E)                                                                                          \
  return launch_transh<E>(backward, ent, rel, nrm, ent_tot, rel_tot, dim, idx, T, p_norm, norm_flag, h_out, r_out, \
                          w_out, t_out, score, dscore, d_ent, d_rel, d_nrm, stream)
  if (dim <= 32) SKG_TRANSH_CASE(1);
  if (dim <= 64) SKG_TRANSH_CASE(2);
  if (dim <= 128) SKG_TRANSH_CASE(4);
  SKG_TRANSH_CASE(8);
#undef SKG_TRANSH_CASE
}

}  // namespace skg

using namespace skg;

extern "C" int skg_transh_backward(const float* ent, const float* rel, const float* nrm, int ent_tot, int rel_tot,
                                   int dim, const int64_t* head, const int64_t* relation, const int64_t* tail,
                                   int64_t num_triples, int p_norm, int norm_flag, const float* dscore, float* d_ent,
                                   float* d_rel, float* d_nrm, cudaStream_t stream) {
  SKG_REQUIRE(num_triples == 0 || (head && relation && tail && dscore && d_ent && d_rel && d_nrm), "null pointer");
  TripleIndex idx{head, relation, tail, nullptr, 0, rel_tot, 0};
  return dispatch_transh(true, ent, rel, nrm, ent_tot, rel_tot, dim, idx, num_triples, p_norm, norm_flag, nullptr,
                         nullptr, nullptr, nullptr, nullptr, dscore, d_ent, d_rel, d_nrm, stream);
}

extern "C" int skg_margin_loss(const float* p_score, const float* n_score, int64_t rows, int64_t cols, float margin,
                               float adv_temperature, int use_adv, float* loss, float* d_p, float* d_n,
                               float grad_scale, cudaStream_t stream) {
  SKG_REQUIRE(rows > 0 && cols > 0 && p_score && n_score && loss, "margin loss needs non-empty scores");
  SKG_REQUIRE((d_p != nullptr) == (d_n != nullptr), "d_p and d_n must both be set or both null");
  SKG_CUDA(cudaMemsetAsync(loss, 0, sizeof(float), stream));
  const float inv_count = use_adv ? 1.f / (float)rows : 1.f / ((float)rows * (float)cols);
  const int threads = cols >= 256 ? 256 : (cols >= 64 ? 64 : 32);
  const int grid = (int)(rows < 4096 ? rows : 4096);
  margin_loss_kernel<<<grid, threads, 0, stream>>>(p_score, n_score, rows, cols, margin, adv_temperature, use_adv, loss,
                                                   d_p, d_n, grad_scale * inv_count);
  int rc = check_launch("margin_loss_kernel");
  if (rc) return rc;
  margin_loss_finalize<<<1, 1, 0, stream>>>(loss, inv_count, margin);
  return check_launch("margin_loss_finalize");
}
