// TransH knowledge-graph training as ONE persistent cooperative kernel (BASELINE configs[3]):
//   reference loop  OpenKE/openke/config/Trainer.py:43-99  (sample -> forward -> MarginLoss -> backward -> SGD, ~25 launches
//                   of the eager path for ~0.1 ms of GPU work)
//   sampler         OpenKE/openke/base/Base.cpp:78-197, Corrupt.h:7-83, Random.h:18-28  (the code of csrc/kgsample.cu: the same
//                   per-worker LCG streams, the same batches bit for bit)
//   scoring         heads/TransH/TransH.py:56-106 (csrc/transh.cuh: the per-triple forward / backward of csrc/transh.cu)
//   loss            heads/MarginLoss.py:28-36 without the adversarial weighting: mean(max(p - n, -margin)) + margin
//   strategy        OpenKE/openke/module/strategy/NegativeSampling.py:13-32 with regul_rate = 0
//   optimiser       plain SGD (Trainer.py: opt_method "sgd"), lr = alpha, no weight decay
//
// `num_steps` optimiser steps per launch; the grid (one CTA per SM, co-resident: cooperative launch) walks every step in
// four phases separated by grid-wide barriers:
//   A  sample the batch (one thread per TRIPLE: every negative jumps to its own position of the LCG stream), reset the per-step
//      accumulators   [+ SGD of the previous step]
//   B  stage the three tables in shared memory, score every triple (one warp per triple)
//   C  margin loss and d loss / d score (one thread per negative)
//   D  backward of every triple with a non-zero gradient into a per-CTA shared-memory copy of the gradient tables,
//      flushed once per CTA with atomics (the human entity row is hit by EVERY triple of the HOI label graph: atomics
//      straight into global memory would serialise on its 50 addresses)
// The SGD update of step s runs at the start of phase A of step s + 1 (and once more after the last step): it touches only
// the tables and the gradient accumulators, which phase A does not read.
#include "common.cuh"
#include "skghoi_b200.h"
#include "transh.cuh"
#include <cooperative_groups.h>

namespace cg = cooperative_groups;

namespace skg {

// ---- the sampler's device code (csrc/kgsample.cu), restated for in-kernel use -----------------------------------------
constexpr unsigned long long kKgLcgA = 25214903917ull, kKgLcgC = 11ull;
struct KgLcg {
  unsigned long long s;
  __device__ __forceinline__ unsigned long long next() { s = s * kKgLcgA + kKgLcgC; return s; }
  __device__ __forceinline__ int64_t below(int64_t x) { return (int64_t)(next() % (unsigned long long)x); }
};
__device__ inline unsigned long long kg_lcg_jump(unsigned long long state, unsigned long long k) {
  unsigned long long acc_m = 1, acc_p = 0, cur_m = kKgLcgA, cur_p = kKgLcgC;
  while (k) {
    if (k & 1) { acc_m *= cur_m; acc_p = acc_p * cur_m + cur_p; }
    cur_p = (cur_m + 1) * cur_p;
    cur_m *= cur_m;
    k >>= 1;
  }
  return acc_m * state + acc_p;
}
__device__ __forceinline__ int64_t kg_corrupt(const int64_t* __restrict__ rel, const int64_t* __restrict__ other,
                                              const int64_t* __restrict__ lef_arr, const int64_t* __restrict__ rig_arr, int64_t e,
                                              int64_t r, int64_t ent_total, int64_t keep, KgLcg& g) {
  int64_t lef = lef_arr[e] - 1, rig = rig_arr[e], mid;
  while (lef + 1 < rig) { mid = (lef + rig) >> 1; if (rel[mid] >= r) rig = mid; else lef = mid; }
  const int64_t ll = rig;
  lef = lef_arr[e]; rig = rig_arr[e] + 1;
  while (lef + 1 < rig) { mid = (lef + rig) >> 1; if (rel[mid] <= r) lef = mid; else rig = mid; }
  const int64_t rr = lef;
  const int64_t free_cnt = ent_total - (rr - ll + 1);
  if (free_cnt <= 0) { g.next(); return keep; }
  const int64_t tmp = g.below(free_cnt);
  if (tmp < other[ll]) return tmp;
  if (tmp > other[rr] - rr + ll - 1) return tmp + rr - ll + 1;
  lef = ll; rig = rr + 1;
  while (lef + 1 < rig) { mid = (lef + rig) >> 1; if (other[mid] - mid + ll - 1 < tmp) lef = mid; else rig = mid; }
  return tmp + lef - ll + 1;
}

struct KgTrainParams {
  skg_kg_train_args a;
  int64_t* batch_h; int64_t* batch_t; int64_t* batch_r;   // [B (1 + neg)]
  float* score; float* dscore;                            // [B (1 + neg)]
  float* grads;                                           // [nE + 2 nR]
};

constexpr int kKgThreads = 256;

template <int EPL>
__global__ void __launch_bounds__(kKgThreads, 1) kg_train_kernel(KgTrainParams p) {
  cg::grid_group grid = cg::this_grid();
  extern __shared__ __align__(16) float smem[];
  const skg_kg_train_args& a = p.a;
  const int dim = a.dim;
  const int nE = (int)a.ent_total * dim, nR = (int)a.rel_total * dim, nTab = nE + 2 * nR;
  float* sTab = smem;                 // ent | rel | nrm
  float* sGrad = smem + nTab;
  __shared__ float red[kKgThreads / 32];
  const int tid = threadIdx.x, lane = tid & 31;
  const int64_t gtid = (int64_t)blockIdx.x * kKgThreads + tid, gthreads = (int64_t)gridDim.x * kKgThreads;
  const int64_t gwarp = gtid >> 5, gwarps = gthreads >> 5;
  const int64_t B = a.batch_size, T3 = B * (1 + a.neg_rate), NN = B * a.neg_rate, T = a.train_total;
  const int W = a.work_threads;
  const float inv = NN > 0 ? 1.f / (float)NN : 0.f;
  const unsigned long long draws = 1ull + (unsigned long long)a.neg_rate * (a.mode == 0 ? 2ull : 1ull);
  const int64_t per = (B % W == 0) ? B / W : B / W + 1;
  unsigned long long* seeds = reinterpret_cast<unsigned long long*>(a.seeds);

  for (int step = 0; step <= a.num_steps; ++step) {
    // ---- SGD of the previous step (plain SGD: p -= lr * g), gradient accumulators back to zero
    if (step > 0) {
      for (int64_t e = gtid; e < nTab; e += gthreads) {
        const float g = p.grads[e];
        if (g != 0.f) {
          float* tab = e < nE ? a.ent + e : (e < nE + nR ? a.rel + (e - nE) : a.nrm + (e - nE - nR));
          *tab -= a.lr * g;
          p.grads[e] = 0.f;
        }
      }
    }
    if (step == a.num_steps) break;
    // ---- phase A: sample (csrc/kgsample.cu), reset the per-step accumulators
    const unsigned long long* seeds_in = seeds + (size_t)(step & 1) * W;
    unsigned long long* seeds_out = seeds + (size_t)((step + 1) & 1) * W;
    // one thread per TRIPLE of the batch (not per positive): negative k of item b starts at a fixed position of its worker's
    // stream -- 1 draw picks the positive, every negative takes `dpn` draws -- so it jumps there directly and re-derives the
    // positive with one extra draw; the 25 corruptions of an item (three dependent binary searches each) then run in
    // parallel instead of back to back in one thread (the per-positive form was ~90 us of a 140 us step)
    const unsigned long long dpn = (a.mode == 0) ? 2ull : 1ull;
    for (int64_t j = gtid; j < T3 || j < W; j += gthreads) {
      if (j < W) {
        int64_t lef = j * per, rig = (j + 1) * per;
        if (rig > B) rig = B;
        const int64_t items = rig > lef ? rig - lef : 0;
        seeds_out[j] = kg_lcg_jump(seeds_in[j], (unsigned long long)items * draws);
      }
      if (j >= T3) continue;
      const int64_t k = j / B, b = j - k * B;          // k = 0: the positive, k >= 1: negative k - 1
      const int64_t id = b / per, lef0 = id * per;
      const unsigned long long item0 = kg_lcg_jump(seeds_in[id], (unsigned long long)(b - lef0) * draws);
      KgLcg g{item0};
      const int64_t i = g.below(T);
      const int64_t h = a.head_hrt[i], r = a.head_hrt[T + i], t = a.head_hrt[2 * T + i];
      if (k == 0) {
        p.batch_h[b] = h; p.batch_t[b] = t; p.batch_r[b] = r;
        p.dscore[b] = 0.f;
        continue;
      }
      g.s = kg_lcg_jump(g.s, (unsigned long long)(k - 1) * dpn);
      int64_t nh = h, nt = t;
      bool corrupt_tail_entity;
      if (a.mode == 0) {
        float prob = 500.f;
        if (a.left_mean) prob = __fdiv_rn(__fmul_rn(1000.f, a.right_mean[r]), __fadd_rn(a.right_mean[r], a.left_mean[r]));
        corrupt_tail_entity = (float)(g.next() % 1000ull) < prob;
      } else {
        corrupt_tail_entity = (a.mode != -1);
      }
      if (corrupt_tail_entity)
        nt = kg_corrupt(a.head_hrt + T, a.head_hrt + 2 * T, a.lef_head, a.rig_head, h, r, a.ent_total, t, g);
      else
        nh = kg_corrupt(a.tail_hrt + T, a.tail_hrt, a.lef_tail, a.rig_tail, t, r, a.ent_total, h, g);
      p.batch_h[j] = nh; p.batch_t[j] = nt; p.batch_r[j] = r;
    }
    if (gtid == 0) a.loss_out[step] = a.margin;      // + mean(...) accumulated in phase C
    grid.sync();
    // ---- phase B: tables -> shared memory (plain loads: they changed inside this kernel), forward scores
    for (int e = tid; e < nTab; e += kKgThreads) {
      sTab[e] = e < nE ? a.ent[e] : (e < nE + nR ? a.rel[e - nE] : a.nrm[e - nE - nR]);
      sGrad[e] = 0.f;
    }
    __syncthreads();
    const float* sE = sTab; const float* sR = sTab + nE; const float* sW = sTab + nE + nR;
    for (int64_t t = gwarp; t < T3; t += gwarps) {
      const int64_t hi = p.batch_h[t], ri = p.batch_r[t], ti = p.batch_t[t];
      float h[EPL], tl[EPL], r[EPL], w[EPL], s[EPL];
#pragma unroll
      for (int e = 0; e < EPL; ++e) {
        const int d = lane + 32 * e;
        const bool in = d < dim;
        h[e] = in ? sE[hi * dim + d] : 0.f; tl[e] = in ? sE[ti * dim + d] : 0.f;
        r[e] = in ? sR[ri * dim + d] : 0.f; w[e] = in ? sW[ri * dim + d] : 0.f;
      }
      TranshTriple st;
      const float sc = transh_triple_fwd<EPL>(h, tl, r, w, a.p_norm, a.norm_flag, s, st);
      if (lane == 0) p.score[t] = sc;
    }
    grid.sync();
    // ---- phase C: MarginLoss (MarginLoss.py:31-35) of p [B] against n [B, neg] (negative k of positive i sits at B + k B + i)
    float part = 0.f;
    for (int64_t j = gtid; j < NN; j += gthreads) {
      const int64_t i = j % B;
      const float d = p.score[i] - p.score[B + j];
      const bool on = d > -a.margin;
      part += on ? d : -a.margin;
      const float g = on ? inv : 0.f;
      p.dscore[B + j] = -g;
      if (on) atomicAdd(p.dscore + i, g);
    }
    part = warp_sum(part);
    if (lane == 0) red[tid >> 5] = part;
    __syncthreads();
    if (tid == 0) {
      float sum = 0.f;
      for (int k = 0; k < kKgThreads / 32; ++k) sum += red[k];
      if (sum != 0.f) atomicAdd(a.loss_out + step, sum * inv);
    }
    grid.sync();
    // ---- phase D: backward into the CTA's shared-memory gradient tables, one flush per CTA
    float* gE = sGrad; float* gR = sGrad + nE; float* gW = sGrad + nE + nR;
    for (int64_t t = gwarp; t < T3; t += gwarps) {
      const float g = p.dscore[t];
      if (g == 0.f) continue;                        // warp-uniform
      const int64_t hi = p.batch_h[t], ri = p.batch_r[t], ti = p.batch_t[t];
      float h[EPL], tl[EPL], r[EPL], w[EPL], s[EPL];
#pragma unroll
      for (int e = 0; e < EPL; ++e) {
        const int d = lane + 32 * e;
        const bool in = d < dim;
        h[e] = in ? sE[hi * dim + d] : 0.f; tl[e] = in ? sE[ti * dim + d] : 0.f;
        r[e] = in ? sR[ri * dim + d] : 0.f; w[e] = in ? sW[ri * dim + d] : 0.f;
      }
      TranshTriple st;
      transh_triple_fwd<EPL>(h, tl, r, w, a.p_norm, a.norm_flag, s, st);
      float g_h[EPL], g_t[EPL], g_r[EPL], g_w[EPL];
      transh_triple_bwd<EPL>(h, tl, r, w, s, st, g, a.p_norm, a.norm_flag, g_h, g_t, g_r, g_w);
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
    __syncthreads();
    for (int e = tid; e < nTab; e += kKgThreads) {
      const float v = sGrad[e];
      if (v != 0.f) atomicAdd(p.grads + e, v);
    }
    grid.sync();
  }
}

static inline int64_t kg_align(int64_t v) { return (v + 255) / 256 * 256; }

}  // namespace skg

using namespace skg;

extern "C" int64_t skg_kg_train_scratch_bytes(int64_t batch_size, int neg_rate, int64_t ent_total, int64_t rel_total, int dim) {
  const int64_t T3 = batch_size * (1 + neg_rate);
  return 3 * kg_align(T3 * 8) + 2 * kg_align(T3 * 4) + kg_align((ent_total + 2 * rel_total) * dim * 4);
}

extern "C" int skg_kg_train_steps(const skg_kg_train_args* args, cudaStream_t stream) {
  SKG_REQUIRE(args, "null pointer argument");
  const skg_kg_train_args& a = *args;
  SKG_REQUIRE(a.head_hrt && a.tail_hrt && a.lef_head && a.rig_head && a.lef_tail && a.rig_tail && a.seeds, "null index pointer");
  SKG_REQUIRE((a.left_mean == nullptr) == (a.right_mean == nullptr), "left_mean and right_mean come together (bern) or not at all");
  SKG_REQUIRE(a.ent && a.rel && a.nrm && a.loss_out && a.scratch, "null table / output pointer");
  SKG_REQUIRE(a.train_total > 0 && a.ent_total > 1 && a.rel_total > 0 && a.work_threads >= 1 && a.batch_size > 0 && a.neg_rate >= 1,
              "bad sizes");
  SKG_REQUIRE(a.mode == 0 || a.mode == -1 || a.mode == 1, "mode must be 0, -1 or 1");
  SKG_REQUIRE(a.dim >= 1 && a.dim <= 256, "TransH dim must be in [1, 256], got %d", a.dim);
  SKG_REQUIRE(a.p_norm == 1 || a.p_norm == 2, "p_norm must be 1 or 2");
  SKG_REQUIRE(a.num_steps >= 0, "negative step count");
  if (a.num_steps == 0) return SKG_OK;
  const int64_t need = skg_kg_train_scratch_bytes(a.batch_size, a.neg_rate, a.ent_total, a.rel_total, a.dim);
  SKG_REQUIRE(a.scratch_bytes >= need, "scratch too small: %lld < %lld bytes", (long long)a.scratch_bytes, (long long)need);
  SKG_REQUIRE((reinterpret_cast<uintptr_t>(a.scratch) & 255u) == 0, "scratch must be 256-byte aligned");
  const int64_t T3 = a.batch_size * (1 + a.neg_rate);
  const int64_t nTab = (a.ent_total + 2 * a.rel_total) * a.dim;
  const size_t smem = (size_t)nTab * 2 * sizeof(float);
  SKG_REQUIRE(smem <= 200 * 1024, "the fused trainer keeps the tables and their gradients in shared memory (%zu bytes needed, 200 KB "
              "available): use the eager Trainer for larger graphs", smem);
  KgTrainParams p;
  p.a = a;
  uint8_t* s = reinterpret_cast<uint8_t*>(a.scratch);
  p.batch_h = reinterpret_cast<int64_t*>(s); s += kg_align(T3 * 8);
  p.batch_t = reinterpret_cast<int64_t*>(s); s += kg_align(T3 * 8);
  p.batch_r = reinterpret_cast<int64_t*>(s); s += kg_align(T3 * 8);
  p.score = reinterpret_cast<float*>(s); s += kg_align(T3 * 4);
  p.dscore = reinterpret_cast<float*>(s); s += kg_align(T3 * 4);
  p.grads = reinterpret_cast<float*>(s);
  SKG_CUDA(cudaMemsetAsync(p.grads, 0, (size_t)nTab * 4, stream));
  const int epl = (a.dim + 31) / 32;
  void* kern = epl <= 2 ? (void*)kg_train_kernel<2> : epl <= 4 ? (void*)kg_train_kernel<4> : (void*)kg_train_kernel<8>;
  SKG_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int per_sm = 0;
  SKG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kKgThreads, smem));
  SKG_REQUIRE(per_sm >= 1, "the fused KG trainer does not fit on an SM");
  // enough warps for one triple each, at most one CTA per SM (all co-resident: grid-wide barriers)
  int64_t want = (T3 * 32 + kKgThreads - 1) / kKgThreads;
  int grid = (int)(want < num_sms() ? want : num_sms());
  if (grid < 1) grid = 1;
  void* kargs[] = {&p};
  SKG_CUDA(cudaLaunchCooperativeKernel(kern, dim3(grid), dim3(kKgThreads), kargs, smem, stream));
  return check_launch("kg_train_kernel");
}
