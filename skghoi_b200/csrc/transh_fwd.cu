// TransH forward scoring (reference heads/TransH/TransH.py:56-106 and transH_head.forward head.py:558-582).
//
// Two kernels, both staged through shared memory and bound by their global stores:
//
//  transh_forward_kernel<EPL>  general triples (explicit head / relation / tail index arrays, or the implicit grid
//      numbering).  The three tables (80x50 + 2x117x50 fp32 = 62.8 KB for the HOI label graph) and the per-relation
//      inverse norms 1/||w||, 1/||r|| live in shared memory.  A warp scores FOUR consecutive triples at a time, one
//      per 8-lane group (dim <= 8*EPL elements per triple, EPL per lane), so a reduction is three shuffle steps and
//      four triples share every instruction.  The optional API-faithful outputs (the RAW gathered rows h_, r, w, t_
//      the fork returns, TransH.py:106) are a second phase: the 4 x dim floats the warp owes each output array are
//      contiguous in global memory, so the warp copies them table -> global as full 128-byte coalesced stores.
//
//  transh_grid_kernel  the in-head form (head.py:570-572): head = human_idx, relation = 0..rel_tot-1, tail = box
//      index, for every grid cell.  score(cell, r) depends only on (tail, r): the [ent_tot, rel_tot] score table is
//      computed ONCE by the general kernel (9 360 triples) and this kernel broadcasts it -- one warp per cell, pure
//      shared-memory -> global copies (4 B / triple score-only, 804 B / triple with the four row outputs).
#include "common.cuh"
#include "skghoi_b200.h"
#include "transh.cuh"

namespace skg {

constexpr int kFwdThreads = 256;
constexpr int kFwdWarps = kFwdThreads / 32;

__device__ __forceinline__ float group8_sum(float v) {
  v += __shfl_xor_sync(0xffffffffu, v, 1);
  v += __shfl_xor_sync(0xffffffffu, v, 2);
  v += __shfl_xor_sync(0xffffffffu, v, 4);
  return v;
}

// global -> shared staging of a table with 16-byte loads and eight of them in flight per thread (a scalar loop leaves one
// L2 round trip per element on the critical path of every CTA)
__device__ __forceinline__ void stage_table(float* __restrict__ dst, const float* __restrict__ src, int n) {
  if (((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 15u) == 0) {
    const int n4 = n >> 2;
    const float4* s4 = reinterpret_cast<const float4*>(src);
    float4* d4 = reinterpret_cast<float4*>(dst);
#pragma unroll 8
    for (int e = threadIdx.x; e < n4; e += blockDim.x) d4[e] = __ldg(s4 + e);
    for (int e = (n4 << 2) + threadIdx.x; e < n; e += blockDim.x) dst[e] = __ldg(src + e);
  } else {
#pragma unroll 8
    for (int e = threadIdx.x; e < n; e += blockDim.x) dst[e] = __ldg(src + e);
  }
}

// F.normalize(x, 2, -1): x / max(||x||, 1e-12)   (TransH.py:57-60, 76)
__device__ __forceinline__ float inv_norm(float sq) { return 1.f / fmaxf(sqrtf(sq), 1e-12f); }

template <int EPL, bool SMEM_TABLES>
__global__ void __launch_bounds__(kFwdThreads, (EPL <= 8 ? 3 : 1)) transh_forward_kernel(
    const float* __restrict__ ent, const float* __restrict__ rel, const float* __restrict__ nrm, int ent_tot, int rel_tot,
    int dim, TripleIndex idx, int64_t T, int p_norm, int norm_flag, float* __restrict__ h_out, float* __restrict__ r_out,
    float* __restrict__ w_out, float* __restrict__ t_out, float* __restrict__ score_out) {
  extern __shared__ __align__(16) float smem[];
  const float* E = ent; const float* R = rel; const float* W = nrm;
  const float* inv_w_tab = nullptr; const float* inv_r_tab = nullptr;
  if (SMEM_TABLES) {
    const int aE = (ent_tot * dim + 3) & ~3, aR = (rel_tot * dim + 3) & ~3;     // 16-byte aligned sub-tables
    float* sE = smem; float* sR = sE + aE; float* sW = sR + aR;
    float* s_iw = sW + aR; float* s_ir = s_iw + rel_tot;
    stage_table(sE, ent, ent_tot * dim); stage_table(sR, rel, rel_tot * dim); stage_table(sW, nrm, rel_tot * dim);
    __syncthreads();
    for (int r = threadIdx.x; r < rel_tot; r += blockDim.x) {   // per-relation inverse norms, once per CTA
      float ww = 0.f, rr = 0.f;
      for (int d = 0; d < dim; ++d) { const float a = sW[r * dim + d], b = sR[r * dim + d]; ww += a * a; rr += b * b; }
      s_iw[r] = inv_norm(ww); s_ir[r] = inv_norm(rr);
    }
    __syncthreads();
    E = sE; R = sR; W = sW; inv_w_tab = s_iw; inv_r_tab = s_ir;
  }
  const int lane = threadIdx.x & 31, grp = lane >> 3, l8 = lane & 7;
  const int64_t warp = (int64_t)blockIdx.x * kFwdWarps + (threadIdx.x >> 5);
  const int64_t nwarps = (int64_t)gridDim.x * kFwdWarps;
  for (int64_t t0 = warp * 4; t0 < T; t0 += nwarps * 4) {
    const int64_t t = t0 + grp;
    const bool live = t < T;
    int64_t hi = 0, ri = 0, ti = 0;
    if (live) idx.get(t, hi, ri, ti);
    float h[EPL], tl[EPL], r[EPL], w[EPL];
    float hw = 0.f, tw = 0.f, ww = 0.f, rr = 0.f;
#pragma unroll
    for (int k = 0; k < EPL; ++k) {
      const int d = l8 + 8 * k;
      const bool in = d < dim;
      h[k] = in ? E[hi * dim + d] : 0.f; tl[k] = in ? E[ti * dim + d] : 0.f;
      r[k] = in ? R[ri * dim + d] : 0.f; w[k] = in ? W[ri * dim + d] : 0.f;
      hw += h[k] * w[k]; tw += tl[k] * w[k];
      if (!SMEM_TABLES) { ww += w[k] * w[k]; rr += r[k] * r[k]; }
    }
    hw = group8_sum(hw); tw = group8_sum(tw);
    float inv_w, inv_r;
    if (SMEM_TABLES) { inv_w = inv_w_tab[ri]; inv_r = inv_r_tab[ri]; }
    else { inv_w = inv_norm(group8_sum(ww)); inv_r = inv_norm(group8_sum(rr)); }
    const float a_h = hw * inv_w, a_t = tw * inv_w;          // sum(e * w_hat)   TransH.py:85
    float hp[EPL], tp[EPL];
    float hh = 0.f, tt = 0.f;
#pragma unroll
    for (int k = 0; k < EPL; ++k) {
      const float wn = w[k] * inv_w;
      hp[k] = h[k] - a_h * wn; tp[k] = tl[k] - a_t * wn;
      hh += hp[k] * hp[k]; tt += tp[k] * tp[k];
    }
    float inv_h = 1.f, inv_t = 1.f;
    if (norm_flag) { inv_h = inv_norm(group8_sum(hh)); inv_t = inv_norm(group8_sum(tt)); } else inv_r = 1.f;
    float acc = 0.f;
#pragma unroll
    for (int k = 0; k < EPL; ++k) {
      const float s = (hp[k] * inv_h + r[k] * inv_r) - tp[k] * inv_t;   // (h + r) - t   TransH.py:68
      acc += (p_norm == 1) ? fabsf(s) : s * s;
    }
    acc = group8_sum(acc);
    if (live && l8 == 0) score_out[t] = (p_norm == 1) ? acc : sqrtf(acc);   // torch.norm(score, p, -1)  TransH.py:70
    if (h_out) {
      // phase 2: the whole warp copies the rows of triples t0 .. t0+3, one triple at a time (coalesced, dim floats each)
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        if (t0 + g >= T) break;                          // warp-uniform
        const int64_t hq = __shfl_sync(0xffffffffu, hi, g * 8), rq = __shfl_sync(0xffffffffu, ri, g * 8),
                      tq = __shfl_sync(0xffffffffu, ti, g * 8);
        const float* eh = E + hq * dim; const float* er = R + rq * dim; const float* ew = W + rq * dim; const float* et = E + tq * dim;
        const int64_t base = (t0 + g) * dim;
        for (int d = lane; d < dim; d += 32) {
          __stcs(h_out + base + d, eh[d]); __stcs(r_out + base + d, er[d]);
          __stcs(w_out + base + d, ew[d]); __stcs(t_out + base + d, et[d]);
        }
      }
    }
  }
}


// Broadcast of the [ent_tot, rel_tot] score table over the grid cells (+ the four row outputs): one warp per cell.
constexpr int kGridThreads = 512;
constexpr int kGridWarps = kGridThreads / 32;

template <bool ROWS>
__global__ void __launch_bounds__(kGridThreads) transh_grid_kernel(
    const float* __restrict__ S, const float* __restrict__ ent, const float* __restrict__ rel, const float* __restrict__ nrm,
    int ent_tot, int rel_tot, int dim, int head_const, const int64_t* __restrict__ cell_tail, int64_t num_cells,
    float* __restrict__ h_out, float* __restrict__ r_out, float* __restrict__ w_out, float* __restrict__ t_out,
    float* __restrict__ score) {
  extern __shared__ __align__(16) float smem[];
  float* sS = smem;             // score-only mode; rows mode reads the score table through L1 (__ldg) and keeps the
                                // 37 KB out of shared memory: 63 KB per CTA -> three resident CTAs per SM instead of two
  const int nS = ent_tot * rel_tot, nE = ent_tot * dim, nR = rel_tot * dim;
  // rows mode: sS | sE | sR | sW ; score-only mode: sS | s_tail (the table pointers then collapse onto the end of sS)
  float* sE = sS + (ROWS ? 0 : ((nS + 3) & ~3)); float* sR = sE + (ROWS ? ((nE + 3) & ~3) : 0); float* sW = sR + (ROWS ? ((nR + 3) & ~3) : 0);
  if (!ROWS) stage_table(sS, S, nS);
  if (ROWS) { stage_table(sE, ent, nE); stage_table(sR, rel, nR); stage_table(sW, nrm, nR); }
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int64_t warp = (int64_t)blockIdx.x * kGridWarps + (threadIdx.x >> 5);
  const int64_t nwarps = (int64_t)gridDim.x * kGridWarps;
  const bool vec2 = ROWS && (dim & 1) == 0 && (((uintptr_t)h_out | (uintptr_t)r_out | (uintptr_t)w_out | (uintptr_t)t_out) & 7u) == 0;
  const int step = (vec2 ? 64 : 32) % dim;
  if (!ROWS) {
    // score only: a CTA owns chunks of kGridThreads consecutive cells = one contiguous, 16-byte aligned slab of the output.
    // The chunk's tails are staged in shared memory (one coalesced load), then every thread writes whole float4s:
    // element q of the slab is relation q % rel_tot of cell q / rel_tot.
    int* s_tail = reinterpret_cast<int*>(sW);
    const float inv_rel = 1.f / (float)rel_tot;
    const bool aligned = (reinterpret_cast<uintptr_t>(score) & 15u) == 0;
    for (int64_t c0 = (int64_t)blockIdx.x * kGridThreads; c0 < num_cells; c0 += (int64_t)gridDim.x * kGridThreads) {
      const int64_t cl = c0 + threadIdx.x;
      int tv = 0;
      if (cl < num_cells) {
        const int64_t t64 = cell_tail ? __ldg(cell_tail + cl) : cl;
        tv = (int)(t64 >= ent_tot ? ent_tot - 1 : (t64 < 0 ? 0 : t64));     // host rejects out-of-range tails (IndexError)
      }
      __syncthreads();                                   // previous chunk's readers are done
      s_tail[threadIdx.x] = tv * rel_tot;
      __syncthreads();
      const int cells_here = (int)(num_cells - c0 < kGridThreads ? num_cells - c0 : kGridThreads);
      const int n_el = cells_here * rel_tot;
      float* dst = score + c0 * rel_tot;                 // c0 is a multiple of 4: 16-byte aligned when `score` is
      for (int q = 4 * threadIdx.x; q < n_el; q += 4 * kGridThreads) {
        int cell = (int)(((float)q + 0.5f) * inv_rel);
        if (cell * rel_tot > q) --cell;
        int r = q - cell * rel_tot;
        if (r >= rel_tot) { r -= rel_tot; ++cell; }
        float v[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          v[j] = (q + j < n_el) ? sS[s_tail[cell] + r] : 0.f;
          if (++r == rel_tot) { r = 0; ++cell; }
        }
        if (aligned && q + 4 <= n_el) {
          __stcs(reinterpret_cast<float4*>(dst + q), make_float4(v[0], v[1], v[2], v[3]));
        } else {
#pragma unroll
          for (int j = 0; j < 4; ++j) if (q + j < n_el) dst[q + j] = v[j];
        }
      }
    }
    return;
  }
  // rows mode, fast path: a CTA owns PAIRS of consecutive cells.  One cell's block of an output array is rel_tot*dim
  // floats (23 400 B for 117 x 50: only 8-byte aligned), a pair's is 16-byte aligned whenever rel_tot*dim is even, so
  // the pair is written as whole float4s -- 512 contiguous bytes per warp store -- each assembled from two float2
  // shared-memory reads (dim even: a float2 never straddles a row, nR even: nor a cell).
  int64_t c_done = 0;
  if ((nR & 1) == 0 && (dim & 1) == 0 &&
      (((uintptr_t)h_out | (uintptr_t)r_out | (uintptr_t)w_out | (uintptr_t)t_out) & 15u) == 0) {
    const int64_t num_pairs = num_cells >> 1;
    const int n4 = nR >> 1;                              // float4s per pair and per output array
    const float* eh = sE + head_const * dim;
    for (int64_t pr = blockIdx.x; pr < num_pairs; pr += gridDim.x) {
      const int64_t c0 = 2 * pr;
      int64_t ta = cell_tail ? __ldg(cell_tail + c0) : c0, tb = cell_tail ? __ldg(cell_tail + c0 + 1) : c0 + 1;
      const int tl0 = (int)(ta >= ent_tot ? ent_tot - 1 : (ta < 0 ? 0 : ta));
      const int tl1 = (int)(tb >= ent_tot ? ent_tot - 1 : (tb < 0 ? 0 : tb));
      for (int q = threadIdx.x; q < 2 * rel_tot; q += kGridThreads)
        __stcs(score + c0 * rel_tot + q, __ldg(S + (q < rel_tot ? tl0 * rel_tot : tl1 * rel_tot - rel_tot) + q));
      const float* et0 = sE + tl0 * dim; const float* et1 = sE + tl1 * dim;
      const int64_t base = c0 * nR;
      float4* ph = reinterpret_cast<float4*>(h_out + base); float4* pr4 = reinterpret_cast<float4*>(r_out + base);
      float4* pw = reinterpret_cast<float4*>(w_out + base); float4* pt = reinterpret_cast<float4*>(t_out + base);
      const int dstart = (4 * (int)threadIdx.x) % dim;   // (4k) % dim, advanced incrementally; nR % dim == 0, so the
      const int dstep = (4 * kGridThreads) % dim;        // wrap into the second cell leaves it unchanged
      // one output array at a time: at any instant the CTA's stores sweep ONE contiguous 2*nR-float region
      for (int k = threadIdx.x; k < n4; k += kGridThreads) {          // r_out, w_out: the tables themselves, twice
        int o0 = 4 * k, o1 = 4 * k + 2;
        if (o0 >= nR) o0 -= nR;
        if (o1 >= nR) o1 -= nR;
        const float2 r0 = *reinterpret_cast<const float2*>(sR + o0), r1 = *reinterpret_cast<const float2*>(sR + o1);
        __stcs(pr4 + k, make_float4(r0.x, r0.y, r1.x, r1.y));
      }
      for (int k = threadIdx.x; k < n4; k += kGridThreads) {
        int o0 = 4 * k, o1 = 4 * k + 2;
        if (o0 >= nR) o0 -= nR;
        if (o1 >= nR) o1 -= nR;
        const float2 w0 = *reinterpret_cast<const float2*>(sW + o0), w1 = *reinterpret_cast<const float2*>(sW + o1);
        __stcs(pw + k, make_float4(w0.x, w0.y, w1.x, w1.y));
      }
      int d0 = dstart;
      for (int k = threadIdx.x; k < n4; k += kGridThreads) {          // h_out: the constant head row, 2*rel_tot times
        int d1 = d0 + 2; if (d1 >= dim) d1 -= dim;
        const float2 a0 = *reinterpret_cast<const float2*>(eh + d0), a1 = *reinterpret_cast<const float2*>(eh + d1);
        __stcs(ph + k, make_float4(a0.x, a0.y, a1.x, a1.y));
        d0 += dstep; if (d0 >= dim) d0 -= dim;
      }
      d0 = dstart;
      for (int k = threadIdx.x; k < n4; k += kGridThreads) {          // t_out: tail row of cell 0, then of cell 1
        int d1 = d0 + 2; if (d1 >= dim) d1 -= dim;
        const float* e0 = (4 * k >= nR) ? et1 : et0; const float* e1 = (4 * k + 2 >= nR) ? et1 : et0;
        const float2 a0 = *reinterpret_cast<const float2*>(e0 + d0), a1 = *reinterpret_cast<const float2*>(e1 + d1);
        __stcs(pt + k, make_float4(a0.x, a0.y, a1.x, a1.y));
        d0 += dstep; if (d0 >= dim) d0 -= dim;
      }
    }
    c_done = num_pairs * 2;                              // an odd last cell falls through to the per-warp path
  }
  // rows mode, general path: one warp per cell -- 4 x rel_tot x dim floats of pure shared-memory -> global copy per cell
  for (int64_t c = c_done + warp; c < num_cells; c += nwarps) {
    int64_t t64 = cell_tail ? __ldg(cell_tail + c) : c;
    const int tl = (int)(t64 >= ent_tot ? ent_tot - 1 : (t64 < 0 ? 0 : t64));
    const float* srow = S + tl * rel_tot;
    float* dst = score + c * rel_tot;
    for (int r = lane; r < rel_tot; r += 32) __stcs(dst + r, __ldg(srow + r));
    const int64_t base = c * nR;
    const float* eh = sE + head_const * dim; const float* et = sE + tl * dim;
    if (vec2) {                                       // nR even, base even: 8-byte aligned float2 stores
      int d = (2 * lane) % dim;
      for (int q = 2 * lane; q < nR; q += 64) {
        const float2 vh = *reinterpret_cast<const float2*>(eh + d), vt = *reinterpret_cast<const float2*>(et + d);
        const float2 vr = *reinterpret_cast<const float2*>(sR + q), vw = *reinterpret_cast<const float2*>(sW + q);
        __stcs(reinterpret_cast<float2*>(h_out + base + q), vh); __stcs(reinterpret_cast<float2*>(r_out + base + q), vr);
        __stcs(reinterpret_cast<float2*>(w_out + base + q), vw); __stcs(reinterpret_cast<float2*>(t_out + base + q), vt);
        d += step; if (d >= dim) d -= dim;
      }
    } else {
      int d = lane % dim;
      for (int q = lane; q < nR; q += 32) {
        __stcs(h_out + base + q, eh[d]); __stcs(r_out + base + q, sR[q]);
        __stcs(w_out + base + q, sW[q]); __stcs(t_out + base + q, et[d]);
        d += step; if (d >= dim) d -= dim;
      }
    }
  }
}

// Rows mode through the bulk-copy engine.  Three of the four row outputs do not depend on the cell at all: r_out and
// w_out repeat the whole relation / normal tables per cell, h_out repeats the constant head row.  A PAIR of cells is a
// 16-byte aligned block of 2 * rel_tot * dim floats in every output array, so the CTA keeps one ready-made image of
// that block per array in shared memory (3 x 46 800 B for 117 x 50) and one elected thread writes each pair with three
// cp.async.bulk stores -- no load / store instructions at all for 3/4 of the bytes, the images are never rewritten, so
// nothing ever waits on the copy engine.  The threads only assemble t_out (the pair's two tail rows) and the scores.
__device__ __forceinline__ void bulk_store(float* gdst, const float* ssrc, uint32_t bytes) {
  const uint32_t saddr = (uint32_t)__cvta_generic_to_shared(ssrc);
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;\n" ::"l"(gdst), "r"(saddr), "r"(bytes) : "memory");
}

__global__ void __launch_bounds__(kGridThreads, 1) transh_grid_rows_bulk_kernel(
    const float* __restrict__ S, const float* __restrict__ ent, const float* __restrict__ rel, const float* __restrict__ nrm,
    int ent_tot, int rel_tot, int dim, int head_const, const int64_t* __restrict__ cell_tail, int64_t num_cells,
    float* __restrict__ h_out, float* __restrict__ r_out, float* __restrict__ w_out, float* __restrict__ t_out,
    float* __restrict__ score) {
  extern __shared__ __align__(16) float smem[];
  const int nE = ent_tot * dim, nR = rel_tot * dim;
  float* sRR = smem;                       // [2 * nR] relation table, twice
  float* sWW = sRR + 2 * nR;               // [2 * nR] normal-vector table, twice
  float* sHH = sWW + 2 * nR;               // [2 * nR] head row, 2 * rel_tot times
  float* sE = sHH + 2 * nR;                // [nE] entity table (tail rows)          (2 * nR floats = a multiple of 16 B)
  for (int e = threadIdx.x; e < nR; e += kGridThreads) {
    const float r = __ldg(rel + e), w = __ldg(nrm + e);
    sRR[e] = r; sRR[e + nR] = r; sWW[e] = w; sWW[e + nR] = w;
  }
  for (int e = threadIdx.x; e < nE; e += kGridThreads) sE[e] = __ldg(ent + e);
  {
    int d = threadIdx.x % dim;
    const int step = kGridThreads % dim;
    for (int e = threadIdx.x; e < 2 * nR; e += kGridThreads) {
      sHH[e] = __ldg(ent + head_const * dim + d);
      d += step; if (d >= dim) d -= dim;
    }
  }
  asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");   // generic-proxy writes -> visible to the bulk engine
  __syncthreads();
  const int64_t num_pairs = num_cells >> 1;
  const int n4 = nR >> 1;                              // float4s per pair and per output array
  const uint32_t pair_bytes = (uint32_t)(2 * nR) * 4u;
  const int dstart = (4 * (int)threadIdx.x) % dim;
  const int dstep = (4 * kGridThreads) % dim;
  for (int64_t pr = blockIdx.x; pr < num_pairs; pr += gridDim.x) {
    const int64_t c0 = 2 * pr;
    const int64_t base = c0 * nR;
    if (threadIdx.x == 0) {
      bulk_store(r_out + base, sRR, pair_bytes);
      bulk_store(w_out + base, sWW, pair_bytes);
      bulk_store(h_out + base, sHH, pair_bytes);
      asm volatile("cp.async.bulk.commit_group;\n" ::: "memory");
    }
    int64_t ta = cell_tail ? __ldg(cell_tail + c0) : c0, tb = cell_tail ? __ldg(cell_tail + c0 + 1) : c0 + 1;
    const int tl0 = (int)(ta >= ent_tot ? ent_tot - 1 : (ta < 0 ? 0 : ta));
    const int tl1 = (int)(tb >= ent_tot ? ent_tot - 1 : (tb < 0 ? 0 : tb));
    for (int q = threadIdx.x; q < 2 * rel_tot; q += kGridThreads)
      __stcs(score + c0 * rel_tot + q, __ldg(S + (q < rel_tot ? tl0 * rel_tot : tl1 * rel_tot - rel_tot) + q));
    const float* et0 = sE + tl0 * dim; const float* et1 = sE + tl1 * dim;
    float4* pt = reinterpret_cast<float4*>(t_out + base);
    int d0 = dstart;
    for (int k = threadIdx.x; k < n4; k += kGridThreads) {            // t_out: tail row of cell 0, then of cell 1
      int d1 = d0 + 2; if (d1 >= dim) d1 -= dim;
      const float* e0 = (4 * k >= nR) ? et1 : et0; const float* e1 = (4 * k + 2 >= nR) ? et1 : et0;
      const float2 a0 = *reinterpret_cast<const float2*>(e0 + d0), a1 = *reinterpret_cast<const float2*>(e1 + d1);
      __stcs(pt + k, make_float4(a0.x, a0.y, a1.x, a1.y));
      d0 += dstep; if (d0 >= dim) d0 -= dim;
    }
  }
  if ((num_cells & 1) && blockIdx.x == gridDim.x - 1) {                // odd last cell: plain stores
    const int64_t c = num_cells - 1;
    const int64_t t64 = cell_tail ? __ldg(cell_tail + c) : c;
    const int tl = (int)(t64 >= ent_tot ? ent_tot - 1 : (t64 < 0 ? 0 : t64));
    for (int q = threadIdx.x; q < rel_tot; q += kGridThreads) score[c * rel_tot + q] = __ldg(S + tl * rel_tot + q);
    const int64_t base = c * nR;
    for (int q = threadIdx.x; q < nR; q += kGridThreads) {
      h_out[base + q] = sHH[q]; r_out[base + q] = sRR[q]; w_out[base + q] = sWW[q]; t_out[base + q] = sE[tl * dim + q % dim];
    }
  }
  if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group.read 0;\n" ::: "memory");   // images stay valid until read
}

template <int EPL>
static int launch_forward(const float* ent, const float* rel, const float* nrm, int ent_tot, int rel_tot, int dim,
                          const TripleIndex& idx, int64_t T, int p_norm, int norm_flag, float* h_out, float* r_out,
                          float* w_out, float* t_out, float* score, cudaStream_t stream) {
  const size_t table_bytes = ((((size_t)ent_tot * dim + 3) & ~(size_t)3) + 2 * (((size_t)rel_tot * dim + 3) & ~(size_t)3) + 2 * (size_t)rel_tot) * sizeof(float);
  // few triples and no row outputs (the [ent_tot, rel_tot] score table of the grid form): staging 60 KB of tables per CTA
  // would be the whole run time -- read the rows through L1/L2 instead and spread one 4-triple step per warp
  const bool small = T <= 32768 && h_out == nullptr;
  const bool fits = table_bytes <= 100 * 1024 && !small;
  const int64_t want = ceil_div64(T, kFwdWarps * 4 * (small ? 1 : 4));
  const int per_sm = fits ? (table_bytes > 72 * 1024 ? 2 : 3) : 8;
  int grid = (int)(want < (int64_t)num_sms() * per_sm ? want : (int64_t)num_sms() * per_sm);
  if (grid < 1) grid = 1;
  if (fits) {
    auto k = transh_forward_kernel<EPL, true>;
    SKG_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)table_bytes));
    k<<<grid, kFwdThreads, table_bytes, stream>>>(ent, rel, nrm, ent_tot, rel_tot, dim, idx, T, p_norm, norm_flag, h_out,
                                                  r_out, w_out, t_out, score);
  } else {
    transh_forward_kernel<EPL, false><<<grid, kFwdThreads, 0, stream>>>(ent, rel, nrm, ent_tot, rel_tot, dim, idx, T, p_norm,
                                                                        norm_flag, h_out, r_out, w_out, t_out, score);
  }
  return check_launch("transh_forward_kernel");
}

static int dispatch_forward(const float* ent, const float* rel, const float* nrm, int ent_tot, int rel_tot, int dim,
                            const TripleIndex& idx, int64_t T, int p_norm, int norm_flag, float* h_out, float* r_out,
                            float* w_out, float* t_out, float* score, cudaStream_t stream) {
  SKG_REQUIRE(dim >= 1 && dim <= 256, "TransH dim must be in [1, 256], got %d", dim);
  SKG_REQUIRE(p_norm == 1 || p_norm == 2, "p_norm must be 1 or 2, got %d", p_norm);
  SKG_REQUIRE(ent_tot > 0 && rel_tot > 0 && T >= 0, "bad table sizes");
  if (T == 0) return SKG_OK;
  SKG_REQUIRE(ent && rel && nrm && score, "null table / score pointer");
  SKG_REQUIRE((h_out != nullptr) == (r_out != nullptr) && (h_out != nullptr) == (w_out != nullptr) &&
                  (h_out != nullptr) == (t_out != nullptr),
              "h_out, r_out, w_out, t_out must be all set or all null");
#define SKG_FWD_CASE(E) \
  return launch_forward<E>(ent, rel, nrm, ent_tot, rel_tot, dim, idx, T, p_norm, norm_flag, h_out, r_out, w_out, t_out, score, stream)
  if (dim <= 32) SKG_FWD_CASE(4);
  if (dim <= 64) SKG_FWD_CASE(8);
  if (dim <= 128) SKG_FWD_CASE(16);
  SKG_FWD_CASE(32);
#undef SKG_FWD_CASE
}

}  // namespace skg

using namespace skg;

extern "C" int skg_transh_forward(const float* ent, const float* rel, const float* nrm, int ent_tot, int rel_tot,
                                  int dim, const int64_t* head, const int64_t* relation, const int64_t* tail,
                                  int64_t num_triples, int p_norm, int norm_flag, float* h_out, float* r_out,
                                  float* w_out, float* t_out, float* score, cudaStream_t stream) {
  SKG_REQUIRE(num_triples == 0 || (head && relation && tail && score), "null index/score pointer");
  TripleIndex idx{head, relation, tail, nullptr, 0, rel_tot, 0};
  return dispatch_forward(ent, rel, nrm, ent_tot, rel_tot, dim, idx, num_triples, p_norm, norm_flag, h_out, r_out, w_out,
                          t_out, score, stream);
}

extern "C" int skg_transh_forward_grid(const float* ent, const float* rel, const float* nrm, int ent_tot, int rel_tot,
                                       int dim, int64_t head_const, const int64_t* cell_tail, int64_t num_cells,
                                       int p_norm, int norm_flag, float* h_out, float* r_out, float* w_out,
                                       float* t_out, float* score, cudaStream_t stream) {
  SKG_REQUIRE(num_cells == 0 || score, "null score pointer");
  SKG_REQUIRE(head_const >= 0 && head_const < ent_tot, "head_const out of range");
  SKG_REQUIRE((h_out != nullptr) == (r_out != nullptr) && (h_out != nullptr) == (w_out != nullptr) &&
                  (h_out != nullptr) == (t_out != nullptr),
              "h_out, r_out, w_out, t_out must be all set or all null");
  if (num_cells == 0) return SKG_OK;
  const bool rows = h_out != nullptr;
  const size_t nS = (size_t)ent_tot * rel_tot, nE = (size_t)ent_tot * dim, nR = (size_t)rel_tot * dim;
  const size_t smem_bytes = (rows ? ((nE + 3) & ~(size_t)3) + 2 * ((nR + 3) & ~(size_t)3) : ((nS + 3) & ~(size_t)3) + (size_t)kGridThreads) * sizeof(float);
  const bool implicit_tail_ok = cell_tail != nullptr || num_cells <= ent_tot;
  if (smem_bytes > 200 * 1024 || num_cells < 4 * (int64_t)ent_tot || !implicit_tail_ok) {
    // big tables or few cells: score every triple directly
    TripleIndex idx{nullptr, nullptr, nullptr, cell_tail, head_const, rel_tot, 1};
    return dispatch_forward(ent, rel, nrm, ent_tot, rel_tot, dim, idx, num_cells * rel_tot, p_norm, norm_flag, h_out, r_out,
                            w_out, t_out, score, stream);
  }
  // 1) score table S[tail, r] = score(head_const, r, tail) for every entity: ent_tot * rel_tot triples, once
  float* S = nullptr;
  {   // keep freed pool memory cached: without a release threshold every cudaFreeAsync hands the block back to the OS
    static thread_local int pool_dev = -1;
    int dev = 0;
    SKG_CUDA(cudaGetDevice(&dev));
    if (pool_dev != dev) {
      cudaMemPool_t pool;
      SKG_CUDA(cudaDeviceGetDefaultMemPool(&pool, dev));
      unsigned long long thr = 64ull << 20;
      unsigned long long cur = 0;
      cudaMemPoolGetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &cur);
      if (cur < thr) SKG_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr));
      pool_dev = dev;
    }
  }
  SKG_CUDA(cudaMallocAsync(&S, nS * sizeof(float), stream));     // stream-ordered pool allocation: no synchronisation
  TripleIndex idx{nullptr, nullptr, nullptr, nullptr, head_const, rel_tot, 1};
  int rc = dispatch_forward(ent, rel, nrm, ent_tot, rel_tot, dim, idx, (int64_t)nS, p_norm, norm_flag, nullptr, nullptr, nullptr,
                            nullptr, S, stream);
  if (rc == SKG_OK) {
    // 2) broadcast over the cells
    const int per_sm = smem_bytes > 110 * 1024 ? 1 : (smem_bytes > 72 * 1024 ? 2 : 3);
    const int64_t want = rows ? ceil_div64(num_cells, 2) : ceil_div64(num_cells, kGridThreads);   // rows: a CTA per cell pair
    const int grid = (int)(want < (int64_t)num_sms() * per_sm ? want : (int64_t)num_sms() * per_sm);
    const size_t bulk_bytes = ((size_t)6 * rel_tot * dim + (size_t)ent_tot * dim) * sizeof(float);
    const bool bulk = rows && num_cells >= 2 && ((rel_tot * dim) & 1) == 0 && (dim & 1) == 0 && bulk_bytes <= 220 * 1024 &&
                      (((uintptr_t)h_out | (uintptr_t)r_out | (uintptr_t)w_out | (uintptr_t)t_out) & 15u) == 0;
    if (bulk) {
      auto k = transh_grid_rows_bulk_kernel;
      cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bulk_bytes);
      const int64_t pairs = num_cells >> 1;
      const int g = (int)(pairs < (int64_t)num_sms() ? pairs : (int64_t)num_sms());
      k<<<g, kGridThreads, bulk_bytes, stream>>>(S, ent, rel, nrm, ent_tot, rel_tot, dim, (int)head_const, cell_tail, num_cells,
                                                h_out, r_out, w_out, t_out, score);
    } else if (rows) {
      auto k = transh_grid_kernel<true>;
      cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes);
      k<<<grid, kGridThreads, smem_bytes, stream>>>(S, ent, rel, nrm, ent_tot, rel_tot, dim, (int)head_const, cell_tail, num_cells,
                                                   h_out, r_out, w_out, t_out, score);
    } else {
      auto k = transh_grid_kernel<false>;
      cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes);
      k<<<grid, kGridThreads, smem_bytes, stream>>>(S, ent, rel, nrm, ent_tot, rel_tot, dim, (int)head_const, cell_tail, num_cells,
                                                   h_out, r_out, w_out, t_out, score);
    }
    rc = check_launch("transh_grid_kernel");
  }
  cudaFreeAsync(S, stream);
  return rc;
}
