// Pair enumeration + pairwise spatial encoding (reference head.py:847-868, ops.py:20-157).
//
// One launch covers every image of the batch: the human x box grid cells of all images are
// numbered consecutively (cell r of image b = cell_off[b] + i*n + j, row-major like
// torch.meshgrid(indexing='ij')), one thread per cell.  The kernel is HBM-write bound
// (46 fp32 = 184 B per cell): each CTA stages a 128-cell tile in shared memory laid out exactly
// like the output rows and drains it with coalesced 16-byte streaming stores issued by all 128
// threads (st.global.cs.v4: consecutive lanes on consecutive addresses, full 128-byte lines
// regardless of the 184-byte row pitch).  kDrainBulk = true selects the earlier drain, one
// cp.async.bulk per tile (1-D TMA, SASS UBLKCP): it left the CTA waiting on the bulk engine's
// shared-memory read (0.64 vs 0.76 of copy peak) and is kept for comparison only.
//
// Valid-pair compaction (torch.nonzero(x != y), head.py:852): in row i exactly the cell j == i is
// dropped (i < n_h <= n), so the row-major rank of a kept cell is p = i*(n-1) + j - (j > i):
// bit-exact with nonzero() ordering, no scan and no host sync.
//
// Arithmetic: the reference computes 22 fp32 divisions and 23 logs per pair; with IEEE division and libm logf
// that is ~1000 instructions per cell and the kernel is issue-bound at a third of the HBM roofline.  Parity for
// the encodings is 1e-5 relative, so divisions use MUFU.RCP (__fdividef, 2 ulp) and logs
// MUFU.LG2 (__logf, |err| < 2^-21 near 1): ~45 MUFU per cell, well under the store time.  NaN / inf placement is
// unchanged (0/0 -> NaN, x/0 -> inf, log(<0) -> NaN, log(0 + eps) finite).  Compiled with -fmad=false so the
// products keep the reference's separate rounding.
#include "common.cuh"
#include "skghoi_b200.h"

namespace skg {

constexpr int kTileCells = 128;
constexpr bool kDrainBulk = false;   // tile drain: false = coalesced st.global.cs.v4 by all threads, true = one cp.async.bulk

struct SegTable {
  const int32_t* box_off;    // [B+1] prefix sum of n
  const int32_t* n_h;        // [B]
  const int64_t* cell_off;   // [B+1] prefix sum of n_h*n
  const int64_t* pair_off;   // [B+1] prefix sum of n_h*(n-1)
  const int32_t* hum_off;    // [B+1] prefix sum of n_h
  const float* hw;           // [B,3] = h, w, h*w (as fp32)
  int B;
};

__device__ __forceinline__ int find_segment(const int64_t* __restrict__ off, int B, int64_t r) {
  int lo = 0, hi = B;  // invariant: off[lo] <= r < off[hi]
  while (hi - lo > 1) {
    int mid = (lo + hi) >> 1;
    if (__ldg(off + mid) <= r) lo = mid; else hi = mid;
  }
  return lo;
}

// Next cell of the same thread (a few tiles further on): walk forward from the previous segment instead of a fresh
// dependent-load binary search (11 serial loads at 2048 images -- it was 60 % of the kernel's stall samples).
__device__ __forceinline__ int advance_segment(const int64_t* __restrict__ off, int B, int b, int64_t r) {
  for (int steps = 0; r >= __ldg(off + b + 1); ++steps) {
    ++b;
    if (steps == 8) {            // many tiny images in between: finish with a bounded binary search
      int lo = b, hi = B;
      while (hi - lo > 1) {
        int mid = (lo + hi) >> 1;
        if (__ldg(off + mid) <= r) lo = mid; else hi = mid;
      }
      return lo;
    }
  }
  return b;
}

// Writes 2*F floats for one (b1, b2) pair.  F == 23: compute_spatial_ratio_encodings (ops.py:107-155),
// F == 18: compute_spatial_encodings (ops.py:46-79).  Returns true when any value is NaN.
__device__ __forceinline__ float fdiv(float a, float b) { return __fdividef(a, b); }

template <int F>
__device__ __forceinline__ bool encode_pair(float4 b1, float4 b2, float h, float w, float hw, float eps,
                                            float* __restrict__ out) {
  const float c1_x = (b1.x + b1.z) * 0.5f, c1_y = (b1.y + b1.w) * 0.5f;     // exact: /2 == *0.5
  const float c2_x = (b2.x + b2.z) * 0.5f, c2_y = (b2.y + b2.w) * 0.5f;
  const float b1_w = b1.z - b1.x, b1_h = b1.w - b1.y;
  const float b2_w = b2.z - b2.x, b2_h = b2.w - b2.y;
  const float d_x = fdiv(fabsf(c2_x - c1_x), b1_w + eps);
  const float d_y = fdiv(fabsf(c2_y - c1_y), b1_h + eps);
  // torchvision box_iou diagonal: no eps, 0/0 -> NaN
  const float area1 = b1_w * b1_h, area2 = b2_w * b2_h;
  const float iw = fmaxf(fminf(b1.z, b2.z) - fmaxf(b1.x, b2.x), 0.f);
  const float ih = fmaxf(fminf(b1.w, b2.w) - fmaxf(b1.y, b2.y), 0.f);
  const float inter = iw * ih;
  const float iou = fdiv(inter, (area1 + area2) - inter);
  const float gx = (c2_x > c1_x) ? 1.f : 0.f, lx = (c2_x < c1_x) ? 1.f : 0.f;
  const float gy = (c2_y > c1_y) ? 1.f : 0.f, ly = (c2_y < c1_y) ? 1.f : 0.f;
  float f[F];
  if (F == 23) {
    const float c1_xw = fdiv(c1_x, w), c1_yh = fdiv(c1_y, h), c2_xw = fdiv(c2_x, w), c2_yh = fdiv(c2_y, h);
    const float b1_ww = fdiv(b1_w, w), b1_hh = fdiv(b1_h, h), b2_ww = fdiv(b2_w, w), b2_hh = fdiv(b2_h, h);
    const float a1 = fdiv(area1, hw), a2 = fdiv(area2, hw);
    f[0] = c1_xw; f[1] = c1_yh; f[2] = c2_xw; f[3] = c2_yh;
    f[4] = fdiv(c1_xw, c2_xw + eps); f[5] = fdiv(c1_yh, c2_yh + eps);
    f[6] = b1_ww; f[7] = b1_hh; f[8] = b2_ww; f[9] = b2_hh;
    f[10] = fdiv(b1_ww, b2_ww + eps); f[11] = fdiv(b1_hh, b2_hh + eps);
    f[12] = a1; f[13] = a2; f[14] = fdiv(a1, a2 + eps);
    f[15] = fdiv(area2, area1 + eps);
    f[16] = fdiv(b1_w, b1_h + eps); f[17] = fdiv(b2_w, b2_h + eps);
    f[18] = iou;
    f[19] = gx * d_x; f[20] = lx * d_x; f[21] = gy * d_y; f[22] = ly * d_y;
  } else {
    f[0] = fdiv(c1_x, w); f[1] = fdiv(c1_y, h); f[2] = fdiv(c2_x, w); f[3] = fdiv(c2_y, h);
    f[4] = fdiv(b1_w, w); f[5] = fdiv(b1_h, h); f[6] = fdiv(b2_w, w); f[7] = fdiv(b2_h, h);
    f[8] = fdiv(area1, hw); f[9] = fdiv(area2, hw);
    f[10] = fdiv(area2, area1 + eps);
    f[11] = fdiv(b1_w, b1_h + eps); f[12] = fdiv(b2_w, b2_h + eps);
    f[13] = iou;
    f[14] = gx * d_x; f[15] = lx * d_x; f[16] = gy * d_y; f[17] = ly * d_y;
  }
  bool has_nan = false;
#pragma unroll
  for (int k = 0; k < F; ++k) {
    const float l = __logf(f[k] + eps);
    out[k] = f[k];
    out[F + k] = l;
    has_nan |= (f[k] != f[k]) | (l != l);
  }
  return has_nan;
}

__device__ __forceinline__ void bulk_store_tile(float* gdst, const float* ssrc, uint32_t bytes) {
  uint32_t saddr = (uint32_t)__cvta_generic_to_shared(ssrc);
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;\n" ::"l"(gdst), "r"(saddr), "r"(bytes)
               : "memory");
  asm volatile("cp.async.bulk.commit_group;\n" ::: "memory");
}

struct PairOutputs {
  float* enc;          // [cells, 2F]
  int64_t* x;          // [cells] human index within image (may be null)
  int64_t* y;          // [cells] box index within image (may be null)
  int64_t* x_keep;     // [pairs] (may be null)
  int64_t* y_keep;     // [pairs] (may be null)
  int32_t* cell2pair;  // [cells] global pair index or -1 (may be null)
  int32_t* cell_hrow;  // [cells] global human-node row (may be null)
  int32_t* cell_orow;  // [cells] global box row (may be null)
  int32_t* cell_img;   // [cells] image index (may be null)
  int32_t* pair2cell;  // [pairs] global cell index (may be null)
  int32_t* nan_flag;   // [B] set to 1 when the image's encoding holds a NaN (may be null)
};

// Capacity mode (boxes_slot != nullptr, SKG_HEAD_CAPACITY): `total_cells` is the CAPACITY the launch was sized for, the
// true cell count is cell_off[B] of the segment table and the box array's address is read from `boxes_slot`, both in device
// memory -- nothing of the launch depends on the batch, so a captured graph serves every batch of the bucket.  Padding
// cells get zero encodings and neutral row maps (cell2pair = -1, node rows 0).
template <int F>
__global__ void __launch_bounds__(kTileCells, 8) pairs_spatial_kernel(const float4* __restrict__ boxes, SegTable seg,
                                                                      int64_t total_cells, float eps, PairOutputs o,
                                                                      const unsigned long long* __restrict__ boxes_slot) {
  // One 128-cell tile per CTA iteration, staged in shared memory in the exact output layout (23.5 KB) and drained by all
  // threads as coalesced 16-byte streaming stores; 8 CTAs per SM (188 KB of shared memory, <= 64 registers) keep 32 warps
  // of MUFU / store work in flight, so the drain of one CTA's tile overlaps the arithmetic of the seven others.
  constexpr int W = 2 * F;
  __shared__ __align__(128) float tile[kTileCells * W];
  const int tid = threadIdx.x;
  const int64_t num_tiles = (total_cells + kTileCells - 1) / kTileCells;
  int64_t true_cells = total_cells;
  if (boxes_slot) {
    boxes = reinterpret_cast<const float4*>(__ldg(boxes_slot));
    const int64_t tc = __ldg(seg.cell_off + seg.B);
    true_cells = tc < total_cells ? tc : total_cells;
  }
  // each CTA owns a CONTIGUOUS run of tiles, so a thread's image index only ever moves forward
  const int64_t t_begin = (int64_t)blockIdx.x * num_tiles / gridDim.x;
  const int64_t t_end = (int64_t)(blockIdx.x + 1) * num_tiles / gridDim.x;
  int b = -1;
  for (int64_t t = t_begin; t < t_end; ++t) {
    const int64_t r0 = t * kTileCells;
    const int64_t r = r0 + tid;
    // index work first: it does not touch the tile, so it overlaps the previous bulk store
    int boff = 0, n = 1, i = 0, j = 0;
    float4 b1 = make_float4(0.f, 0.f, 0.f, 0.f), b2 = b1;
    float h = 1.f, w = 1.f, hw = 1.f;
    const bool live = r < true_cells;
    if (!live && r < total_cells) {                  // capacity padding
      if (o.x) o.x[r] = 0;
      if (o.y) o.y[r] = 0;
      if (o.cell_hrow) o.cell_hrow[r] = 0;
      if (o.cell_orow) o.cell_orow[r] = 0;
      if (o.cell_img) o.cell_img[r] = 0;
      if (o.cell2pair) o.cell2pair[r] = -1;
    }
    if (live) {
      b = (b < 0) ? find_segment(seg.cell_off, seg.B, r) : advance_segment(seg.cell_off, seg.B, b, r);
      boff = __ldg(seg.box_off + b);
      n = __ldg(seg.box_off + b + 1) - boff;
      const int64_t c = r - __ldg(seg.cell_off + b);
      i = (int)(c / n); j = (int)(c - (int64_t)i * n);
      b1 = __ldg(boxes + boff + i);
      b2 = __ldg(boxes + boff + j);
      h = __ldg(seg.hw + 3 * b); w = __ldg(seg.hw + 3 * b + 1); hw = __ldg(seg.hw + 3 * b + 2);
      if (o.x) o.x[r] = i;
      if (o.y) o.y[r] = j;
      if (o.cell_hrow) o.cell_hrow[r] = __ldg(seg.hum_off + b) + i;
      if (o.cell_orow) o.cell_orow[r] = boff + j;
      if (o.cell_img) o.cell_img[r] = b;
      const bool valid = (i != j);
      const int64_t p = __ldg(seg.pair_off + b) + (int64_t)i * (n - 1) + j - (j > i ? 1 : 0);
      if (o.cell2pair) o.cell2pair[r] = valid ? (int32_t)p : -1;
      if (valid) {
        if (o.x_keep) o.x_keep[p] = i;
        if (o.y_keep) o.y_keep[p] = j;
        if (o.pair2cell) o.pair2cell[p] = (int32_t)r;
      }
    }
    // the bulk store of the previous iteration must have finished READING the tile
    if (tid == 0) asm volatile("cp.async.bulk.wait_group.read 0;\n" ::: "memory");
    __syncthreads();
    if (live) {
      const bool has_nan = encode_pair<F>(b1, b2, h, w, hw, eps, &tile[tid * W]);
      if (has_nan && o.nan_flag) atomicOr(o.nan_flag + b, 1);
    } else if (r < total_cells) {
#pragma unroll
      for (int k = 0; k < W; ++k) tile[tid * W + k] = 0.f;
    }
    // make the generic-proxy smem writes visible to the async (TMA) proxy, then drain the tile
    asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
    __syncthreads();
    const int64_t cells_here = (total_cells - r0 < kTileCells) ? (total_cells - r0) : kTileCells;
    float* gdst = o.enc + r0 * W;
    const uint32_t bytes = (uint32_t)(cells_here * W * sizeof(float));
    if (kDrainBulk && (bytes & 15u) == 0 && ((reinterpret_cast<uintptr_t>(gdst) & 15u) == 0)) {
      if (tid == 0) bulk_store_tile(gdst, tile, bytes);
    } else if ((bytes & 15u) == 0 && ((reinterpret_cast<uintptr_t>(gdst) & 15u) == 0)) {
      // fire-and-forget 16-byte stores, consecutive lanes on consecutive addresses (full 128-byte lines)
      const float4* src = reinterpret_cast<const float4*>(tile);
      float4* dst = reinterpret_cast<float4*>(gdst);
      const int n4 = (int)(bytes >> 4);
#pragma unroll 4
      for (int e = tid; e < n4; e += kTileCells) __stcs(dst + e, src[e]);
      __syncthreads();
    } else {  // ragged tail: plain coalesced stores
      for (int e = tid; e < cells_here * W; e += kTileCells) gdst[e] = tile[e];
      __syncthreads();
    }
  }
  if (tid == 0) asm volatile("cp.async.bulk.wait_group 0;\n" ::: "memory");
}

// ops.compute_spatial_encodings / compute_spatial_ratio_encodings on explicit (boxes_1, boxes_2) rows.
template <int F>
__global__ void __launch_bounds__(kTileCells) spatial_rows_kernel(const float4* __restrict__ b1s,
                                                                  const float4* __restrict__ b2s,
                                                                  const int64_t* __restrict__ row_off,
                                                                  const float* __restrict__ hw, int num_lists,
                                                                  int64_t rows, float eps, float* __restrict__ enc,
                                                                  int32_t* __restrict__ nan_flag) {
  constexpr int W = 2 * F;
  __shared__ __align__(128) float tile[kTileCells * W];
  const int tid = threadIdx.x;
  const int64_t num_tiles = (rows + kTileCells - 1) / kTileCells;
  for (int64_t t = blockIdx.x; t < num_tiles; t += gridDim.x) {
    const int64_t r0 = t * kTileCells, r = r0 + tid;
    __syncthreads();
    if (r < rows) {
      const int b = find_segment(row_off, num_lists, r);
      const bool has_nan = encode_pair<F>(__ldg(b1s + r), __ldg(b2s + r), __ldg(hw + 3 * b), __ldg(hw + 3 * b + 1),
                                          __ldg(hw + 3 * b + 2), eps, &tile[tid * W]);
      if (has_nan && nan_flag) atomicOr(nan_flag + b, 1);
    }
    __syncthreads();
    const int64_t here = (rows - r0 < kTileCells) ? (rows - r0) : kTileCells;
    for (int e = tid; e < here * W; e += kTileCells) enc[r0 * W + e] = tile[e];
  }
}

// head.py:866-868: only when the image's encoding holds a NaN, torch.nan_to_num the whole tensor
// (NaN -> 0, +inf -> FLT_MAX, -inf -> -FLT_MAX); images without NaN keep their infs.
__global__ void nan_guard_kernel(float* __restrict__ enc, const int64_t* __restrict__ cell_off,
                                 const int32_t* __restrict__ nan_flag, int B, int W) {
  const int b = blockIdx.y;
  if (b >= B || nan_flag[b] == 0) return;
  const int64_t lo = cell_off[b] * W, hi = cell_off[b + 1] * W;
  for (int64_t e = lo + blockIdx.x * (int64_t)blockDim.x + threadIdx.x; e < hi; e += (int64_t)gridDim.x * blockDim.x) {
    float v = enc[e];
    if (v != v) v = 0.f;
    else if (isinf(v)) v = v > 0 ? 3.4028234663852886e38f : -3.4028234663852886e38f;
    enc[e] = v;
  }
}

}  // namespace skg

using namespace skg;

extern "C" int skg_pairs_spatial_encode(const float* boxes, const int32_t* box_off, const int32_t* n_h,
                                        const int64_t* cell_off, const int64_t* pair_off, const int32_t* hum_off,
                                        const float* hw, int num_images, int64_t total_cells, int num_features,
                                        float eps, int nan_guard, float* enc, int64_t* x, int64_t* y,
                                        int64_t* x_keep, int64_t* y_keep, int32_t* cell2pair, int32_t* cell_hrow,
                                        int32_t* cell_orow, int32_t* cell_img, int32_t* pair2cell, int32_t* nan_flag,
                                        cudaStream_t stream) {
  return skg::pairs_spatial_encode_impl(boxes, nullptr, box_off, n_h, cell_off, pair_off, hum_off, hw, num_images, total_cells,
                                        num_features, eps, nan_guard, enc, x, y, x_keep, y_keep, cell2pair, cell_hrow, cell_orow,
                                        cell_img, pair2cell, nan_flag, stream);
}

int skg::pairs_spatial_encode_impl(const float* boxes, const unsigned long long* boxes_slot, const int32_t* box_off, const int32_t* n_h,
                                   const int64_t* cell_off, const int64_t* pair_off, const int32_t* hum_off,
                                   const float* hw, int num_images, int64_t total_cells, int num_features,
                                   float eps, int nan_guard, float* enc, int64_t* x, int64_t* y,
                                   int64_t* x_keep, int64_t* y_keep, int32_t* cell2pair, int32_t* cell_hrow,
                                   int32_t* cell_orow, int32_t* cell_img, int32_t* pair2cell, int32_t* nan_flag,
                                   cudaStream_t stream) {
  SKG_REQUIRE(num_features == 18 || num_features == 23, "num_features must be 18 or 23, got %d", num_features);
  SKG_REQUIRE(num_images >= 0 && total_cells >= 0, "negative sizes");
  SKG_REQUIRE(!nan_guard || nan_flag, "nan_guard needs a nan_flag buffer");
  if (total_cells == 0 || num_images == 0) return SKG_OK;
  SKG_REQUIRE((boxes || boxes_slot) && box_off && n_h && cell_off && pair_off && hum_off && hw && enc, "null pointer argument");
  SKG_REQUIRE((reinterpret_cast<uintptr_t>(boxes) & 15u) == 0, "boxes must be 16-byte aligned");
  SegTable seg{box_off, n_h, cell_off, pair_off, hum_off, hw, num_images};
  PairOutputs o{enc, x, y, x_keep, y_keep, cell2pair, cell_hrow, cell_orow, cell_img, pair2cell, nan_flag};
  if (nan_flag) SKG_CUDA(cudaMemsetAsync(nan_flag, 0, sizeof(int32_t) * num_images, stream));
  const int64_t tiles = ceil_div64(total_cells, kTileCells);
  // 8 resident 128-thread CTAs per SM (23.5 KB smem each); grid is a multiple of the SM count
  const int grid = (int)(tiles < (int64_t)num_sms() * 8 ? tiles : (int64_t)num_sms() * 8);
  if (num_features == 23)
    pairs_spatial_kernel<23><<<grid, kTileCells, 0, stream>>>((const float4*)boxes, seg, total_cells, eps, o, boxes_slot);
  else
    pairs_spatial_kernel<18><<<grid, kTileCells, 0, stream>>>((const float4*)boxes, seg, total_cells, eps, o, boxes_slot);
  int rc = check_launch("pairs_spatial_kernel");
  if (rc) return rc;
  if (nan_guard) {
    const int gx = 4;  // clean images exit at once; a flagged image is at most a few thousand rows
    for (int b0 = 0; b0 < num_images; b0 += 65535) {
      int nb = num_images - b0 < 65535 ? num_images - b0 : 65535;
      nan_guard_kernel<<<dim3(gx, nb), 256, 0, stream>>>(enc, cell_off + b0, nan_flag + b0, nb, 2 * num_features);
    }
    rc = check_launch("nan_guard_kernel");
  }
  return rc;
}

extern "C" int skg_spatial_encode_rows(const float* boxes_1, const float* boxes_2, const int64_t* row_off,
                                       const float* hw, int num_lists, int64_t rows, int num_features, float eps,
                                       float* enc, int32_t* nan_flag, cudaStream_t stream) {
  SKG_REQUIRE(num_features == 18 || num_features == 23, "num_features must be 18 or 23, got %d", num_features);
  if (rows == 0) return SKG_OK;
  SKG_REQUIRE(boxes_1 && boxes_2 && row_off && hw && enc && num_lists > 0, "null pointer argument");
  SKG_REQUIRE(((reinterpret_cast<uintptr_t>(boxes_1) | reinterpret_cast<uintptr_t>(boxes_2)) & 15u) == 0,
              "box arrays must be 16-byte aligned");
  if (nan_flag) SKG_CUDA(cudaMemsetAsync(nan_flag, 0, sizeof(int32_t) * num_lists, stream));
  const int64_t tiles = ceil_div64(rows, kTileCells);
  const int grid = (int)(tiles < (int64_t)num_sms() * 8 ? tiles : (int64_t)num_sms() * 8);
  if (num_features == 23)
    spatial_rows_kernel<23><<<grid, kTileCells, 0, stream>>>((const float4*)boxes_1, (const float4*)boxes_2, row_off,
                                                            hw, num_lists, rows, eps, enc, nan_flag);
  else
    spatial_rows_kernel<18><<<grid, kTileCells, 0, stream>>>((const float4*)boxes_1, (const float4*)boxes_2, row_off,
                                                            hw, num_lists, rows, eps, enc, nan_flag);
  return check_launch("spatial_rows_kernel");
}
