// Bandwidth-bound glue kernels of the interaction head that sit between the tensor-core GEMMs:
// casts / packing, the fp32 first spatial layer, node-input assembly with the TransH entity rows,
// global average pooling, softmax-weighted message aggregation, LayerNorm, and the branch product
// of the final attention head.  (Reference heads/adamixer_transH_spatial_r50_head.py:811, 884-925,
// 966-973; each kernel cites the lines it covers.)
#include "common.cuh"
#include "skghoi_b200.h"

namespace skg {

__device__ __forceinline__ uint32_t pack2(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ float2 unpack2(uint32_t u) {
  __nv_bfloat162 v = *reinterpret_cast<__nv_bfloat162*>(&u);
  return __bfloat1622float2(v);
}

// dst[r, c] = bf16(relu?(src[r, c] + bias[c])) for c < cols, 0 for cols <= c < cols_pad
// Capacity mode (SKG_HEAD_CAPACITY): `src_slot` holds the source ADDRESS and `rows_dev` the true row count in device
// memory (the step's segment table), rows in [*rows_dev, rows) are written as zeros -- the launch itself depends on the
// capacity only, so a captured graph serves every batch of the bucket.
__global__ void cast_pad_bf16_kernel(const float* __restrict__ src, int64_t ld_src, const float* __restrict__ bias,
                                     int relu, int64_t rows, int cols, __nv_bfloat16* __restrict__ dst, int64_t ld_dst,
                                     int cols_pad, const unsigned long long* __restrict__ src_slot,
                                     const int32_t* __restrict__ rows_dev) {
  const int64_t quads_per_row = cols_pad >> 2;   // cols_pad % 4 == 0
  const int64_t total = rows * quads_per_row;
  if (src_slot) src = reinterpret_cast<const float*>(__ldg(src_slot));
  const int64_t rows_true = rows_dev ? (int64_t)__ldg(rows_dev) : rows;
  for (int64_t q = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; q < total; q += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = q / quads_per_row;
    const int c = (int)(q - r * quads_per_row) << 2;
    if (r >= rows_true) { *reinterpret_cast<uint2*>(dst + r * ld_dst + c) = make_uint2(0u, 0u); continue; }
    float v[4];
    const float* s = src + r * ld_src + c;
    if (c + 4 <= cols && ((ld_src & 3) == 0) && ((reinterpret_cast<uintptr_t>(src) & 15u) == 0)) {
      const float4 f = __ldg(reinterpret_cast<const float4*>(s));
      v[0] = f.x; v[1] = f.y; v[2] = f.z; v[3] = f.w;
    } else {
#pragma unroll
      for (int j = 0; j < 4; ++j) v[j] = (c + j < cols) ? __ldg(s + j) : 0.f;
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (c + j < cols) {
        if (bias) v[j] += __ldg(bias + c + j);
        if (relu) v[j] = fmaxf(v[j], 0.f);
      }
    }
    *reinterpret_cast<uint2*>(dst + r * ld_dst + c) = make_uint2(pack2(v[0], v[1]), pack2(v[2], v[3]));
  }
}

// spatial_head layer 0 (head.py:663-664): out = bf16(relu(enc[C, in] @ W[out, in]^T + b)) in fp32 CUDA-core math --
// the 46 hand-crafted features span [-23, 1e3], so this layer is kept out of bf16.
template <int IN, int OUT>
__global__ void __launch_bounds__(OUT) spatial_fc1_kernel(const float* __restrict__ enc, const float* __restrict__ W,
                                                          const float* __restrict__ b, int64_t cells,
                                                          __nv_bfloat16* __restrict__ out) {
  constexpr int ROWS = 32;
  __shared__ float sW[IN][OUT + 1];     // transposed weights: sW[k][o]
  __shared__ float sX[ROWS][IN];
  const int o = threadIdx.x;
  for (int e = threadIdx.x; e < IN * OUT; e += OUT) { const int oo = e / IN, k = e - oo * IN; sW[k][oo] = __ldg(W + e); }
  const float bias = __ldg(b + o);
  for (int64_t r0 = (int64_t)blockIdx.x * ROWS; r0 < cells; r0 += (int64_t)gridDim.x * ROWS) {
    __syncthreads();
    const int nrow = (int)((cells - r0 < ROWS) ? (cells - r0) : ROWS);
    for (int e = threadIdx.x; e < nrow * IN; e += OUT) sX[e / IN][e % IN] = __ldg(enc + r0 * IN + e);
    __syncthreads();
    for (int r = 0; r < nrow; ++r) {
      float acc = bias;
#pragma unroll
      for (int k = 0; k < IN; ++k) acc = fmaf(sX[r][k], sW[k][o], acc);
      out[(r0 + r) * OUT + o] = __float2bfloat16(fmaxf(acc, 0.f));
    }
  }
}

// Inputs of fc_head / fc_tail (head.py:884-885): cat(node_enc[row], E[entity]) as bf16, zero padded to ld.
//   humans : row = box_off[b] + i            entity = human_idx          (head_stack == E[human_idx], head.py:571,877-882)
//   objects: row = box_off[b] + j            entity = j (box index!)     (tail_stack == E[y],        head.py:572)
__global__ void node_inputs_kernel(const __nv_bfloat16* __restrict__ node_enc, int enc_dim, const float* __restrict__ ent,
                                   int ent_dim, int ent_tot, int human_idx, const int32_t* __restrict__ box_off,
                                   const int32_t* __restrict__ hum_off, int B, int total_boxes, int total_humans,
                                   __nv_bfloat16* __restrict__ h_in, __nv_bfloat16* __restrict__ o_in, int ld) {
  const int row = blockIdx.x;                       // [0, total_humans) humans, then objects
  const bool is_h = row < total_humans;
  const int local = is_h ? row : row - total_humans;
  const int32_t* off = is_h ? hum_off : box_off;
  int lo = 0, hi = B;
  while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (__ldg(off + mid) <= local) lo = mid; else hi = mid; }
  const int idx_in_img = local - __ldg(off + lo);
  const int src_row = __ldg(box_off + lo) + idx_in_img;
  int entity = is_h ? human_idx : idx_in_img;
  if (entity >= ent_tot) entity = ent_tot - 1;      // host rejects n > ent_tot (reference raises IndexError)
  __nv_bfloat16* dst = (is_h ? h_in : o_in) + (int64_t)local * ld;
  const __nv_bfloat16* src = node_enc + (int64_t)src_row * enc_dim;
  for (int c = threadIdx.x; c < ld; c += blockDim.x) {
    __nv_bfloat16 v;
    if (c < enc_dim) v = src[c];
    else if (c < enc_dim + ent_dim) v = __float2bfloat16(__ldg(ent + (int64_t)entity * ent_dim + (c - enc_dim)));
    else v = __float2bfloat16(0.f);
    dst[c] = v;
  }
}

// Per-box preparation for the fused fc_head / fc_tail GEMM (head.py:875-885): the TransH entity columns of the
// node input, o_in[row, enc_dim : ld] = bf16(E[index of the box inside its image]) zero padded (tail_stack == E[y],
// head.py:572; the human side's constant E[human_idx] term is folded into the bias by the host), and the
// box -> human-row map (human rows are the first n_h boxes of every image, -1 otherwise).
__global__ void node_prep_kernel(const float* __restrict__ ent, int ent_dim, int ent_tot, const int32_t* __restrict__ box_off,
                                 const int32_t* __restrict__ n_h, const int32_t* __restrict__ hum_off, int B,
                                 __nv_bfloat16* __restrict__ o_in, int ld, int enc_dim, int32_t* __restrict__ box2hum) {
  const int row = blockIdx.x;
  if (row >= __ldg(box_off + B)) {                    // capacity padding: no box behind this row
    for (int c = enc_dim + threadIdx.x; c < ld; c += blockDim.x) o_in[(int64_t)row * ld + c] = __float2bfloat16(0.f);
    if (threadIdx.x == 0 && box2hum) box2hum[row] = -1;
    return;
  }
  int lo = 0, hi = B;
  while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (__ldg(box_off + mid) <= row) lo = mid; else hi = mid; }
  const int idx = row - __ldg(box_off + lo);
  int entity = idx < ent_tot ? idx : ent_tot - 1;     // host rejects n > ent_tot (reference raises IndexError)
  for (int c = enc_dim + threadIdx.x; c < ld; c += blockDim.x)
    o_in[(int64_t)row * ld + c] = __float2bfloat16(c < enc_dim + ent_dim ? __ldg(ent + (int64_t)entity * ent_dim + (c - enc_dim)) : 0.f);
  if (threadIdx.x == 0 && box2hum) box2hum[row] = idx < __ldg(n_h + lo) ? __ldg(hum_off + lo) + idx : -1;
}

// bias_out[r] = bias_in[r] + W_ent[r, :] . ent[entity, :]   (one warp per row)
__global__ void fc_bias_fold_kernel(const float* __restrict__ bias_in, const float* __restrict__ W_ent, const float* __restrict__ ent,
                                    int ent_dim, int entity, int rows, float* __restrict__ bias_out) {
  const int lane = threadIdx.x & 31;
  const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (row >= rows) return;
  float s = 0.f;
  for (int k = lane; k < ent_dim; k += 32) s = fmaf(__ldg(W_ent + (int64_t)row * ent_dim + k), __ldg(ent + (int64_t)entity * ent_dim + k), s);
  s = warp_sum(s);
  if (lane == 0) bias_out[row] = __ldg(bias_in + row) + s;
}

// global_features = avg_pool(features['3']).flatten(1) (head.py:811): one warp per (image, channel)
__global__ void global_pool_kernel(const float* __restrict__ f, int64_t planes, int hw, __nv_bfloat16* __restrict__ out_bf16,
                                   float* __restrict__ out_f32, const unsigned long long* __restrict__ f_slot) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  if (warp >= planes) return;
  if (f_slot) f = reinterpret_cast<const float*>(__ldg(f_slot));   // capacity mode: the address lives in the segment table
  const float* p = f + warp * hw;
  float s = 0.f;
  for (int i = lane; i < hw; i += 32) s += __ldg(p + i);
  s = warp_sum(s);
  if (lane == 0) {
    const float m = s / (float)hw;
    if (out_bf16) out_bf16[warp] = __float2bfloat16(m);
    if (out_f32) out_f32[warp] = m;
  }
}

// Message aggregation (head.py:897, 909-910, 916-921) for every image at once.
//   adjacency[i, j] = sum_t dot_part[cell, t] + adj_bias                      (adjacency Linear(1024->1))
//   Z_h[i, :] = sum_j softmax_j(adjacency[i, :])[j] * H_os[cell(i, j), :]     (to be multiplied by W3 of obj_to_sub)
//   Z_o[j, :] = sum_i softmax_i(adjacency[:, j])[i] * H_so[cell(i, j), :]     (                        sub_to_obj)
// fc_3 of MessageMBF has no outer ReLU and softmax weights sum to 1, so it commutes with the weighted sum and
// is applied afterwards to n_h (resp. n) rows instead of n_h*n.  One CTA per output row, 128 threads x 8 columns.
constexpr int kMaxPerImage = 1024;
__global__ void __launch_bounds__(128) aggregate_kernel(const float* __restrict__ dot_part, int num_part, float adj_bias,
                                                        const __nv_bfloat16* __restrict__ H_os, const __nv_bfloat16* __restrict__ H_so,
                                                        int64_t ld_h, const int32_t* __restrict__ box_off,
                                                        const int32_t* __restrict__ hum_off, const int64_t* __restrict__ cell_off,
                                                        int B, int total_humans, int width, __nv_bfloat16* __restrict__ Z_h,
                                                        __nv_bfloat16* __restrict__ Z_o, float* __restrict__ adjacency) {
  __shared__ float sw[kMaxPerImage];
  __shared__ float red[4];
  const int row = blockIdx.x;
  const bool is_h = row < total_humans;
  const int local = is_h ? row : row - total_humans;
  const int32_t* off = is_h ? hum_off : box_off;
  if (local >= __ldg(off + B)) return;             // capacity padding (SKG_HEAD_CAPACITY): no node behind this row
  int lo = 0, hi = B;
  while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (__ldg(off + mid) <= local) lo = mid; else hi = mid; }
  const int b = lo;
  const int n = __ldg(box_off + b + 1) - __ldg(box_off + b);
  const int n_h = __ldg(hum_off + b + 1) - __ldg(hum_off + b);
  const int idx = local - __ldg(off + b);
  const int64_t c0 = __ldg(cell_off + b);
  const int cnt = is_h ? n : n_h;                 // softmax length
  if (cnt == 0) {                                  // object row of an image without humans: zero message
    for (int c = threadIdx.x * 8; c < width; c += 128 * 8) *reinterpret_cast<uint4*>((is_h ? Z_h : Z_o) + (int64_t)local * width + c) = make_uint4(0, 0, 0, 0);
    return;
  }
  // cell(i, j) = c0 + i*n + j ; human row: fixed i = idx, j runs ; object row: fixed j = idx, i runs
  const int64_t base = is_h ? c0 + (int64_t)idx * n : c0 + idx;
  const int64_t step = is_h ? 1 : n;
  float mx = -INFINITY;
  for (int t = threadIdx.x; t < cnt; t += 128) {
    const int64_t cell = base + t * step;
    float a = adj_bias;
    for (int k = 0; k < num_part; ++k) a += __ldg(dot_part + cell * num_part + k);
    sw[t] = a;
    if (is_h && adjacency) adjacency[cell] = a;
    mx = fmaxf(mx, a);
  }
  mx = warp_max(mx);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = mx;
  __syncthreads();
  mx = fmaxf(fmaxf(red[0], red[1]), fmaxf(red[2], red[3]));
  __syncthreads();
  float s = 0.f;
  for (int t = threadIdx.x; t < cnt; t += 128) { const float e = expf(sw[t] - mx); sw[t] = e; s += e; }
  s = warp_sum(s);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  const float inv = 1.f / (red[0] + red[1] + red[2] + red[3]);
  const __nv_bfloat16* H = is_h ? H_os : H_so;
  for (int c = threadIdx.x * 8; c < width; c += 128 * 8) {
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    // eight message rows in flight per thread (the rows of an object's column are n * ld_h apart: every load is its own
    // L2 round trip), accumulated in the original order
    for (int t0 = 0; t0 < cnt; t0 += 8) {
      uint4 u[8];
#pragma unroll
      for (int k = 0; k < 8; ++k)
        u[k] = (t0 + k < cnt) ? __ldg(reinterpret_cast<const uint4*>(H + (base + (t0 + k) * step) * ld_h + c)) : make_uint4(0, 0, 0, 0);
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        if (t0 + k >= cnt) break;
        const float w = sw[t0 + k] * inv;
        const float2 a = unpack2(u[k].x), b2 = unpack2(u[k].y), c2 = unpack2(u[k].z), d = unpack2(u[k].w);
        acc[0] = fmaf(w, a.x, acc[0]); acc[1] = fmaf(w, a.y, acc[1]); acc[2] = fmaf(w, b2.x, acc[2]); acc[3] = fmaf(w, b2.y, acc[3]);
        acc[4] = fmaf(w, c2.x, acc[4]); acc[5] = fmaf(w, c2.y, acc[5]); acc[6] = fmaf(w, d.x, acc[6]); acc[7] = fmaf(w, d.y, acc[7]);
      }
    }
    *reinterpret_cast<uint4*>((is_h ? Z_h : Z_o) + (int64_t)local * width + c) =
        make_uint4(pack2(acc[0], acc[1]), pack2(acc[2], acc[3]), pack2(acc[4], acc[5]), pack2(acc[6], acc[7]));
  }
}

// LayerNorm over the last dimension (head.py:912-914, 923-925; eps 1e-5, affine): one warp per row.
// Rows [0, rows_a) use (w_a, b_a) = norm_h, rows [rows_a, rows) use (w_b, b_b) = norm_o.
__global__ void layernorm_kernel(const float* __restrict__ x, int64_t rows, int64_t rows_a, int width,
                                 const float* __restrict__ w_a, const float* __restrict__ b_a, const float* __restrict__ w_b,
                                 const float* __restrict__ b_b, float eps, float* __restrict__ out_f32,
                                 __nv_bfloat16* __restrict__ out_bf16) {
  const int lane = threadIdx.x & 31;
  const int64_t row = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  if (row >= rows) return;
  const float* xr = x + row * width;
  float s = 0.f;
  for (int c = lane; c < width; c += 32) s += xr[c];
  const float mean = warp_sum(s) / (float)width;
  float v = 0.f;
  for (int c = lane; c < width; c += 32) { const float d = xr[c] - mean; v += d * d; }
  const float rstd = rsqrtf(warp_sum(v) / (float)width + eps);
  const float* w = row < rows_a ? w_a : w_b;
  const float* b = row < rows_a ? b_a : b_b;
  for (int c = lane; c < width; c += 32) {
    const float y = (xr[c] - mean) * rstd * __ldg(w + c) + __ldg(b + c);
    if (out_f32) out_f32[row * width + c] = y;
    if (out_bf16) out_bf16[row * width + c] = __float2bfloat16(y);
  }
}

// LayerNorm, one 256-thread CTA per row, the row held in registers (VPT float4s per thread): one pass over global
// memory instead of three, every load in flight at once.  Same two-pass mean / variance formula as above.
template <int VPT>
__global__ void __launch_bounds__(256) layernorm_row_kernel(const float* __restrict__ x, int64_t rows_a, int width,
                                                            const float* __restrict__ w_a, const float* __restrict__ b_a,
                                                            const float* __restrict__ w_b, const float* __restrict__ b_b,
                                                            float eps, float* __restrict__ out_f32,
                                                            __nv_bfloat16* __restrict__ out_bf16) {
  __shared__ float red[8];
  const int64_t row = blockIdx.x;
  const int n4 = width >> 2;
  const float4* xr = reinterpret_cast<const float4*>(x + row * width);
  float4 v[VPT];
  float s = 0.f;
#pragma unroll
  for (int j = 0; j < VPT; ++j) {
    const int q = threadIdx.x + 256 * j;
    v[j] = q < n4 ? __ldg(xr + q) : make_float4(0.f, 0.f, 0.f, 0.f);
    s += (v[j].x + v[j].y) + (v[j].z + v[j].w);
  }
  s = warp_sum(s);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  float tot = 0.f;
#pragma unroll
  for (int k = 0; k < 8; ++k) tot += red[k];
  const float mean = tot / (float)width;
  __syncthreads();
  float sq = 0.f;
#pragma unroll
  for (int j = 0; j < VPT; ++j) {
    if (threadIdx.x + 256 * j < n4) {
      const float a = v[j].x - mean, b = v[j].y - mean, c = v[j].z - mean, d = v[j].w - mean;
      sq += (a * a + b * b) + (c * c + d * d);
    }
  }
  sq = warp_sum(sq);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = sq;
  __syncthreads();
  tot = 0.f;
#pragma unroll
  for (int k = 0; k < 8; ++k) tot += red[k];
  const float rstd = rsqrtf(tot / (float)width + eps);
  const float4* w = reinterpret_cast<const float4*>(row < rows_a ? w_a : w_b);
  const float4* b = reinterpret_cast<const float4*>(row < rows_a ? b_a : b_b);
#pragma unroll
  for (int j = 0; j < VPT; ++j) {
    const int q = threadIdx.x + 256 * j;
    if (q >= n4) continue;
    const float4 g = __ldg(w + q), o = __ldg(b + q);
    const float4 y = make_float4((v[j].x - mean) * rstd * g.x + o.x, (v[j].y - mean) * rstd * g.y + o.y,
                                 (v[j].z - mean) * rstd * g.z + o.z, (v[j].w - mean) * rstd * g.w + o.w);
    if (out_f32) reinterpret_cast<float4*>(out_f32 + row * width)[q] = y;
    if (out_bf16) reinterpret_cast<uint2*>(out_bf16 + row * width)[q] = make_uint2(pack2(y.x, y.y), pack2(y.z, y.w));
  }
}

// Branch product of the final attention head (head.py:966-970 with MBF head.py:469-474):
//   H[cell, :] = bf16(relu(S_raw[cell, :] * (U_h[hrow[cell], :] + U_o[orow[cell], :])))
__global__ void hadamard_cells_kernel(const __nv_bfloat16* __restrict__ S_raw, int64_t ld_s, const float* __restrict__ U_h,
                                      int64_t ld_uh, const float* __restrict__ U_o, int64_t ld_uo,
                                      const int32_t* __restrict__ hrow, const int32_t* __restrict__ orow, int64_t cells,
                                      int width, __nv_bfloat16* __restrict__ H, int64_t ld_hh) {
  const int chunks = width >> 3;
  const int64_t total = cells * chunks;
  for (int64_t q = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; q < total; q += (int64_t)gridDim.x * blockDim.x) {
    const int64_t cell = q / chunks;
    const int c = (int)(q - cell * chunks) << 3;
    const uint4 u = __ldg(reinterpret_cast<const uint4*>(S_raw + cell * ld_s + c));
    const float* ph = U_h + (int64_t)__ldg(hrow + cell) * ld_uh + c;
    const float* po = U_o + (int64_t)__ldg(orow + cell) * ld_uo + c;
    const float4 h0 = __ldg(reinterpret_cast<const float4*>(ph)), h1 = __ldg(reinterpret_cast<const float4*>(ph + 4));
    const float4 o0 = __ldg(reinterpret_cast<const float4*>(po)), o1 = __ldg(reinterpret_cast<const float4*>(po + 4));
    const float2 a = unpack2(u.x), b = unpack2(u.y), c2 = unpack2(u.z), d = unpack2(u.w);
    const uint4 r = make_uint4(pack2(fmaxf(a.x * (h0.x + o0.x), 0.f), fmaxf(a.y * (h0.y + o0.y), 0.f)),
                               pack2(fmaxf(b.x * (h0.z + o0.z), 0.f), fmaxf(b.y * (h0.w + o0.w), 0.f)),
                               pack2(fmaxf(c2.x * (h1.x + o1.x), 0.f), fmaxf(c2.y * (h1.y + o1.y), 0.f)),
                               pack2(fmaxf(d.x * (h1.z + o1.z), 0.f), fmaxf(d.y * (h1.w + o1.w), 0.f)));
    *reinterpret_cast<uint4*>(H + cell * ld_hh + c) = r;
  }
}

static inline int grid_for(int64_t work_items, int threads, int per_sm = 8) {
  const int64_t want = ceil_div64(work_items, threads);
  const int64_t cap = (int64_t)num_sms() * per_sm;
  return (int)(want < cap ? (want < 1 ? 1 : want) : cap);
}

}  // namespace skg

using namespace skg;

extern "C" int skg_cast_pad_bf16(const float* src, int64_t ld_src, const float* bias, int relu, int64_t rows, int cols,
                                 void* dst, int64_t ld_dst, int cols_pad, cudaStream_t stream) {
  if (rows == 0) return SKG_OK;
  SKG_REQUIRE(src && dst && cols > 0 && cols_pad >= cols && cols_pad % 4 == 0 && ld_dst >= cols_pad && ld_dst % 4 == 0,
              "bad cast_pad arguments");
  cast_pad_bf16_kernel<<<grid_for(rows * (cols_pad / 4), 256), 256, 0, stream>>>(src, ld_src, bias, relu, rows, cols,
                                                                              (__nv_bfloat16*)dst, ld_dst, cols_pad, nullptr, nullptr);
  return check_launch("cast_pad_bf16_kernel");
}

namespace skg {
int cast_pad_bf16_indirect(const unsigned long long* src_slot, int64_t ld_src, const int32_t* rows_dev, int64_t rows_cap, int cols,
                           void* dst, int64_t ld_dst, int cols_pad, cudaStream_t stream) {
  if (rows_cap == 0) return SKG_OK;
  SKG_REQUIRE(src_slot && rows_dev && dst && cols > 0 && cols_pad >= cols && cols_pad % 4 == 0 && ld_dst >= cols_pad && ld_dst % 4 == 0,
              "bad cast_pad arguments");
  cast_pad_bf16_kernel<<<grid_for(rows_cap * (cols_pad / 4), 256), 256, 0, stream>>>(nullptr, ld_src, nullptr, 0, rows_cap, cols,
                                                                                  (__nv_bfloat16*)dst, ld_dst, cols_pad, src_slot, rows_dev);
  return check_launch("cast_pad_bf16_kernel");
}
int global_pool_indirect(const unsigned long long* f_slot, int64_t planes, int hw, void* out_bf16, float* out_f32, cudaStream_t stream) {
  if (planes == 0) return SKG_OK;
  SKG_REQUIRE(f_slot && hw > 0 && (out_bf16 || out_f32), "bad global_pool arguments");
  global_pool_kernel<<<(int)ceil_div64(planes * 32, 256), 256, 0, stream>>>(nullptr, planes, hw, (__nv_bfloat16*)out_bf16, out_f32, f_slot);
  return check_launch("global_pool_kernel");
}
}  // namespace skg

extern "C" int skg_spatial_fc1(const float* enc, const float* weight, const float* bias, int64_t cells, int in_features,
                               int out_features, void* out_bf16, cudaStream_t stream) {
  if (cells == 0) return SKG_OK;
  SKG_REQUIRE(enc && weight && bias && out_bf16, "null pointer argument");
  SKG_REQUIRE(in_features == 46 && out_features == 128, "spatial_fc1 supports 46 -> 128 (head.py:663), got %d -> %d",
              in_features, out_features);
  const int grid = grid_for(ceil_div64(cells, 32), 1, 4);
  spatial_fc1_kernel<46, 128><<<grid, 128, 0, stream>>>(enc, weight, bias, cells, (__nv_bfloat16*)out_bf16);
  return check_launch("spatial_fc1_kernel");
}

extern "C" int skg_node_inputs(const void* node_enc_bf16, int enc_dim, const float* ent, int ent_dim, int ent_tot,
                               int human_idx, const int32_t* box_off, const int32_t* hum_off, int num_images,
                               int total_boxes, int total_humans, void* h_in, void* o_in, int ld, cudaStream_t stream) {
  if (total_boxes + total_humans == 0) return SKG_OK;
  SKG_REQUIRE(node_enc_bf16 && ent && box_off && hum_off && h_in && o_in, "null pointer argument");
  SKG_REQUIRE(ld >= enc_dim + ent_dim && human_idx >= 0 && human_idx < ent_tot, "bad node_inputs sizes");
  node_inputs_kernel<<<total_boxes + total_humans, 128, 0, stream>>>((const __nv_bfloat16*)node_enc_bf16, enc_dim, ent,
                                                                   ent_dim, ent_tot, human_idx, box_off, hum_off, num_images,
                                                                   total_boxes, total_humans, (__nv_bfloat16*)h_in,
                                                                   (__nv_bfloat16*)o_in, ld);
  return check_launch("node_inputs_kernel");
}

extern "C" int skg_node_prep(const float* ent, int ent_dim, int ent_tot, const int32_t* box_off, const int32_t* n_h,
                             const int32_t* hum_off, int num_images, int total_boxes, void* o_in, int ld, int enc_dim,
                             int32_t* box2hum, cudaStream_t stream) {
  if (total_boxes == 0) return SKG_OK;
  SKG_REQUIRE(ent && box_off && n_h && hum_off && o_in, "null pointer argument");
  SKG_REQUIRE(ld >= enc_dim + ent_dim && ent_tot > 0, "bad node_prep sizes");
  node_prep_kernel<<<total_boxes, 64, 0, stream>>>(ent, ent_dim, ent_tot, box_off, n_h, hum_off, num_images,
                                                   (__nv_bfloat16*)o_in, ld, enc_dim, box2hum);
  return check_launch("node_prep_kernel");
}

extern "C" int skg_fc_bias_fold(const float* bias_in, const float* W_ent, const float* ent, int ent_dim, int entity, int rows,
                                float* bias_out, cudaStream_t stream) {
  if (rows == 0) return SKG_OK;
  SKG_REQUIRE(bias_in && W_ent && ent && bias_out && ent_dim > 0 && entity >= 0, "bad fc_bias_fold arguments");
  fc_bias_fold_kernel<<<ceil_div(rows * 32, 256), 256, 0, stream>>>(bias_in, W_ent, ent, ent_dim, entity, rows, bias_out);
  return check_launch("fc_bias_fold_kernel");
}

extern "C" int skg_global_pool(const float* features, int64_t planes, int hw, void* out_bf16, float* out_f32,
                               cudaStream_t stream) {
  if (planes == 0) return SKG_OK;
  SKG_REQUIRE(features && hw > 0 && (out_bf16 || out_f32), "bad global_pool arguments");
  global_pool_kernel<<<(int)ceil_div64(planes * 32, 256), 256, 0, stream>>>(features, planes, hw, (__nv_bfloat16*)out_bf16, out_f32, nullptr);
  return check_launch("global_pool_kernel");
}

extern "C" int skg_message_aggregate(const float* dot_part, int num_part, float adj_bias, const void* H_os, const void* H_so,
                                     int64_t ld_h, const int32_t* box_off, const int32_t* hum_off, const int64_t* cell_off,
                                     int num_images, int total_boxes, int total_humans, int max_per_image, int width,
                                     void* Z_h, void* Z_o, float* adjacency, cudaStream_t stream) {
  if (total_boxes + total_humans == 0) return SKG_OK;
  SKG_REQUIRE(dot_part && H_os && H_so && box_off && hum_off && cell_off && Z_h && Z_o, "null pointer argument");
  SKG_REQUIRE(width % 8 == 0 && ld_h % 8 == 0 && num_part >= 1, "width and ld_h must be multiples of 8");
  SKG_REQUIRE(max_per_image <= kMaxPerImage, "at most %d boxes per image are supported, got %d", kMaxPerImage, max_per_image);
  aggregate_kernel<<<total_boxes + total_humans, 128, 0, stream>>>(dot_part, num_part, adj_bias, (const __nv_bfloat16*)H_os,
                                                                 (const __nv_bfloat16*)H_so, ld_h, box_off, hum_off, cell_off,
                                                                 num_images, total_humans, width, (__nv_bfloat16*)Z_h,
                                                                 (__nv_bfloat16*)Z_o, adjacency);
  return check_launch("aggregate_kernel");
}

extern "C" int skg_layernorm(const float* x, int64_t rows, int64_t rows_a, int width, const float* w_a, const float* b_a,
                             const float* w_b, const float* b_b, float eps, float* out_f32, void* out_bf16,
                             cudaStream_t stream) {
  if (rows == 0) return SKG_OK;
  SKG_REQUIRE(x && w_a && b_a && (rows_a >= rows || (w_b && b_b)) && (out_f32 || out_bf16), "bad layernorm arguments");
  const float* wb = w_b ? w_b : w_a; const float* bb = b_b ? b_b : b_a;
  const uintptr_t al = (uintptr_t)x | (uintptr_t)w_a | (uintptr_t)b_a | (uintptr_t)wb | (uintptr_t)bb | (uintptr_t)out_f32;
  if (width % 4 == 0 && width <= 4096 && (al & 15u) == 0 && ((uintptr_t)out_bf16 & 7u) == 0 && rows <= 0x7fffffff) {
    if (width <= 1024)
      layernorm_row_kernel<1><<<(int)rows, 256, 0, stream>>>(x, rows_a, width, w_a, b_a, wb, bb, eps, out_f32, (__nv_bfloat16*)out_bf16);
    else
      layernorm_row_kernel<4><<<(int)rows, 256, 0, stream>>>(x, rows_a, width, w_a, b_a, wb, bb, eps, out_f32, (__nv_bfloat16*)out_bf16);
    return check_launch("layernorm_row_kernel");
  }
  layernorm_kernel<<<(int)ceil_div64(rows * 32, 256), 256, 0, stream>>>(x, rows, rows_a, width, w_a, b_a, wb, bb, eps, out_f32,
                                                                       (__nv_bfloat16*)out_bf16);
  return check_launch("layernorm_kernel");
}

extern "C" int skg_hadamard_cells(const void* S_raw, int64_t ld_s, const float* U_h, int64_t ld_uh, const float* U_o,
                                  int64_t ld_uo, const int32_t* hrow, const int32_t* orow, int64_t cells, int width,
                                  void* H, int64_t ld_hh, cudaStream_t stream) {
  if (cells == 0) return SKG_OK;
  SKG_REQUIRE(S_raw && U_h && U_o && hrow && orow && H, "null pointer argument");
  SKG_REQUIRE(width % 8 == 0 && ld_s % 8 == 0 && ld_hh % 8 == 0 && ld_uh % 4 == 0 && ld_uo % 4 == 0, "bad leading dimensions");
  hadamard_cells_kernel<<<grid_for(cells * (width / 8), 256), 256, 0, stream>>>((const __nv_bfloat16*)S_raw, ld_s, U_h, ld_uh, U_o,
                                                                             ld_uo, hrow, orow, cells, width,
                                                                             (__nv_bfloat16*)H, ld_hh);
  return check_launch("hadamard_cells_kernel");
}
