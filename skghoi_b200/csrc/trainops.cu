// Backward-pass kernels of the interaction head that sit between the tensor-core GEMMs (training step, BASELINE
// configs[2]; reference: autograd through heads/adamixer_transH_spatial_r50_head.py:431-530, 884-973 and the losses
// head.py:153-205).  All of them are HBM-bound passes over [cells, R] activations or tiny node-row reductions:
//   transpose_bf16        G[M, N] -> G^T[N, Mp] (+ fused column sums = the bias gradient) for the dW GEMMs
//   head_loss_backward    d(hoi_loss, interactiveness_loss)/d logits straight from the post-processing tables
//   hadamard_backward     backward of the five branch products (fc_1(app) * fc_2(sp)) of the four MBFs, with the
//                         per-node segment sums of the appearance-side gradients
//   aggregate_backward    backward of the softmax-weighted message sums (head.py:907-921)
//   adjacency_backward    backward of adjacency = Linear(R -> 1)(relu(...)) (head.py:897)
//   layernorm_backward    LayerNorm + residual + ReLU gate of the message (head.py:909-925)
//   spatial_fc1_backward  weight gradient of the fp32 first spatial layer (head.py:663)
//   block_copy            parameter packing (per-branch fp32 -> concatenated bf16 operands, + transposed copies)
//                         and gradient unpacking (packed fp32 gradients -> per-branch parameter layout)
#include "common.cuh"
#include "skghoi_b200.h"

namespace skg {

__device__ __forceinline__ uint32_t t_pack2(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ float2 t_unpack2(uint32_t u) {
  __nv_bfloat162 v = *reinterpret_cast<__nv_bfloat162*>(&u);
  return __bfloat1622float2(v);
}
__device__ __forceinline__ void red_add_f32(float* p, float v) { asm volatile("red.global.add.f32 [%0], %1;\n" ::"l"(p), "f"(v) : "memory"); }
__device__ __forceinline__ void red_add_v4(float* p, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};\n" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// ------------------------------------------------------------------------------------------------------------------
// out[c, r] = in[row_map ? row_map[r] : r, c]   (bf16, 64 x 64 tiles through shared memory; rows with a negative map
// entry read as zero); colsum[c] += sum_r in[r, c] (fp32 atomics, optional: the bias gradient rides along for free)
// ------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) transpose_bf16_kernel(const __nv_bfloat16* __restrict__ in, int64_t ld_in, int64_t rows,
                                                             int cols, const int32_t* __restrict__ row_map,
                                                             __nv_bfloat16* __restrict__ out, int64_t ld_out,
                                                             float* __restrict__ colsum) {
  __shared__ __nv_bfloat16 tile[64][66];
  const int64_t r0 = (int64_t)blockIdx.x * 64;
  const int c0 = blockIdx.y * 64;
  // load: 64 rows x 64 cols, each thread two 8-element (16 B) chunks when aligned, scalar otherwise
  const bool vec = ((ld_in & 7) == 0) && ((reinterpret_cast<uintptr_t>(in) & 15u) == 0) && (c0 + 64 <= cols);
  for (int e = threadIdx.x; e < 512; e += 256) {
    const int r = e >> 3, cc = (e & 7) << 3;
    const int64_t gr = r0 + r;
    int64_t src = -1;
    if (gr < rows) src = row_map ? (int64_t)__ldg(row_map + gr) : gr;
    if (src >= 0 && vec) {
      const uint4 u = __ldg(reinterpret_cast<const uint4*>(in + src * ld_in + c0 + cc));
      const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) *reinterpret_cast<uint32_t*>(&tile[r][cc + 2 * k]) = w[k];
    } else {
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const int c = c0 + cc + k;
        tile[r][cc + k] = (src >= 0 && c < cols) ? in[src * ld_in + c] : __float2bfloat16(0.f);
      }
    }
  }
  __syncthreads();
  if (colsum && threadIdx.x < 64 && c0 + (int)threadIdx.x < cols) {
    float s = 0.f;
#pragma unroll 8
    for (int r = 0; r < 64; ++r) s += __bfloat162float(tile[r][threadIdx.x]);
    red_add_f32(colsum + c0 + threadIdx.x, s);
  }
  // store: out row = input column; 64 output rows x 64 output cols (input rows), 8 consecutive input rows per thread
  const bool vout = ((ld_out & 7) == 0) && ((reinterpret_cast<uintptr_t>(out) & 15u) == 0);
  for (int e = threadIdx.x; e < 512; e += 256) {
    const int c = e >> 3, rr = (e & 7) << 3;
    if (c0 + c >= cols) continue;
    __nv_bfloat16* dst = out + (int64_t)(c0 + c) * ld_out + r0 + rr;
    if (vout && r0 + rr + 8 <= ld_out) {      // pad columns up to ld_out are written (zeros from the guarded load)
      uint32_t w[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const uint16_t lo = *reinterpret_cast<const uint16_t*>(&tile[rr + 2 * k][c]);
        const uint16_t hi = *reinterpret_cast<const uint16_t*>(&tile[rr + 2 * k + 1][c]);
        w[k] = (uint32_t)lo | ((uint32_t)hi << 16);
      }
      *reinterpret_cast<uint4*>(dst) = make_uint4(w[0], w[1], w[2], w[3]);
    } else {
#pragma unroll
      for (int k = 0; k < 8; ++k)
        if (r0 + rr + k < ld_out) dst[k] = tile[rr + k][c];
    }
  }
}

// ------------------------------------------------------------------------------------------------------------------
// d(hoi_loss * g_hoi + interactiveness_loss * g_int) / d logits, one warp per valid pair (mirrors postprocess_kernel):
//   score_e = sigmoid(l[p, v]) * (s_h^pw * s_o^pw) * w_p.detach(),  w_p = sigmoid(l[p, num_cls])          head.py:316
//   hoi   = sum_e FL(score_e, label_e; gamma 0.2) / n_p[0]   (head.py:153-177)
//   inter = sum_p FL(w_p, unary_p; gamma 2.0)     / n_p[1]   (head.py:180-205)
// dlogits is a bf16 [rows, ld_d] GEMM operand (zeroed by the caller); the pair's row is logit_row[p] (cell row) or p.
// ------------------------------------------------------------------------------------------------------------------
struct LossTables {
  const int32_t* verb_off; const int32_t* verb_idx; const int32_t* pair2cell; const int32_t* cell_hrow; const int32_t* cell_orow;
  const int32_t* cell_img; const int32_t* hum_off; const int32_t* box_off; const int64_t* labels; const float* scores;
};

__device__ __forceinline__ float focal_dx(float xv, float yv, float alpha, float gamma, float eps) {
  const float a = fabsf(1.f - yv - alpha);
  const float u = fabsf(yv - xv) + eps;
  const float bce = -(yv * fmaxf(logf(xv), -100.f) + (1.f - yv) * fmaxf(logf(1.f - xv), -100.f));
  const float ug = powf(u, gamma);
  const float sgn = (xv > yv) ? 1.f : ((xv < yv) ? -1.f : 0.f);
  const float dbce = (xv - yv) / fmaxf((1.f - xv) * xv, 1e-12f);
  return a * (gamma * powf(u, gamma - 1.f) * sgn * bce + ug * dbce);
}

__global__ void head_loss_backward_kernel(const float* __restrict__ logits, int ld_logits, int num_cls, LossTables t,
                                          const int32_t* __restrict__ logit_row, int64_t pairs, float pw,
                                          const float* __restrict__ gt_labels, const float* __restrict__ n_p,
                                          const float* __restrict__ g_hoi, const float* __restrict__ g_int, float alpha,
                                          float gamma_hoi, float gamma_int, float eps, __nv_bfloat16* __restrict__ dlogits,
                                          int ld_d) {
  const int lane = threadIdx.x & 31;
  const int64_t p = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  if (p >= pairs) return;
  const int cell = __ldg(t.pair2cell + p);
  const int64_t lrow = logit_row ? (int64_t)__ldg(logit_row + p) : p;
  const int hr = __ldg(t.cell_hrow + cell), orow = __ldg(t.cell_orow + cell), b = __ldg(t.cell_img + cell);
  const int boff = __ldg(t.box_off + b);
  const int i = hr - __ldg(t.hum_off + b);
  const float s_h = __ldg(t.scores + boff + i), s_o = __ldg(t.scores + orow);
  const float prior = powf(s_h, pw) * powf(s_o, pw);
  const float* lr = logits + lrow * ld_logits;
  const float w = 1.f / (1.f + expf(-__ldg(lr + num_cls)));
  const float gh = (g_hoi ? __ldg(g_hoi) : 1.f) / __ldg(n_p), gi = (g_int ? __ldg(g_int) : 1.f) / __ldg(n_p + 1);
  __nv_bfloat16* dr = dlogits + lrow * ld_d;
  float unary = 0.f;
  for (int v = lane; v < num_cls; v += 32) unary += gt_labels[p * num_cls + v];
  unary = fminf(warp_sum(unary), 1.f);
  if (lane == 0) dr[num_cls] = __float2bfloat16(gi * focal_dx(w, unary, alpha, gamma_int, eps) * w * (1.f - w));
  if (!(s_h > 0.f)) return;                 // no entries for this pair (nonzero(prior[0]) is empty, head.py:315)
  const int64_t lab_o = __ldg(t.labels + orow);
  const int vo = __ldg(t.verb_off + lab_o), cnt = __ldg(t.verb_off + lab_o + 1) - vo;
  for (int k = lane; k < cnt; k += 32) {
    const int v = __ldg(t.verb_idx + vo + k);
    const float s = 1.f / (1.f + expf(-__ldg(lr + v)));
    const float x = (s * prior) * w;
    dr[v] = __float2bfloat16(gh * focal_dx(x, gt_labels[p * num_cls + v], alpha, gamma_hoi, eps) * (prior * w) * s * (1.f - s));
  }
}

// ------------------------------------------------------------------------------------------------------------------
// Backward of the branch products of the four MBFs (the Hadamard epilogue of the fused fc_2 GEMM and the final
// attention head's product).  Forward, per cell c with human row hr, box row orow, image b (R columns each):
//   pre_att = S_att * (U_h[hr] + U_o[orow])      pre_fin = S_att * (U2h[hr] + U2o[orow])
//   pre_os  = S_os * V_os[orow]                  pre_so  = S_so * V_so[hr]              pre_g = S_g * GT[b]
// Given the gradients dP_* of the five pre-activations (ReLU gates already applied):
//   dS[c] = [dP_att*(U_h+U_o) + dP_fin*(U2h+U2o) | dP_os*V_os | dP_so*V_so | dP_g*GT]            (bf16 [cells, 4R])
//   dhproj[hr]  = [sum_c dP_att*S_att | sum_c dP_so*S_so]     doproj[orow] = [sum_c dP_att*S_att | sum_c dP_os*S_os]
//   dU2h[hr] = sum_c dP_fin*S_att    dU2o[orow] = sum_c dP_fin*S_att    dGT[b] = sum_c dP_g*S_g
// One CTA per (human row, 512-column chunk): the human's n cells are contiguous, its sums stay in registers; the
// box-row and image sums go through fp32 red.global.add (n_h-way contention per address).  The chain rule forces two
// passes: dP_att / dP_os / dP_so depend on dU2h / dU2o through the LayerNorm, the messages and the adjacency.
// ------------------------------------------------------------------------------------------------------------------
struct HadBwd {
  const __nv_bfloat16* S; int64_t ld_s;          // [cells, 4R] raw fc_2 outputs (att | os | so | g)
  const __nv_bfloat16 *dP_att, *dP_fin, *dP_os, *dP_so, *dP_g; int64_t ld_dp;   // [cells, R] each
  const float* hproj; const float* oproj;        // [NH, 2R] (U_h | V_so), [NB, 2R] (U_o | V_os)
  const float* U2h; const float* U2o; const float* GT;   // [NH, R], [NB, R], [B, R]
  __nv_bfloat16* dS; int64_t ld_ds;
  float *dhproj, *doproj, *dU2h, *dU2o, *dGT;
  const int32_t* box_off; const int32_t* hum_off; const int64_t* cell_off;
  int B, R;
};

// PHASE 1 (final pass, before the node side is back-propagated): dS_g, dU2h, dU2o, dGT from dP_fin / dP_g.
// PHASE 2 (after dP_att / dP_os / dP_so are known): dS_att, dS_os, dS_so, dhproj, doproj.
template <int PHASE>
__global__ void __launch_bounds__(128) hadamard_backward_kernel(HadBwd a) {
  const int hr = blockIdx.x;
  const int col = blockIdx.y * 512 + threadIdx.x * 4;
  if (col >= a.R) return;
  const int R = a.R;
  int lo = 0, hi = a.B;
  while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (__ldg(a.hum_off + mid) <= hr) lo = mid; else hi = mid; }
  const int b = lo;
  const int n = __ldg(a.box_off + b + 1) - __ldg(a.box_off + b);
  const int i = hr - __ldg(a.hum_off + b);
  const int64_t c_base = __ldg(a.cell_off + b) + (int64_t)i * n;
  const int o_base = __ldg(a.box_off + b);
  auto ld4 = [](const float* p, float (&v)[4]) { const float4 f = __ldg(reinterpret_cast<const float4*>(p)); v[0] = f.x; v[1] = f.y; v[2] = f.z; v[3] = f.w; };
  auto ldb = [](const __nv_bfloat16* p, float (&v)[4]) {
    const uint2 u = __ldg(reinterpret_cast<const uint2*>(p));
    float2 f = t_unpack2(u.x); v[0] = f.x; v[1] = f.y; f = t_unpack2(u.y); v[2] = f.x; v[3] = f.y;
  };
  auto stb = [](__nv_bfloat16* p, const float (&v)[4]) { *reinterpret_cast<uint2*>(p) = make_uint2(t_pack2(v[0], v[1]), t_pack2(v[2], v[3])); };
  float uh[4], vso[4], u2h[4], gt[4];
  ld4(a.hproj + (int64_t)hr * 2 * R + col, uh); ld4(a.hproj + (int64_t)hr * 2 * R + R + col, vso);
  ld4(a.U2h + (int64_t)hr * R + col, u2h); ld4(a.GT + (int64_t)b * R + col, gt);
  float s0[4] = {0, 0, 0, 0}, s1[4] = {0, 0, 0, 0};      // phase 1: dU2h, dGT   phase 2: dU_h, dV_so
#pragma unroll 2
  for (int j = 0; j < n; ++j) {
    const int64_t c = c_base + j;
    const int orow = o_base + j;
    const __nv_bfloat16* S = a.S + c * a.ld_s + col;
    __nv_bfloat16* d = a.dS + c * a.ld_ds + col;
    float sa[4], pf[4];
    ldb(S, sa); ldb(a.dP_fin + c * a.ld_dp + col, pf);
    if (PHASE == 1) {
      float sg[4], pg[4], t_fin[4], d_g[4];
      ldb(S + 3 * R, sg); ldb(a.dP_g + c * a.ld_dp + col, pg);
#pragma unroll
      for (int k = 0; k < 4; ++k) { t_fin[k] = pf[k] * sa[k]; d_g[k] = pg[k] * gt[k]; s0[k] += t_fin[k]; s1[k] += pg[k] * sg[k]; }
      stb(d + 3 * R, d_g);
      red_add_v4(a.dU2o + (int64_t)orow * R + col, t_fin[0], t_fin[1], t_fin[2], t_fin[3]);
    } else {
      float so_[4], ss[4], pa[4], po[4], ps[4], uo[4], vos[4], u2o[4], d_att[4], d_os[4], d_so[4], t_att[4], t_os[4];
      ldb(S + R, so_); ldb(S + 2 * R, ss);
      ldb(a.dP_att + c * a.ld_dp + col, pa); ldb(a.dP_os + c * a.ld_dp + col, po); ldb(a.dP_so + c * a.ld_dp + col, ps);
      ld4(a.oproj + (int64_t)orow * 2 * R + col, uo); ld4(a.oproj + (int64_t)orow * 2 * R + R + col, vos);
      ld4(a.U2o + (int64_t)orow * R + col, u2o);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        d_att[k] = pa[k] * (uh[k] + uo[k]) + pf[k] * (u2h[k] + u2o[k]);
        d_os[k] = po[k] * vos[k];
        d_so[k] = ps[k] * vso[k];
        t_att[k] = pa[k] * sa[k]; t_os[k] = po[k] * so_[k];
        s0[k] += t_att[k]; s1[k] += ps[k] * ss[k];
      }
      stb(d, d_att); stb(d + R, d_os); stb(d + 2 * R, d_so);
      red_add_v4(a.doproj + (int64_t)orow * 2 * R + col, t_att[0], t_att[1], t_att[2], t_att[3]);
      red_add_v4(a.doproj + (int64_t)orow * 2 * R + R + col, t_os[0], t_os[1], t_os[2], t_os[3]);
    }
  }
  if (PHASE == 1) {
    *reinterpret_cast<float4*>(a.dU2h + (int64_t)hr * R + col) = make_float4(s0[0], s0[1], s0[2], s0[3]);
    red_add_v4(a.dGT + (int64_t)b * R + col, s1[0], s1[1], s1[2], s1[3]);
  } else {
    *reinterpret_cast<float4*>(a.dhproj + (int64_t)hr * 2 * R + col) = make_float4(s0[0], s0[1], s0[2], s0[3]);
    *reinterpret_cast<float4*>(a.dhproj + (int64_t)hr * 2 * R + R + col) = make_float4(s1[0], s1[1], s1[2], s1[3]);
  }
}

// ------------------------------------------------------------------------------------------------------------------
// Backward of the message aggregation (forward: aggregate_kernel in headops.cu, head.py:907-921).  One CTA per node
// row (humans first, then boxes), four warps, one warp per cell of the row at a time:
//   alpha_t = softmax_t(adjacency)                      delta_t = dZ[row] . H[cell_t]
//   dP[cell_t, :] = alpha_t * dZ[row, :] * (H[cell_t, :] > 0)      (gradient of the branch pre-activation, ReLU gated)
//   dadj_part[cell_t] = alpha_t * (delta_t - sum_u alpha_u delta_u)
// human rows read H_os and write dP_os / dadj_h; box rows read H_so and write dP_so / dadj_o.
// ------------------------------------------------------------------------------------------------------------------
constexpr int kAggMax = 1024;
template <int VPL>   // bf16 values per lane = width / 32 (multiple of 8)
__global__ void __launch_bounds__(128) aggregate_backward_kernel(const float* __restrict__ adjacency, const __nv_bfloat16* __restrict__ dZ_h,
                                                                 const __nv_bfloat16* __restrict__ dZ_o, const __nv_bfloat16* __restrict__ H_os,
                                                                 const __nv_bfloat16* __restrict__ H_so, int64_t ld_h,
                                                                 const int32_t* __restrict__ box_off, const int32_t* __restrict__ hum_off,
                                                                 const int64_t* __restrict__ cell_off, int B, int total_humans,
                                                                 __nv_bfloat16* __restrict__ dP_os, __nv_bfloat16* __restrict__ dP_so,
                                                                 int64_t ld_dp, float* __restrict__ dadj_h, float* __restrict__ dadj_o) {
  constexpr int width = VPL * 32;
  __shared__ float sw[kAggMax];      // alpha
  __shared__ float sd[kAggMax];      // delta
  __shared__ float red[4];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x;
  const bool is_h = row < total_humans;
  const int local = is_h ? row : row - total_humans;
  const int32_t* off = is_h ? hum_off : box_off;
  int lo = 0, hi = B;
  while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (__ldg(off + mid) <= local) lo = mid; else hi = mid; }
  const int b = lo;
  const int n = __ldg(box_off + b + 1) - __ldg(box_off + b);
  const int n_h = __ldg(hum_off + b + 1) - __ldg(hum_off + b);
  const int idx = local - __ldg(off + b);
  const int64_t c0 = __ldg(cell_off + b);
  const int cnt = is_h ? n : n_h;
  if (cnt == 0) return;
  const int64_t base = is_h ? c0 + (int64_t)idx * n : c0 + idx;
  const int64_t step = is_h ? 1 : n;
  // softmax weights of the row
  float mx = -INFINITY;
  for (int t = threadIdx.x; t < cnt; t += 128) { const float v = __ldg(adjacency + base + t * step); sw[t] = v; mx = fmaxf(mx, v); }
  mx = warp_max(mx);
  if (lane == 0) red[warp] = mx;
  __syncthreads();
  mx = fmaxf(fmaxf(red[0], red[1]), fmaxf(red[2], red[3]));
  __syncthreads();
  float s = 0.f;
  for (int t = threadIdx.x; t < cnt; t += 128) { const float e = expf(sw[t] - mx); sw[t] = e; s += e; }
  s = warp_sum(s);
  if (lane == 0) red[warp] = s;
  __syncthreads();
  const float inv = 1.f / (red[0] + red[1] + red[2] + red[3]);
  // the row's upstream gradient, VPL values per lane in 8-element chunks (chunk q covers columns (q*32 + lane)*8 ..)
  const __nv_bfloat16* dz = (is_h ? dZ_h : dZ_o) + (int64_t)local * width;
  float g[VPL];
#pragma unroll
  for (int q = 0; q < VPL / 8; ++q) {
    const uint4 u = __ldg(reinterpret_cast<const uint4*>(dz + (q * 32 + lane) * 8));
    const uint32_t w4[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) { const float2 f = t_unpack2(w4[k]); g[q * 8 + 2 * k] = f.x; g[q * 8 + 2 * k + 1] = f.y; }
  }
  const __nv_bfloat16* H = is_h ? H_os : H_so;
  __nv_bfloat16* dP = is_h ? dP_os : dP_so;
  for (int t = warp; t < cnt; t += 4) {
    const int64_t cell = base + t * step;
    const float alpha = sw[t] * inv;
    float dot = 0.f;
#pragma unroll
    for (int q = 0; q < VPL / 8; ++q) {
      const uint4 u = __ldg(reinterpret_cast<const uint4*>(H + cell * ld_h + (q * 32 + lane) * 8));
      const uint32_t w4[4] = {u.x, u.y, u.z, u.w};
      uint32_t o4[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float2 f = t_unpack2(w4[k]);
        const float g0 = g[q * 8 + 2 * k], g1 = g[q * 8 + 2 * k + 1];
        dot = fmaf(g0, f.x, dot); dot = fmaf(g1, f.y, dot);
        o4[k] = t_pack2(f.x > 0.f ? alpha * g0 : 0.f, f.y > 0.f ? alpha * g1 : 0.f);
      }
      *reinterpret_cast<uint4*>(dP + cell * ld_dp + (q * 32 + lane) * 8) = make_uint4(o4[0], o4[1], o4[2], o4[3]);
    }
    dot = warp_sum(dot);
    if (lane == 0) sd[t] = dot;
  }
  __syncthreads();
  float acc = 0.f;
  for (int t = threadIdx.x; t < cnt; t += 128) acc += sw[t] * inv * sd[t];
  acc = warp_sum(acc);
  __syncthreads();
  if (lane == 0) red[warp] = acc;
  __syncthreads();
  const float tot = red[0] + red[1] + red[2] + red[3];
  float* dadj = is_h ? dadj_h : dadj_o;
  for (int t = threadIdx.x; t < cnt; t += 128) dadj[base + t * step] = sw[t] * inv * (sd[t] - tot);
}

// ------------------------------------------------------------------------------------------------------------------
// Backward of adjacency[c] = wts[c, :] . w_adj + b_adj with wts = relu(...) kept as bf16 (head.py:896-897):
//   dwts[c, :] = (dadj_h[c] + dadj_o[c]) * w_adj * (wts[c, :] > 0)   dw_adj += sum_c dadj[c] * wts[c, :]   db_adj += sum dadj
// 64 cells per CTA, 128 threads x 8 columns per pass.
// ------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) adjacency_backward_kernel(const float* __restrict__ dadj_h, const float* __restrict__ dadj_o,
                                                                 const __nv_bfloat16* __restrict__ wts, int64_t ld_w,
                                                                 const float* __restrict__ w_adj, int64_t cells, int width,
                                                                 __nv_bfloat16* __restrict__ dwts, int64_t ld_dw,
                                                                 float* __restrict__ dw_adj, float* __restrict__ db_adj) {
  const int64_t c_begin = (int64_t)blockIdx.x * 64;
  const int64_t c_end = c_begin + 64 < cells ? c_begin + 64 : cells;
  __shared__ float sda[64];
  for (int t = threadIdx.x; t < 64; t += 128) sda[t] = (c_begin + t < cells) ? __ldg(dadj_h + c_begin + t) + __ldg(dadj_o + c_begin + t) : 0.f;
  __syncthreads();
  if (threadIdx.x == 0) {
    float s = 0.f;
    for (int t = 0; t < 64; ++t) s += sda[t];
    red_add_f32(db_adj, s);
  }
  for (int col = threadIdx.x * 8; col < width; col += 128 * 8) {
    float wa[8], acc[8];
    const float4 w0 = __ldg(reinterpret_cast<const float4*>(w_adj + col)), w1 = __ldg(reinterpret_cast<const float4*>(w_adj + col + 4));
    wa[0] = w0.x; wa[1] = w0.y; wa[2] = w0.z; wa[3] = w0.w; wa[4] = w1.x; wa[5] = w1.y; wa[6] = w1.z; wa[7] = w1.w;
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k] = 0.f;
    for (int64_t c = c_begin; c < c_end; ++c) {
      const float da = sda[c - c_begin];
      const uint4 u = __ldg(reinterpret_cast<const uint4*>(wts + c * ld_w + col));
      const uint32_t w4[4] = {u.x, u.y, u.z, u.w};
      uint32_t o4[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float2 f = t_unpack2(w4[k]);
        acc[2 * k] = fmaf(da, f.x, acc[2 * k]); acc[2 * k + 1] = fmaf(da, f.y, acc[2 * k + 1]);
        o4[k] = t_pack2(f.x > 0.f ? da * wa[2 * k] : 0.f, f.y > 0.f ? da * wa[2 * k + 1] : 0.f);
      }
      *reinterpret_cast<uint4*>(dwts + c * ld_dw + col) = make_uint4(o4[0], o4[1], o4[2], o4[3]);
    }
    red_add_v4(dw_adj + col, acc[0], acc[1], acc[2], acc[3]);
    red_add_v4(dw_adj + col + 4, acc[4], acc[5], acc[6], acc[7]);
  }
}

// ------------------------------------------------------------------------------------------------------------------
// Backward of y = LayerNorm(x) * gamma + beta with x = base + relu(msg) (head.py:909-914, 921-925), one CTA per row:
//   dx = rstd * (dy*gamma - mean(dy*gamma) - xhat * mean(dy*gamma*xhat))      dgamma += dy * xhat      dbeta += dy
//   dx_f32 (the residual path's gradient, consumed by the node-row dX GEMMs)    dmsg_bf16 = dx * (x - base > 0)
// ------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) layernorm_backward_kernel(const float* __restrict__ x, const float* __restrict__ dy,
                                                                 const float* __restrict__ base_a, const float* __restrict__ base_b,
                                                                 int64_t rows_a, int width, const float* __restrict__ w_a,
                                                                 const float* __restrict__ w_b, float eps, float* __restrict__ dx,
                                                                 __nv_bfloat16* __restrict__ dmsg, float* __restrict__ dw_a,
                                                                 float* __restrict__ db_a, float* __restrict__ dw_b,
                                                                 float* __restrict__ db_b) {
  __shared__ float red[8];
  __shared__ float bc[2];
  const int64_t row = blockIdx.x;
  const bool is_a = row < rows_a;
  const float* xr = x + row * width;
  const float* dyr = dy + row * width;
  const float* base = is_a ? base_a + row * width : base_b + (row - rows_a) * width;
  const float* w = is_a ? w_a : w_b;
  float* dw = is_a ? dw_a : dw_b;
  float* db = is_a ? db_a : db_b;
  auto block_sum = [&](float v) {
    v = warp_sum(v);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    float t = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) t += red[k];
    return t;
  };
  float s = 0.f;
  for (int c = threadIdx.x; c < width; c += 256) s += xr[c];
  const float mean = block_sum(s) / (float)width;
  float v = 0.f;
  for (int c = threadIdx.x; c < width; c += 256) { const float d = xr[c] - mean; v += d * d; }
  const float rstd = rsqrtf(block_sum(v) / (float)width + eps);
  float s1 = 0.f, s2 = 0.f;
  for (int c = threadIdx.x; c < width; c += 256) {
    const float xh = (xr[c] - mean) * rstd, gy = dyr[c] * __ldg(w + c);
    s1 += gy; s2 += gy * xh;
  }
  const float m1 = block_sum(s1) / (float)width;
  const float m2 = block_sum(s2) / (float)width;
  for (int c = threadIdx.x; c < width; c += 256) {
    const float xv = xr[c], xh = (xv - mean) * rstd, dyv = dyr[c];
    const float g = rstd * (dyv * __ldg(w + c) - m1 - xh * m2);
    dx[row * width + c] = g;
    dmsg[row * width + c] = __float2bfloat16((xv - base[c]) > 0.f ? g : 0.f);
    red_add_f32(dw + c, dyv * xh);
    red_add_f32(db + c, dyv);
  }
  (void)bc;
}

// ------------------------------------------------------------------------------------------------------------------
// Weight / bias gradient of the fp32 first spatial layer (forward: spatial_fc1_kernel, head.py:663-664):
//   dW[o, f] += sum_c g[c, o] * enc[c, f]      db[o] += sum_c g[c, o]       (g = bf16 gradient of the pre-activation)
// 64 cells per CTA staged in shared memory, thread o keeps its IN accumulators in registers, fp32 atomics at the end.
// ------------------------------------------------------------------------------------------------------------------
template <int IN, int OUT>
__global__ void __launch_bounds__(OUT) spatial_fc1_backward_kernel(const float* __restrict__ enc, const __nv_bfloat16* __restrict__ g,
                                                                   int64_t cells, float* __restrict__ dW, float* __restrict__ db) {
  constexpr int ROWS = 64;
  __shared__ float sX[ROWS][IN];
  __shared__ __nv_bfloat16 sG[ROWS][OUT + 2];
  const int o = threadIdx.x;
  float acc[IN];
#pragma unroll
  for (int k = 0; k < IN; ++k) acc[k] = 0.f;
  float accb = 0.f;
  for (int64_t r0 = (int64_t)blockIdx.x * ROWS; r0 < cells; r0 += (int64_t)gridDim.x * ROWS) {
    const int nrow = (int)((cells - r0 < ROWS) ? (cells - r0) : ROWS);
    __syncthreads();
    for (int e = threadIdx.x; e < nrow * IN; e += OUT) sX[e / IN][e % IN] = __ldg(enc + r0 * IN + e);
    for (int e = threadIdx.x; e < nrow * OUT; e += OUT) sG[e / OUT][e % OUT] = g[(r0 + e / OUT) * OUT + e % OUT];
    __syncthreads();
    for (int r = 0; r < nrow; ++r) {
      const float gv = __bfloat162float(sG[r][o]);
      accb += gv;
#pragma unroll
      for (int k = 0; k < IN; ++k) acc[k] = fmaf(gv, sX[r][k], acc[k]);
    }
  }
#pragma unroll
  for (int k = 0; k < IN; ++k) red_add_f32(dW + o * IN + k, acc[k]);
  red_add_f32(db + o, accb);
}

// hum2box[box2hum[k]] = k for the human boxes (inverse of the map node_prep_kernel builds)
// n_p of the focal losses straight from the GT labels: one warp per valid pair
__global__ void loss_normalisers_kernel(const float* __restrict__ gt, LossTables t, int64_t pairs, int num_cls, float* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const int64_t p = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  if (p >= pairs) return;
  const int cell = __ldg(t.pair2cell + p);
  const int hr = __ldg(t.cell_hrow + cell), orow = __ldg(t.cell_orow + cell), b = __ldg(t.cell_img + cell);
  const int boff = __ldg(t.box_off + b);
  const int i = hr - __ldg(t.hum_off + b);
  float s = 0.f;
  for (int v = lane; v < num_cls; v += 32) s += gt[p * num_cls + v];
  s = warp_sum(s);
  float e = 0.f;
  if (__ldg(t.scores + boff + i) > 0.f) {           // the pair's entries exist only for a positive human score (head.py:315)
    const int64_t lab_o = __ldg(t.labels + orow);
    const int v0 = __ldg(t.verb_off + lab_o), v1 = __ldg(t.verb_off + lab_o + 1);
    for (int k = v0 + lane; k < v1; k += 32) e += (gt[p * num_cls + __ldg(t.verb_idx + k)] != 0.f) ? 1.f : 0.f;
  }
  e = warp_sum(e);
  if (lane == 0) {
    if (e != 0.f) atomicAdd(out, e);
    if (s != 0.f) atomicAdd(out + 1, 1.f);
  }
}

// ---- TransH positive / negative selection (head.py:926-963, device form): per image, the positives of the GT label block
// in flat order and as many zero-label entries drawn uniformly without replacement (rejection sampling from a counter-hash
// generator -- uniform over the image's zero labels, the distribution of randperm(#zeros)[:#pos], head.py:939).
__global__ void __launch_bounds__(256) select_count_kernel(const float* __restrict__ gt, const int64_t* __restrict__ pair_off,
                                                           int num_cls, int32_t* __restrict__ n_pos, int32_t* __restrict__ total) {
  __shared__ int red[8];
  const int b = blockIdx.x;
  const int64_t lo = __ldg(pair_off + b) * num_cls, hi = __ldg(pair_off + b + 1) * num_cls;
  int c = 0;
  for (int64_t e = lo + threadIdx.x; e < hi; e += 256) c += (__ldg(gt + e) != 0.f) ? 1 : 0;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = c;
  __syncthreads();
  if (threadIdx.x == 0) {
    int s = 0;
    for (int k = 0; k < 8; ++k) s += red[k];
    n_pos[b] = s;
    atomicAdd(total, s);
  }
}

__device__ __forceinline__ uint64_t splitmix64(uint64_t& x) {
  uint64_t z = (x += 0x9E3779B97F4A7C15ull);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

constexpr int kSelMaxChosen = 2048;
__global__ void __launch_bounds__(256) select_fill_kernel(const float* __restrict__ gt, const int64_t* __restrict__ pair_off, int num_images,
                                                          int num_cls, const int32_t* __restrict__ n_pos, unsigned long long seed,
                                                          int64_t* __restrict__ pos_flat, int64_t* __restrict__ neg_flat) {
  __shared__ int cnt[256];
  __shared__ int64_t chosen[kSelMaxChosen];
  const int b = blockIdx.x, tid = threadIdx.x;
  const int k = __ldg(n_pos + b);
  if (k == 0) return;
  int64_t out0 = 0;
  for (int i = 0; i < b; ++i) out0 += __ldg(n_pos + i);
  const int64_t lo = __ldg(pair_off + b) * num_cls, hi = __ldg(pair_off + b + 1) * num_cls, len = hi - lo;
  // positives in flat order: contiguous chunk per thread, block-wide exclusive scan of the chunk counts
  const int64_t c0 = lo + len * tid / 256, c1 = lo + len * (tid + 1) / 256;
  int c = 0;
  for (int64_t e = c0; e < c1; ++e) c += (__ldg(gt + e) != 0.f) ? 1 : 0;
  cnt[tid] = c;
  __syncthreads();
  if (tid == 0) { int run = 0; for (int i = 0; i < 256; ++i) { const int v = cnt[i]; cnt[i] = run; run += v; } }
  __syncthreads();
  int64_t w = out0 + cnt[tid];
  for (int64_t e = c0; e < c1; ++e) if (__ldg(gt + e) != 0.f) pos_flat[w++] = e;
  // negatives: one thread, k draws, rejecting positives and repeats
  if (tid == 0) {
    uint64_t st = seed ^ (0xD1B54A32D192ED03ull * (uint64_t)(b + 1));
    for (int j = 0; j < k; ++j) {
      int64_t pick = lo;
      for (int tries = 0; tries < 4096; ++tries) {
        const int64_t cand = lo + (int64_t)(splitmix64(st) % (uint64_t)len);
        pick = cand;
        if (__ldg(gt + cand) != 0.f) continue;
        bool dup = false;
        const int seen = j < kSelMaxChosen ? j : kSelMaxChosen;
        for (int q = 0; q < seen; ++q) if (chosen[q] == cand) { dup = true; break; }
        if (!dup) break;
      }
      if (j < kSelMaxChosen) chosen[j] = pick;
      neg_flat[out0 + j] = pick;
    }
  }
}

__global__ void hum2box_kernel(const int32_t* __restrict__ box2hum, int total_boxes, int32_t* __restrict__ hum2box) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= total_boxes) return;
  const int h = __ldg(box2hum + k);
  if (h >= 0) hum2box[h] = k;
}

// ------------------------------------------------------------------------------------------------------------------
// Descriptor-driven 2-D block copy: parameter packing and gradient unpacking in ONE launch each.
// ------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) block_copy_kernel(const skg_block_desc* __restrict__ descs, const int32_t* __restrict__ tiles,
                                                         int num_tiles) {
  __shared__ float tile[32][33];
  const int t = blockIdx.x;
  if (t >= num_tiles) return;
  const skg_block_desc d = descs[__ldg(tiles + 3 * t)];
  const int r0 = __ldg(tiles + 3 * t + 1), c0 = __ldg(tiles + 3 * t + 2);
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;     // 32 x 8
  const float scale = (d.flags & SKG_BLOCK_SCALE) ? d.scale : 1.f;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int r = r0 + ty + 8 * k, c = c0 + tx;
    float v = 0.f;
    if (r < d.rows && c < d.cols) {
      const int64_t si = (int64_t)r * d.ld_src + c;
      v = (d.flags & SKG_BLOCK_SRC_BF16) ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(d.src)[si])
                                         : reinterpret_cast<const float*>(d.src)[si];
      v *= scale;
      const int64_t di = (int64_t)r * d.ld_dst + c;
      if (d.dst) {
        if (d.flags & SKG_BLOCK_DST_BF16) reinterpret_cast<__nv_bfloat16*>(d.dst)[di] = __float2bfloat16(v);
        else if (d.flags & SKG_BLOCK_ACCUMULATE) red_add_f32(reinterpret_cast<float*>(d.dst) + di, v);
        else reinterpret_cast<float*>(d.dst)[di] = v;
      }
    }
    tile[ty + 8 * k][tx] = v;
  }
  if (!d.dst_t) return;                      // uniform per CTA
  __syncthreads();
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int c = c0 + ty + 8 * k, r = r0 + tx;     // transposed copy: dst_t[c, r] (bf16)
    if (r < d.rows && c < d.cols)
      reinterpret_cast<__nv_bfloat16*>(d.dst_t)[(int64_t)c * d.ld_dst_t + r] = __float2bfloat16(tile[tx][ty + 8 * k]);
  }
}

}  // namespace skg

using namespace skg;

extern "C" int skg_transpose_bf16(const void* in, int64_t ld_in, int64_t rows, int cols, const int32_t* row_map, void* out,
                                  int64_t ld_out, float* colsum, cudaStream_t stream) {
  if (rows == 0 || cols == 0) return SKG_OK;
  SKG_REQUIRE(in && out && ld_in >= cols && ld_out >= rows, "bad transpose arguments (rows %lld cols %d ld_in %lld ld_out %lld)",
              (long long)rows, cols, (long long)ld_in, (long long)ld_out);
  dim3 grid((unsigned)ceil_div64(rows, 64), (unsigned)ceil_div(cols, 64));
  SKG_REQUIRE(grid.y <= 65535, "too many columns for one transpose launch");
  transpose_bf16_kernel<<<grid, 256, 0, stream>>>((const __nv_bfloat16*)in, ld_in, rows, cols, row_map, (__nv_bfloat16*)out, ld_out, colsum);
  return check_launch("transpose_bf16_kernel");
}

extern "C" int skg_head_loss_backward(const float* logits, int ld_logits, int num_cls, const skg_post_tables* tables,
                                      const int32_t* logit_row, int64_t pairs, float prior_power, const float* gt_labels,
                                      const float* n_p, const float* g_hoi, const float* g_int, void* dlogits_bf16, int ld_d,
                                      cudaStream_t stream) {
  if (pairs == 0) return SKG_OK;
  SKG_REQUIRE(logits && tables && gt_labels && n_p && dlogits_bf16 && ld_d > num_cls, "bad loss_backward arguments");
  LossTables t{tables->verb_off, tables->verb_idx, tables->pair2cell, tables->cell_hrow, tables->cell_orow, tables->cell_img,
               tables->hum_off, tables->box_off, tables->labels, tables->scores};
  SKG_REQUIRE(t.verb_off && t.verb_idx && t.pair2cell && t.cell_hrow && t.cell_orow && t.cell_img && t.hum_off && t.box_off && t.labels && t.scores,
              "incomplete post tables");
  head_loss_backward_kernel<<<(int)ceil_div64(pairs * 32, 256), 256, 0, stream>>>(logits, ld_logits, num_cls, t, logit_row, pairs,
                                                                                  prior_power, gt_labels, n_p, g_hoi, g_int, 0.5f, 0.2f,
                                                                                  2.0f, 1e-6f, (__nv_bfloat16*)dlogits_bf16, ld_d);
  return check_launch("head_loss_backward_kernel");
}

extern "C" int skg_loss_normalisers(const float* gt_labels, const skg_post_tables* tables, int64_t pairs, int num_cls, float* out,
                                    cudaStream_t stream) {
  SKG_REQUIRE(out, "null output pointer");
  SKG_CUDA(cudaMemsetAsync(out, 0, 2 * sizeof(float), stream));
  if (pairs == 0) return SKG_OK;
  SKG_REQUIRE(gt_labels && tables && num_cls > 0, "bad loss_normalisers arguments");
  LossTables t{tables->verb_off, tables->verb_idx, tables->pair2cell, tables->cell_hrow, tables->cell_orow, tables->cell_img,
               tables->hum_off, tables->box_off, tables->labels, tables->scores};
  SKG_REQUIRE(t.verb_off && t.verb_idx && t.pair2cell && t.cell_hrow && t.cell_orow && t.cell_img && t.hum_off && t.box_off && t.labels && t.scores,
              "incomplete post tables");
  loss_normalisers_kernel<<<(int)ceil_div64(pairs * 32, 256), 256, 0, stream>>>(gt_labels, t, pairs, num_cls, out);
  return check_launch("loss_normalisers_kernel");
}

extern "C" int skg_select_count(const float* gt_labels, const int64_t* pair_off, int num_images, int num_cls, int32_t* n_pos,
                                cudaStream_t stream) {
  SKG_REQUIRE(n_pos && num_images >= 0, "bad select_count arguments");
  SKG_CUDA(cudaMemsetAsync(n_pos, 0, sizeof(int32_t) * (num_images + 1), stream));
  if (num_images == 0) return SKG_OK;
  SKG_REQUIRE(gt_labels && pair_off && num_cls > 0, "null pointer argument");
  select_count_kernel<<<num_images, 256, 0, stream>>>(gt_labels, pair_off, num_cls, n_pos, n_pos + num_images);
  return check_launch("select_count_kernel");
}

extern "C" int skg_select_fill(const float* gt_labels, const int64_t* pair_off, int num_images, int num_cls, const int32_t* n_pos,
                               unsigned long long seed, int64_t* pos_flat, int64_t* neg_flat, cudaStream_t stream) {
  if (num_images == 0) return SKG_OK;
  SKG_REQUIRE(gt_labels && pair_off && n_pos && pos_flat && neg_flat && num_cls > 0, "null pointer argument");
  select_fill_kernel<<<num_images, 256, 0, stream>>>(gt_labels, pair_off, num_images, num_cls, n_pos, seed, pos_flat, neg_flat);
  return check_launch("select_fill_kernel");
}

extern "C" int skg_hadamard_backward(const skg_hadamard_bwd* a, int total_humans, int phase, cudaStream_t stream) {
  SKG_REQUIRE(a && (phase == 1 || phase == 2), "hadamard_backward: bad arguments (phase must be 1 or 2)");
  if (total_humans == 0) return SKG_OK;
  SKG_REQUIRE(a->R % 4 == 0 && a->ld_s % 4 == 0 && a->ld_dp % 4 == 0 && a->ld_ds % 4 == 0, "hadamard_backward: widths must be multiples of 4");
  HadBwd k;
  k.S = (const __nv_bfloat16*)a->S; k.ld_s = a->ld_s;
  k.dP_att = (const __nv_bfloat16*)a->dP_att; k.dP_fin = (const __nv_bfloat16*)a->dP_fin; k.dP_os = (const __nv_bfloat16*)a->dP_os;
  k.dP_so = (const __nv_bfloat16*)a->dP_so; k.dP_g = (const __nv_bfloat16*)a->dP_g; k.ld_dp = a->ld_dp;
  k.hproj = a->hproj; k.oproj = a->oproj; k.U2h = a->U2h; k.U2o = a->U2o; k.GT = a->GT;
  k.dS = (__nv_bfloat16*)a->dS; k.ld_ds = a->ld_ds;
  k.dhproj = a->dhproj; k.doproj = a->doproj; k.dU2h = a->dU2h; k.dU2o = a->dU2o; k.dGT = a->dGT;
  k.box_off = a->box_off; k.hum_off = a->hum_off; k.cell_off = a->cell_off; k.B = a->num_images; k.R = a->R;
  SKG_REQUIRE(k.S && k.dP_att && k.dP_fin && k.dP_os && k.dP_so && k.dP_g && k.hproj && k.oproj && k.U2h && k.U2o && k.GT && k.dS && k.dhproj &&
                  k.doproj && k.dU2h && k.dU2o && k.dGT && k.box_off && k.hum_off && k.cell_off,
              "hadamard_backward: null pointer");
  dim3 grid((unsigned)total_humans, (unsigned)ceil_div(a->R, 512));
  if (phase == 1) hadamard_backward_kernel<1><<<grid, 128, 0, stream>>>(k);
  else hadamard_backward_kernel<2><<<grid, 128, 0, stream>>>(k);
  return check_launch("hadamard_backward_kernel");
}

extern "C" int skg_message_aggregate_backward(const float* adjacency, const void* dZ_h, const void* dZ_o, const void* H_os,
                                              const void* H_so, int64_t ld_h, const int32_t* box_off, const int32_t* hum_off,
                                              const int64_t* cell_off, int num_images, int total_boxes, int total_humans,
                                              int max_per_image, int width, void* dP_os, void* dP_so, int64_t ld_dp,
                                              float* dadj_h, float* dadj_o, cudaStream_t stream) {
  if (total_boxes + total_humans == 0) return SKG_OK;
  SKG_REQUIRE(adjacency && dZ_h && dZ_o && H_os && H_so && box_off && hum_off && cell_off && dP_os && dP_so && dadj_h && dadj_o,
              "null pointer argument");
  SKG_REQUIRE(max_per_image <= kAggMax, "at most %d boxes per image are supported, got %d", kAggMax, max_per_image);
  SKG_REQUIRE(ld_h % 8 == 0 && ld_dp % 8 == 0, "leading dimensions must be multiples of 8");
  const int rows = total_boxes + total_humans;
#define SKG_AGG_BWD(V)                                                                                                     \
  aggregate_backward_kernel<V><<<rows, 128, 0, stream>>>(adjacency, (const __nv_bfloat16*)dZ_h, (const __nv_bfloat16*)dZ_o, \
                                                        (const __nv_bfloat16*)H_os, (const __nv_bfloat16*)H_so, ld_h, box_off, hum_off, \
                                                        cell_off, num_images, total_humans, (__nv_bfloat16*)dP_os,         \
                                                        (__nv_bfloat16*)dP_so, ld_dp, dadj_h, dadj_o)
  if (width == 1024) SKG_AGG_BWD(32);
  else if (width == 512) SKG_AGG_BWD(16);
  else if (width == 256) SKG_AGG_BWD(8);
  else if (width == 2048) SKG_AGG_BWD(64);
  else { set_error("message_aggregate_backward supports widths 256 / 512 / 1024 / 2048, got %d", width); return SKG_ERR_UNSUPPORTED; }
#undef SKG_AGG_BWD
  return check_launch("aggregate_backward_kernel");
}

extern "C" int skg_adjacency_backward(const float* dadj_h, const float* dadj_o, const void* wts, int64_t ld_w, const float* w_adj,
                                      int64_t cells, int width, void* dwts, int64_t ld_dw, float* dw_adj, float* db_adj,
                                      cudaStream_t stream) {
  if (cells == 0) return SKG_OK;
  SKG_REQUIRE(dadj_h && dadj_o && wts && w_adj && dwts && dw_adj && db_adj, "null pointer argument");
  SKG_REQUIRE(width % 8 == 0 && ld_w % 8 == 0 && ld_dw % 8 == 0, "width and leading dimensions must be multiples of 8");
  adjacency_backward_kernel<<<(int)ceil_div64(cells, 64), 128, 0, stream>>>(dadj_h, dadj_o, (const __nv_bfloat16*)wts, ld_w, w_adj, cells,
                                                                           width, (__nv_bfloat16*)dwts, ld_dw, dw_adj, db_adj);
  return check_launch("adjacency_backward_kernel");
}

extern "C" int skg_layernorm_backward(const float* x, const float* dy, const float* base_a, const float* base_b, int64_t rows,
                                      int64_t rows_a, int width, const float* w_a, const float* w_b, float eps, float* dx,
                                      void* dmsg_bf16, float* dw_a, float* db_a, float* dw_b, float* db_b, cudaStream_t stream) {
  if (rows == 0) return SKG_OK;
  SKG_REQUIRE(x && dy && base_a && w_a && dx && dmsg_bf16 && dw_a && db_a && (rows_a >= rows || (base_b && w_b && dw_b && db_b)),
              "bad layernorm_backward arguments");
  layernorm_backward_kernel<<<(int)rows, 256, 0, stream>>>(x, dy, base_a, base_b, rows_a, width, w_a, w_b ? w_b : w_a, eps, dx,
                                                          (__nv_bfloat16*)dmsg_bf16, dw_a, db_a, dw_b ? dw_b : dw_a, db_b ? db_b : db_a);
  return check_launch("layernorm_backward_kernel");
}

extern "C" int skg_spatial_fc1_backward(const float* enc, const void* g_bf16, int64_t cells, int in_features, int out_features,
                                        float* dW, float* db, cudaStream_t stream) {
  if (cells == 0) return SKG_OK;
  SKG_REQUIRE(enc && g_bf16 && dW && db, "null pointer argument");
  SKG_REQUIRE(in_features == 46 && out_features == 128, "spatial_fc1_backward supports 46 -> 128 (head.py:663), got %d -> %d", in_features,
              out_features);
  const int64_t chunks = ceil_div64(cells, 64);
  const int grid = (int)(chunks < 2 * num_sms() ? chunks : 2 * num_sms());
  spatial_fc1_backward_kernel<46, 128><<<grid, 128, 0, stream>>>(enc, (const __nv_bfloat16*)g_bf16, cells, dW, db);
  return check_launch("spatial_fc1_backward_kernel");
}

extern "C" int skg_hum2box(const int32_t* box2hum, int total_boxes, int32_t* hum2box, cudaStream_t stream) {
  if (total_boxes == 0) return SKG_OK;
  SKG_REQUIRE(box2hum && hum2box, "null pointer argument");
  hum2box_kernel<<<ceil_div(total_boxes, 256), 256, 0, stream>>>(box2hum, total_boxes, hum2box);
  return check_launch("hum2box_kernel");
}

extern "C" int skg_block_copy(const skg_block_desc* descs_dev, const int32_t* tiles_dev, int num_tiles, cudaStream_t stream) {
  if (num_tiles == 0) return SKG_OK;
  SKG_REQUIRE(descs_dev && tiles_dev && num_tiles > 0, "bad block_copy arguments");
  block_copy_kernel<<<num_tiles, 256, 0, stream>>>(descs_dev, tiles_dev, num_tiles);
  return check_launch("block_copy_kernel");
}
