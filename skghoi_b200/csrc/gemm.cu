// bf16 GEMM with fused epilogues on the 5th-generation tensor cores (tcgen05 + TMEM + TMA).
//
//   C[M, N] = epilogue( A[M, K] (bf16, K-major)  x  B[N, K]^T (bf16, K-major, nn.Linear layout) )
//
// This one kernel carries every dense contraction of the interaction head (reference
// head.py:431-530 MultiBranchFusion / MessageMBF, head.py:635-701 box_head / spatial_head /
// fc_head / fc_tail, models.py:176-177 predictor / suppressor).  Persistent, warp-specialised:
//   warp 0      : TMA producer  (cp.async.bulk.tensor 2D, 128B swizzle, mbarrier complete_tx)
//   warp 1      : MMA issuer    (one elected lane issues tcgen05.mma M=128 x N=BN x K=16, fp32 accum in TMEM)
//   warp 2      : TMEM allocator (2 accumulator buffers so the epilogue of tile i overlaps the MMAs of tile i+1)
//   warps 4..7  : epilogue      (tcgen05.ld 32x32b -> registers -> fused math -> global)
// The epilogue is described per column group (skg_epilogue_group): bias, Hadamard product with
// row-gathered fp32 tables (the MBF "fc_1(app) * fc_2(sp)" branch product, gathered through the
// cell -> human / box row maps so the [pairs, 1024] appearance term is never materialised), ReLU,
// residual add, a fused row-dot (the adjacency Linear(1024 -> 1)), row scatter (valid-pair
// compaction) and bf16 / fp32 stores.
#include "common.cuh"
#include "skghoi_b200.h"
#include <cuda.h>

namespace skg {

constexpr int BM = 128;            // UMMA M (cta_group::1)
constexpr int BK = 64;             // 64 bf16 = 128 B = one swizzle-128B atom row
constexpr int UMMA_K = 16;
constexpr int kGemmThreads = 256;  // 8 warps: 0 TMA, 1 MMA, 2 TMEM alloc, 3 idle, 4..7 epilogue
constexpr int kEpiWarp0 = 4;
constexpr int kMaxGroups = 4;

struct EpiGroups {
  skg_epilogue_group g[kMaxGroups];
  int num;
};

// ---------------------------------------------------------------------------------------------
// PTX wrappers
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];\n" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE;\n"
      "bra WAIT_LOOP;\n"
      "DONE:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void tma_load_2d(const CUtensorMap* map, uint64_t* bar, void* smem_dst, int c_inner, int c_outer) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];\n" ::"r"(
          smem_u32(smem_dst)),
      "l"(map), "r"(smem_u32(bar)), "r"(c_inner), "r"(c_outer)
      : "memory");
}
__device__ __forceinline__ void tcgen05_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory"); }
__device__ __forceinline__ void tcgen05_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory"); }
__device__ __forceinline__ void tcgen05_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory");
}

// K-major, 128B-swizzled operand tile: rows of 64 bf16 (128 B), 8-row groups 1024 B apart.
// Descriptor fields (see cute/arch/mma_sm100_desc.hpp): start>>4 [0,14), LBO>>4 [16,30) (=1, unused for
// swizzled K-major), SBO>>4 [32,46) (=64), version=1 [46,48), layout_type=2 (SWIZZLE_128B) [61,64).
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
// instruction descriptor: c=F32 [4,6), a=BF16 [7,10), b=BF16 [10,13), K-major both, N>>3 [17,23), M>>4 [24,29)
__host__ __device__ constexpr uint32_t make_idesc(int m, int n) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}

__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}

// ---------------------------------------------------------------------------------------------
// kernel
// ---------------------------------------------------------------------------------------------
template <int BN>
struct GemmSmem {
  static constexpr int kStages = (BN == 256) ? 4 : 6;
  static constexpr int kABytes = BM * BK * 2;
  static constexpr int kBBytes = BN * BK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kTotal = kStages * kStageBytes + 1024 /*align slack*/ + 256 /*barriers*/;
};

template <int BN>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, int M, int N, int K,
                 const __grid_constant__ EpiGroups epi) {
  using S = GemmSmem<BN>;
  constexpr int kStages = S::kStages;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kStages * S::kStageBytes);
  uint64_t* full_bar = bars;                     // [kStages]  TMA -> MMA
  uint64_t* empty_bar = bars + kStages;          // [kStages]  MMA -> TMA
  uint64_t* tmem_full = bars + 2 * kStages;      // [2]        MMA -> epilogue
  uint64_t* tmem_empty = bars + 2 * kStages + 2; // [2]        epilogue -> MMA
  uint32_t* tmem_base_slot = reinterpret_cast<uint32_t*>(bars + 2 * kStages + 4);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int num_m = (M + BM - 1) / BM, num_n = (N + BN - 1) / BN;
  const int num_tiles = num_m * num_n;
  const int num_k = (K + BK - 1) / BK;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];\n" ::"l"(&tmA) : "memory");
    asm volatile("prefetch.tensormap [%0];\n" ::"l"(&tmB) : "memory");
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kStages; ++s) { mbar_init(full_bar + s, 1); mbar_init(empty_bar + s, 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(tmem_full + a, 1); mbar_init(tmem_empty + a, 4); }
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  if (warp == 2) {   // 2 x BN fp32 accumulator columns (power of two >= 32)
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(smem_u32(tmem_base_slot)), "r"(2 * BN));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::: "memory");
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_base_slot;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int m0 = (tile / num_n) * BM, n0 = (tile % num_n) * BN;
        for (int kb = 0; kb < num_k; ++kb) {
          mbar_wait(empty_bar + stage, phase ^ 1);
          uint8_t* sa = smem + stage * S::kStageBytes;
          uint8_t* sb = sa + S::kABytes;
          mbar_arrive_expect_tx(full_bar + stage, S::kStageBytes);
          tma_load_2d(&tmA, full_bar + stage, sa, kb * BK, m0);
          tma_load_2d(&tmB, full_bar + stage, sb, kb * BK, n0);
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc(BM, BN);
      int stage = 0; uint32_t phase = 0;
      int acc = 0; uint32_t acc_phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        mbar_wait(tmem_empty + acc, acc_phase ^ 1);      // epilogue has drained this accumulator
        tcgen05_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(acc * BN);
        for (int kb = 0; kb < num_k; ++kb) {
          mbar_wait(full_bar + stage, phase);
          tcgen05_fence_after();
          const uint32_t sa = smem_u32(smem + stage * S::kStageBytes);
          const uint32_t sb = sa + S::kABytes;
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            const uint64_t da = make_smem_desc(sa + k * UMMA_K * 2);
            const uint64_t db = make_smem_desc(sb + k * UMMA_K * 2);
            umma_bf16(d_tmem, da, db, idesc, (kb | k) != 0 ? 1u : 0u);
          }
          tcgen05_commit(empty_bar + stage);              // smem slot free once these MMAs retire
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
        tcgen05_commit(tmem_full + acc);                  // accumulator complete -> epilogue
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else if (warp >= kEpiWarp0) {
    // ===================== epilogue =====================
    const int quad = warp & 3;                            // TMEM lane quadrant this warp may access
    int acc = 0; uint32_t acc_phase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      const int m_blk = tile / num_n, n_blk = tile % num_n;
      const int m0 = m_blk * BM, n0 = n_blk * BN;
      int gi = 0;
      for (int q = 1; q < epi.num; ++q) if (n0 >= epi.g[q].n_begin) gi = q;
      const skg_epilogue_group& G = epi.g[gi];
      const int row = m0 + quad * 32 + lane;
      const bool row_ok = row < M;
      const int flags = G.flags;
      int rp = row, rq = row, rr = row, ro = row;
      if (row_ok) {
        if (G.row_p) rp = __ldg(G.row_p + row);
        if (G.row_q) rq = __ldg(G.row_q + row);
        if (G.row_r) rr = __ldg(G.row_r + row);
        if (G.out_row) ro = __ldg(G.out_row + row);
      }
      const bool store_ok = row_ok && ro >= 0;
      float dot = 0.f;
      mbar_wait(tmem_full + acc, acc_phase);
      tcgen05_fence_after();
      const uint32_t t_row = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(acc * BN);
#pragma unroll 1
      for (int c0 = 0; c0 < BN; c0 += 32) {
        uint32_t raw[32];
        tmem_ld32(t_row + c0, raw);
        const int col = n0 + c0;                 // global column of raw[0]
        if (col >= N) break;                     // warp-uniform
        const int lc = col - G.n_begin;          // column inside the group (table / output column)
        float v[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(raw[j]);
        if (G.bias) {
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            const float4 b = __ldg(reinterpret_cast<const float4*>(G.bias + col + j));
            v[j] += b.x; v[j + 1] += b.y; v[j + 2] += b.z; v[j + 3] += b.w;
          }
        }
        if (store_ok && G.raw_bf16) {
          uint4* dst = reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(G.raw_bf16) + (int64_t)ro * G.ld_raw + lc);
#pragma unroll
          for (int j = 0; j < 32; j += 8)
            dst[j >> 3] = make_uint4(pack_bf16(v[j], v[j + 1]), pack_bf16(v[j + 2], v[j + 3]), pack_bf16(v[j + 4], v[j + 5]),
                                     pack_bf16(v[j + 6], v[j + 7]));
        }
        if (row_ok && G.mul_p) {
          const float4* p = reinterpret_cast<const float4*>(G.mul_p + (int64_t)rp * G.ld_p + lc);
          if (G.mul_q) {
            const float4* q = reinterpret_cast<const float4*>(G.mul_q + (int64_t)rq * G.ld_q + lc);
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
              const float4 a = __ldg(p + (j >> 2)), b = __ldg(q + (j >> 2));
              v[j] *= (a.x + b.x); v[j + 1] *= (a.y + b.y); v[j + 2] *= (a.z + b.z); v[j + 3] *= (a.w + b.w);
            }
          } else {
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
              const float4 a = __ldg(p + (j >> 2));
              v[j] *= a.x; v[j + 1] *= a.y; v[j + 2] *= a.z; v[j + 3] *= a.w;
            }
          }
        }
        if (flags & SKG_EPI_RELU) {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], 0.f);
        }
        if (row_ok && G.residual) {
          const float4* r4 = reinterpret_cast<const float4*>(G.residual + (int64_t)rr * G.ld_r + lc);
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            const float4 a = __ldg(r4 + (j >> 2));
            v[j] += a.x; v[j + 1] += a.y; v[j + 2] += a.z; v[j + 3] += a.w;
          }
        }
        if (G.dot_w) {
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            const float4 a = __ldg(reinterpret_cast<const float4*>(G.dot_w + lc + j));
            dot += v[j] * a.x + v[j + 1] * a.y + v[j + 2] * a.z + v[j + 3] * a.w;
          }
        }
        if (store_ok && G.out_bf16) {
          uint4* dst = reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(G.out_bf16) + (int64_t)ro * G.ld_ob + lc);
#pragma unroll
          for (int j = 0; j < 32; j += 8)
            dst[j >> 3] = make_uint4(pack_bf16(v[j], v[j + 1]), pack_bf16(v[j + 2], v[j + 3]), pack_bf16(v[j + 4], v[j + 5]),
                                     pack_bf16(v[j + 6], v[j + 7]));
        }
        if (store_ok && G.lo_bf16) {   // residual of the bf16 rounding: v ~= hi + lo to ~16 mantissa bits
          uint4* dst = reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(G.lo_bf16) + (int64_t)ro * G.ld_lo + lc);
#pragma unroll
          for (int j = 0; j < 32; j += 8) {
            float l[8];
#pragma unroll
            for (int t = 0; t < 8; ++t) l[t] = v[j + t] - __bfloat162float(__float2bfloat16(v[j + t]));
            dst[j >> 3] = make_uint4(pack_bf16(l[0], l[1]), pack_bf16(l[2], l[3]), pack_bf16(l[4], l[5]), pack_bf16(l[6], l[7]));
          }
        }
        if (store_ok && G.out_f32) {
          float* dst = G.out_f32 + (int64_t)ro * G.ld_of + lc;
          if (col + 32 <= N && (G.ld_of & 3) == 0) {
#pragma unroll
            for (int j = 0; j < 32; j += 4) *reinterpret_cast<float4*>(dst + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) if (col + j < N) dst[j] = v[j];
          }
        }
      }
      if (G.dot_out && store_ok) G.dot_out[(int64_t)ro * G.ld_dot + (n0 - G.n_begin) / BN] = dot;
      // accumulator drained: hand the TMEM buffer back to the MMA warp
      tcgen05_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tmem_empty + acc);
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  }

  tcgen05_fence_before();
  __syncthreads();
  if (warp == 2) {
    tcgen05_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" ::"r"(tmem_base), "r"(2 * BN));
  }
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                    const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode_fn() {
  static PFN_encodeTiled fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess)
      return nullptr;
    fn = reinterpret_cast<PFN_encodeTiled>(p);
  }
  return fn;
}

// 2-D bf16 row-major [rows, cols] (cols contiguous, row pitch ld elements), box = [box_rows, 64 cols], 128B swizzle
static int make_tmap(CUtensorMap* map, const void* ptr, int64_t rows, int64_t cols, int64_t ld, int box_rows) {
  PFN_encodeTiled enc = get_encode_fn();
  if (!enc) { set_error("cuTensorMapEncodeTiled entry point not available"); return SKG_ERR_CUDA; }
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
  cuuint32_t box[2] = {(cuuint32_t)BK, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled failed (%d) rows=%lld cols=%lld ld=%lld", (int)r, (long long)rows, (long long)cols, (long long)ld); return SKG_ERR_CUDA; }
  return SKG_OK;
}

template <int BN>
static int launch_gemm(const CUtensorMap& tmA, const CUtensorMap& tmB, int M, int N, int K, const EpiGroups& epi,
                       int max_ctas, cudaStream_t stream) {
  using S = GemmSmem<BN>;
  auto kern = gemm_bf16_kernel<BN>;
  static bool attr_set = false;
  if (!attr_set) {
    SKG_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, S::kTotal));
    attr_set = true;
  }
  const int tiles = ceil_div(M, BM) * ceil_div(N, BN);
  int grid = tiles < num_sms() ? tiles : num_sms();
  if (max_ctas > 0 && grid > max_ctas) grid = max_ctas;
  kern<<<grid, kGemmThreads, S::kTotal, stream>>>(tmA, tmB, M, N, K, epi);
  return check_launch("gemm_bf16_kernel");
}

}  // namespace skg

using namespace skg;

extern "C" int skg_gemm_bf16(const void* A, int64_t lda, const void* B, int64_t ldb, int M, int N, int K,
                             const skg_epilogue_group* groups, int num_groups, int block_n, cudaStream_t stream) {
  SKG_REQUIRE(M >= 0 && N > 0 && K > 0, "bad GEMM shape M=%d N=%d K=%d", M, N, K);
  if (M == 0) return SKG_OK;
  SKG_REQUIRE(A && B && groups, "null pointer argument");
  SKG_REQUIRE(num_groups >= 1 && num_groups <= kMaxGroups, "num_groups must be in [1, %d]", kMaxGroups);
  SKG_REQUIRE((lda % 8) == 0 && (ldb % 8) == 0 && lda >= K && ldb >= K, "lda/ldb must be >= K and multiples of 8 elements");
  SKG_REQUIRE(((reinterpret_cast<uintptr_t>(A) | reinterpret_cast<uintptr_t>(B)) & 15u) == 0, "A and B must be 16-byte aligned");
  if (block_n == 0) block_n = (N % 256 == 0) ? 256 : 128;
  SKG_REQUIRE(block_n == 128 || block_n == 256, "block_n must be 128 or 256");
  EpiGroups epi;
  epi.num = num_groups;
  for (int i = 0; i < num_groups; ++i) {
    const skg_epilogue_group& g = groups[i];
    SKG_REQUIRE(g.n_begin % block_n == 0 && g.n_begin < N, "group %d: n_begin must be a multiple of block_n inside N", i);
    SKG_REQUIRE(i == 0 ? g.n_begin == 0 : g.n_begin > groups[i - 1].n_begin, "groups must be sorted by n_begin, first at 0");
    SKG_REQUIRE(!g.mul_q || g.mul_p, "group %d: mul_q requires mul_p", i);
    SKG_REQUIRE(!g.out_bf16 || g.ld_ob % 8 == 0, "group %d: ld_ob must be a multiple of 8", i);
    SKG_REQUIRE(!g.raw_bf16 || g.ld_raw % 8 == 0, "group %d: ld_raw must be a multiple of 8", i);
    SKG_REQUIRE(!g.lo_bf16 || g.ld_lo % 8 == 0, "group %d: ld_lo must be a multiple of 8", i);
    SKG_REQUIRE(!g.mul_p || g.ld_p % 4 == 0, "group %d: ld_p must be a multiple of 4", i);
    SKG_REQUIRE(!g.mul_q || g.ld_q % 4 == 0, "group %d: ld_q must be a multiple of 4", i);
    SKG_REQUIRE(!g.residual || g.ld_r % 4 == 0, "group %d: ld_r must be a multiple of 4", i);
    epi.g[i] = g;
  }
  // bias / tables are read in 32-column chunks up to the padded width: callers pad them to block_n multiples
  CUtensorMap tmA, tmB;
  int rc = make_tmap(&tmA, A, M, K, lda, BM);
  if (rc) return rc;
  rc = make_tmap(&tmB, B, N, K, ldb, block_n);
  if (rc) return rc;
  if (block_n == 256) return launch_gemm<256>(tmA, tmB, M, N, K, epi, 0, stream);
  return launch_gemm<128>(tmA, tmB, M, N, K, epi, 0, stream);
}
