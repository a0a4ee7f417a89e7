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
//   warps 4..11 : epilogue      (tcgen05.ld 32x32b -> registers -> smem transpose -> fused math -> coalesced global);
//                 the gather-epilogue variants move registers from warps 0..3 to these warps (setmaxnreg 40 / 232)
// The epilogue is described per column group (skg_epilogue_group): bias, Hadamard product with
// row-gathered fp32 tables (the MBF "fc_1(app) * fc_2(sp)" branch product, gathered through the
// cell -> human / box row maps so the [pairs, 1024] appearance term is never materialised), ReLU,
// residual add, a fused row-dot (the adjacency Linear(1024 -> 1)), row scatter (valid-pair
// compaction) and bf16 / fp32 stores.
#include "common.cuh"
#include "skghoi_b200.h"
#include <cuda.h>
#include <stdlib.h>

namespace skg {

constexpr int BM = 128;            // UMMA M (cta_group::1)
constexpr int BK = 64;             // 64 bf16 = 128 B = one swizzle-128B atom row
constexpr int UMMA_K = 16;
#ifndef SKG_EPI_WARPS
#define SKG_EPI_WARPS 8
#endif
constexpr int kEpiWarps = SKG_EPI_WARPS;   // kEpiParts per TMEM lane quadrant (8 or 16)
// gathered operands software-pipelined one block ahead (costs ~40 registers).  Measured on B200 (tools/gemm_bench.py --epi,
// fused fc_2 shape 4096 x 4096 x 1024, flushed single launches): 8 warps + prefetch 49.9 us, 16 warps without it (the
// register budget of 640 threads leaves no room) 58.2 us -- latency hiding by prefetch beats latency hiding by occupancy
constexpr bool kEpiPrefetch = kEpiWarps <= 8;
#ifndef SKG_EPI_L1_PREFETCH
#define SKG_EPI_L1_PREFETCH 1
#endif
constexpr bool kEpiL1Prefetch = SKG_EPI_L1_PREFETCH != 0;
constexpr int kEpiParts = kEpiWarps / 4;   // column parts of a tile, one epilogue warp each per quadrant
constexpr int kGemmThreads = 32 * (4 + kEpiWarps);  // warps: 0 TMA, 1 MMA, 2 TMEM alloc, 3 idle, 4..(3 + kEpiWarps) epilogue
// epilogue features (compile-time mask FEAT of gemm_bf16_kernel; bias / ReLU / out_row are always available)
#ifndef SKG_EPI_SETMAXNREG
#define SKG_EPI_SETMAXNREG 1
#endif
constexpr bool kEpiSetMaxNReg = SKG_EPI_SETMAXNREG != 0;   // 0: every warp keeps the launch's register count (A/B builds)
enum : uint32_t { F_P = 1, F_Q = 2, F_RAW = 4, F_RES = 8, F_DOT = 16, F_OB = 32, F_OF = 64, F_LO = 128, F_MASK = 256, F_ALL = 511 };
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
// ---- CTA pair (cta_group::2): the two CTAs of a cluster run ONE M = 256 MMA; each loads its own 128 rows of A and HALF of
// the B tile, so the operand traffic per output element drops by a third (the fused fc_2 shape is bound by the L2 -> SM
// operand stream: 403 MB per launch at ~12 TB/s of L2 read throughput with 128 x 256 tiles per CTA)
constexpr uint32_t kPeerBitMask = 0xFEFFFFFFu;   // shared::cluster address of the same offset in the pair's even (leader) CTA
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;\n" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;\n" ::: "memory");
}
// TMA load whose completion bytes are credited to the LEADER CTA's mbarrier (executed by both CTAs of the pair)
__device__ __forceinline__ void tma_load_2d_pair(const CUtensorMap* map, uint64_t* bar, void* smem_dst, int c_inner, int c_outer) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];\n" ::"r"(
          smem_u32(smem_dst)),
      "l"(map), "r"(smem_u32(bar) & kPeerBitMask), "r"(c_inner), "r"(c_outer)
      : "memory");
}
__device__ __forceinline__ void umma_bf16_pair(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// commit of the pair's MMAs: arrives on the barrier at this offset in BOTH CTAs
__device__ __forceinline__ void tcgen05_commit_pair(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;\n" ::"r"(
                   smem_u32(bar)),
               "h"((uint16_t)3)
               : "memory");
}
// arrive on the LEADER CTA's barrier at this offset (from either CTA of the pair)
__device__ __forceinline__ void mbar_arrive_leader(uint64_t* bar) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];\n" ::"r"(smem_u32(bar) & kPeerBitMask) : "memory");
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

__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
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

// explicit shared-space accesses: the 1024-byte re-alignment of the dynamic smem base goes through uintptr_t, after which
// the compiler no longer knows the address space and would emit generic LD / ST (long-scoreboard, behind the global queue)
__device__ __forceinline__ void sts_f4(uint32_t addr, float4 v) {
  asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};\n" ::"r"(addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ float4 lds_f4(uint32_t addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];\n" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr) : "memory");
  return v;
}

__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}

// ---------------------------------------------------------------------------------------------
// kernel
// ---------------------------------------------------------------------------------------------
// GATHER: the epilogue multiplies by row-gathered tables.  Those launches trade one pipeline stage (48 KB) for L1 capacity:
// the carve-out drops from 228 KB to 164 KB, the tile's table rows (32 KB at n = 32) then stay L1-resident across the four
// lane quadrants; plain launches keep the deeper pipeline (measured: 4 -> 3 stages costs the mainloop 3 us at K = 1024).
template <int BN, bool GATHER, bool TWO = false>
struct GemmSmem {
#ifndef SKG_PAIR_STAGES
#define SKG_PAIR_STAGES 5
#endif
#ifndef SKG_PAIR_STAGES_GATHER
#define SKG_PAIR_STAGES_GATHER 4
#endif
  static constexpr int kStages = TWO ? (GATHER ? SKG_PAIR_STAGES_GATHER : SKG_PAIR_STAGES) : (BN == 256) ? (GATHER ? 3 : 4) : (BN == 128 ? 6 : 7);
  static constexpr int kABytes = BM * BK * 2;
  static constexpr int kBRows = TWO ? BN / 2 : BN;     // CTA pair: each CTA stages half of the B tile
  static constexpr int kBBytes = kBRows * BK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStagingBytes = kEpiWarps * 2048;   // one swizzled 32 x 16 fp32 patch per epilogue warp
  static constexpr int kTotal = kStages * kStageBytes + 1024 /*align slack*/ + 256 /*barriers*/ + kStagingBytes;
};

template <int BN, uint32_t FEAT, bool SK, bool TWO = false>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, int M, int N, int K,
                 const __grid_constant__ EpiGroups epi, int splits, float* __restrict__ sk_ws, int* __restrict__ sk_cnt, int a_k_wrap) {
  static_assert(!TWO || (BN == 256 && !SK), "the CTA-pair variant is the 256 x 256 pair tile without split-K");
  using S = GemmSmem<BN, (FEAT & (F_P | F_Q)) != 0 && FEAT != F_ALL, TWO>;
  constexpr int TM = TWO ? 2 * BM : BM;                 // rows of a work item (pair tile: 256, 128 per CTA)
  const uint32_t cta_rank = TWO ? cluster_ctarank() : 0u;
  const bool leader = cta_rank == 0;
  const int first_item = TWO ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
  const int item_stride = TWO ? (int)(gridDim.x >> 1) : (int)gridDim.x;
  constexpr int kStages = S::kStages;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kStages * S::kStageBytes);
  uint64_t* full_bar = bars;                     // [kStages]  TMA -> MMA
  uint64_t* empty_bar = bars + kStages;          // [kStages]  MMA -> TMA
  uint64_t* tmem_full = bars + 2 * kStages;      // [2]        MMA -> epilogue
  uint64_t* tmem_empty = bars + 2 * kStages + 2; // [2]        epilogue -> MMA
  uint32_t* tmem_base_slot = reinterpret_cast<uint32_t*>(bars + 2 * kStages + 4);
  float* staging = reinterpret_cast<float*>(smem + kStages * S::kStageBytes + 256);
  volatile int* sk_flag = reinterpret_cast<volatile int*>(tmem_base_slot + 1);   // split-K: 'this CTA owns the epilogue'

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int num_m = (M + TM - 1) / TM, num_n = (N + BN - 1) / BN;
  const int num_tiles = num_m * num_n;
  const int num_k = (K + BK - 1) / BK;
  // split-K: work item = (tile, split); the splits of one tile sit on adjacent CTAs and reduce through fp32
  // red.global.add into a zeroed scratch tile; the last split to arrive runs the epilogue and re-zeroes it.
  const int num_items = num_tiles * splits;

  // programmatic dependent launch: the next kernel of the chain may be scheduled right away (its prologue overlaps
  // this kernel's tail); this kernel's own prologue below touches nothing a predecessor wrote
  asm volatile("griddepcontrol.launch_dependents;\n" ::: "memory");
  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];\n" ::"l"(&tmA) : "memory");
    asm volatile("prefetch.tensormap [%0];\n" ::"l"(&tmB) : "memory");
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kStages; ++s) { mbar_init(full_bar + s, 1); mbar_init(empty_bar + s, 1); }
    // pair: the leader's tmem_empty barriers collect the epilogue warps of BOTH CTAs
    for (int a = 0; a < 2; ++a) { mbar_init(tmem_full + a, 1); mbar_init(tmem_empty + a, TWO ? 2 * kEpiWarps : kEpiWarps); }
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  if (warp == 2) {   // 2 x BN fp32 accumulator columns (power of two >= 32); pair: the same warp of both CTAs, same columns
    if (TWO) {
      asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(smem_u32(tmem_base_slot)), "r"(2 * BN));
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;\n" ::: "memory");
    } else {
      asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(smem_u32(tmem_base_slot)), "r"(2 * BN));
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::: "memory");
    }
  }
  tcgen05_fence_before();
  if (TWO) cluster_sync_all(); else __syncthreads();   // pair: nobody signals a peer barrier before it is initialised
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_base_slot;
  asm volatile("griddepcontrol.wait;\n" ::: "memory");   // every predecessor grid has completed, its writes are visible

  // Register reallocation between the warpgroups for the gather epilogues (the launch gives every thread 168): the TMA / MMA
  // warpgroup runs two one-lane loops and hands registers to the two epilogue warpgroups, whose software-pipelined table
  // rows spilled at 168 (same-box A/B on the fused fc_2 launch, 4096 x 4096 x 1024: 48.2 -> 46.5 us).  The setmaxnreg sits
  // INSIDE the two role branches: ptxas budgets each branch by the setmaxnreg that dominates it.  Plain epilogues do not use
  // it: they gain nothing (37.8 us either way), and with it on every launch the C2 step measured 2-3 % slower.
  constexpr bool kSetMax = kEpiSetMaxNReg && kEpiWarps == 8 && (FEAT & F_P) != 0;
  // the pool must balance or setmaxnreg.inc waits forever: a kernel that contains setmaxnreg is launched with the largest
  // count its launch bounds allow (65536 / 384 threads -> 168), warps 0..3 give back (168 - 40) x 128 registers and warps
  // 4..11 take (232 - 168) x 256
  static_assert(kGemmThreads != 384 || (168 - 40) * 128 >= (232 - 168) * 256, "setmaxnreg: the epilogue asks for more than the producer frees");
  if (warp < kEpiWarp0) {
  if (kSetMax) asm volatile("setmaxnreg.dec.sync.aligned.u32 40;\n");
  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      for (int item = first_item; item < num_items; item += item_stride) {
        const int tile = item / splits, split = item - tile * splits;
        const int m0 = (tile / num_n) * TM + (int)cta_rank * BM, n0 = (tile % num_n) * BN;
        const int kb0 = (int)((long long)split * num_k / splits), kb1 = (int)((long long)(split + 1) * num_k / splits);
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(empty_bar + stage, phase ^ 1);
          uint8_t* sa = smem + stage * S::kStageBytes;
          uint8_t* sb = sa + S::kABytes;
          // a_k_wrap: A holds only a_k_wrap columns and is read cyclically along K (split-bf16 predictor: [hi | lo | hi])
          const int ka = (a_k_wrap > 0) ? (kb * BK) % a_k_wrap : kb * BK;
          if (TWO) {
            // both CTAs load (own A rows, own half of the B tile); all four copies complete on the LEADER's full barrier
            if (leader) mbar_arrive_expect_tx(full_bar + stage, 2 * S::kStageBytes);
            tma_load_2d_pair(&tmA, full_bar + stage, sa, ka, m0);
            tma_load_2d_pair(&tmB, full_bar + stage, sb, kb * BK, n0 + (int)cta_rank * S::kBRows);
          } else {
            mbar_arrive_expect_tx(full_bar + stage, S::kStageBytes);
            tma_load_2d(&tmA, full_bar + stage, sa, ka, m0);
            tma_load_2d(&tmB, full_bar + stage, sb, kb * BK, n0);
          }
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0 && leader) {
      constexpr uint32_t idesc = make_idesc(TM, BN);
      int stage = 0; uint32_t phase = 0;
      int acc = 0; uint32_t acc_phase = 0;
      for (int item = first_item; item < num_items; item += item_stride) {
        const int split = item % splits;
        const int kb0 = (int)((long long)split * num_k / splits), kb1 = (int)((long long)(split + 1) * num_k / splits);
        mbar_wait(tmem_empty + acc, acc_phase ^ 1);      // epilogue has drained this accumulator
        tcgen05_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(acc * BN);
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(full_bar + stage, phase);
          tcgen05_fence_after();
          const uint32_t sa = smem_u32(smem + stage * S::kStageBytes);
          const uint32_t sb = sa + S::kABytes;
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            const uint64_t da = make_smem_desc(sa + k * UMMA_K * 2);
            const uint64_t db = make_smem_desc(sb + k * UMMA_K * 2);
            if (TWO) umma_bf16_pair(d_tmem, da, db, idesc, (kb > kb0 || k != 0) ? 1u : 0u);
            else umma_bf16(d_tmem, da, db, idesc, (kb > kb0 || k != 0) ? 1u : 0u);
          }
          if (TWO) tcgen05_commit_pair(empty_bar + stage); // pair: frees the slot in BOTH CTAs
          else tcgen05_commit(empty_bar + stage);          // smem slot free once these MMAs retire
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
        if (TWO) tcgen05_commit_pair(tmem_full + acc);    // accumulator complete -> the epilogue warps of both CTAs
        else tcgen05_commit(tmem_full + acc);             // accumulator complete -> epilogue
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  }
  } else {
    if (kSetMax) asm volatile("setmaxnreg.inc.sync.aligned.u32 232;\n");
    // ===================== epilogue =====================
    // TMEM hands every thread one accumulator ROW (lane = row).  Touching global memory in that layout makes
    // each warp-wide access hit 32 different 128-byte lines, which the L1 serialises; so every 32x16 block is
    // transposed through a private, XOR-swizzled 2 KB shared-memory patch and processed with
    // lane = (row % 8, 4 columns): a warp-wide float4 access then covers 8 rows x 64 contiguous bytes.
    // kEpiWarps epilogue warps (kEpiParts per TMEM lane quadrant, 1 / kEpiParts of the tile's columns each: 8 warps = two
    // per scheduler with the gathered operands software-pipelined in registers); FEAT removes at compile time every
    // epilogue feature the launch does not use.
    const int ew = warp - kEpiWarp0;
    const int quad = warp & 3;                            // TMEM lane quadrant this warp may access
    const int part = ew >> 2;                             // which quarter of the tile's columns this warp drains
    const uint32_t patch = smem_u32(staging) + (uint32_t)ew * 2048u;   // 32 rows x 4 float4 (shared-space address)
    const int sub = lane >> 2;                            // row inside an 8-row group
    const int cq = (lane & 3) * 4;                        // first of this lane's 4 columns inside a 16-column block
    constexpr int kBlocks = BN / (16 * kEpiParts);        // 16-column blocks per warp (even, or 1)
    static_assert(kBlocks == 1 || kBlocks % 2 == 0, "the pipelined epilogue walks the blocks in pairs");
    int acc = 0; uint32_t acc_phase = 0;
    for (int item = first_item; item < num_items; item += item_stride) {
      const int tile = item / splits;
      const int m_blk = tile / num_n, n_blk = tile % num_n;
      const int m0 = m_blk * TM + (int)cta_rank * BM, n0 = n_blk * BN;
      int gi = 0;
      for (int q = 1; q < epi.num; ++q) if (n0 >= epi.g[q].n_begin) gi = q;
      const skg_epilogue_group& G = epi.g[gi];
      // hoist the group descriptor out of the inner loops (it lives in the indexed constant bank)
      const bool relu = (G.flags & SKG_EPI_RELU) != 0;
      const int n_begin = G.n_begin;
      const float* __restrict__ g_bias = G.bias;
      const float* __restrict__ g_p = (FEAT & F_P) ? G.mul_p : nullptr;
      const float* __restrict__ g_q = (FEAT & F_Q) ? G.mul_q : nullptr;
      const float* __restrict__ g_res = (FEAT & F_RES) ? G.residual : nullptr;
      const float* __restrict__ g_dw = (FEAT & F_DOT) ? G.dot_w : nullptr;
      __nv_bfloat16* g_ob = (FEAT & F_OB) ? reinterpret_cast<__nv_bfloat16*>(G.out_bf16) : nullptr;
      __nv_bfloat16* g_raw = (FEAT & F_RAW) ? reinterpret_cast<__nv_bfloat16*>(G.raw_bf16) : nullptr;
      __nv_bfloat16* g_lo = (FEAT & F_LO) ? reinterpret_cast<__nv_bfloat16*>(G.lo_bf16) : nullptr;
      float* g_of = (FEAT & F_OF) ? G.out_f32 : nullptr;
      const __nv_bfloat16* __restrict__ g_mask = (FEAT & F_MASK) ? reinterpret_cast<const __nv_bfloat16*>(G.mask_bf16) : nullptr;
      const int ld_mask = G.ld_mask;
      const int ld_p = G.ld_p, ld_q = G.ld_q, ld_r = G.ld_r, ld_ob = G.ld_ob, ld_raw = G.ld_raw, ld_lo = G.ld_lo, ld_of = G.ld_of;
      // per-row maps: loaded by the lane that owns the row, then handed to the lanes that process it
      int rp_i[4], rq_i[4], rr_i[4], ro_i[4];
      {
        const int row = m0 + quad * 32 + lane;
        const bool row_ok = row < M;
        int rp = row, rq = row, rr = row, ro = row;
        if (row_ok) {
          if ((FEAT & F_P) && G.row_p) rp = __ldg(G.row_p + row);
          if ((FEAT & F_Q) && G.row_q) rq = __ldg(G.row_q + row);
          if ((FEAT & F_RES) && G.row_r) rr = __ldg(G.row_r + row);
          if (G.out_row) ro = __ldg(G.out_row + row);
        } else {
          rp = rq = rr = 0; ro = -1;
        }
        // The gathered table rows of this tile are first-touch L2 hits (~1 us under load) and the one-block-ahead register
        // prefetch below covers ~300 cycles of them: pull the tile's table lines into L1 NOW, one row per lane, while the
        // warp would otherwise sit waiting for the accumulator (ncu: stall_long_sb on the first use of the gathered operands
        // was 20 % of the kernel's active samples)
        if (kEpiL1Prefetch && row_ok) {
          constexpr int kLines = (BN / kEpiParts) / 32;          // 128-byte lines of this warp's column part per row
          const int lc0 = n0 - n_begin + part * (BN / kEpiParts);
          if (n0 + part * (BN / kEpiParts) < N) {
#pragma unroll
            for (int j = 0; j < kLines; ++j) {
              if ((FEAT & F_P) && g_p) asm volatile("prefetch.global.L1 [%0];\n" ::"l"(g_p + (int64_t)rp * ld_p + lc0 + j * 32));
              if ((FEAT & F_Q) && g_q) asm volatile("prefetch.global.L1 [%0];\n" ::"l"(g_q + (int64_t)rq * ld_q + lc0 + j * 32));
              if ((FEAT & F_RES) && g_res) asm volatile("prefetch.global.L1 [%0];\n" ::"l"(g_res + (int64_t)rr * ld_r + lc0 + j * 32));
            }
          }
        }
#pragma unroll
        for (int it = 0; it < 4; ++it) {
          const int src = it * 8 + sub;
          rp_i[it] = __shfl_sync(0xffffffffu, rp, src); rq_i[it] = __shfl_sync(0xffffffffu, rq, src);
          rr_i[it] = __shfl_sync(0xffffffffu, rr, src); ro_i[it] = __shfl_sync(0xffffffffu, ro, src);
        }
      }
      const uint32_t t_row = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(acc * BN);
      const bool split_mode = SK && splits > 1;
      float* ws_rows = split_mode ? sk_ws + (size_t)tile * (BM * BN) + (size_t)(quad * 32 + sub) * BN : nullptr;
      bool is_last = false;
      // pass 0 reads the accumulator from TMEM: without split-K it runs the fused epilogue; with split-K it only adds
      // the partial tile into the scratch tile.  pass 1 (last-arriving split only) runs the epilogue from the scratch tile.
#pragma unroll 1
      for (int pass = 0; pass < (SK ? 2 : 1); ++pass) {
        if (pass == 1 && !is_last) break;
        const bool do_math = (pass == 1) || !split_mode;
        float dot[4] = {0.f, 0.f, 0.f, 0.f};
        bool waited = false;
        // The gathered operands of a 16-column block (bias, dot weights, Hadamard / residual table rows) do not depend on
        // the accumulator.  They are software-pipelined one block ahead in registers: with one warp per scheduler
        // quadrant-part there is no other work to hide an L2 round trip behind, and shared memory is full of pipeline
        // stages, so the L1 that would otherwise catch these re-reads is down to a few KB.
        struct Tabs { float4 b4, w4, tp[4], tq[4], tr[4]; uint2 tm[4]; };
        auto load_tabs = [&](int blk, Tabs& T) {
          const int c0 = part * (BN / kEpiParts) + blk * 16;
          const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
          T.b4 = z4; T.w4 = z4;
          if (!do_math || n0 + c0 >= N) return;
          const int col = n0 + c0 + cq, lc = col - n_begin;
          if (g_bias) T.b4 = __ldg(reinterpret_cast<const float4*>(g_bias + col));
          if ((FEAT & F_DOT) && g_dw) T.w4 = __ldg(reinterpret_cast<const float4*>(g_dw + lc));
          if ((FEAT & F_P) && g_p) {
#pragma unroll
            for (int it = 0; it < 4; ++it) T.tp[it] = __ldg(reinterpret_cast<const float4*>(g_p + (int64_t)rp_i[it] * ld_p + lc));
          }
          if ((FEAT & F_Q) && g_q) {
#pragma unroll
            for (int it = 0; it < 4; ++it) T.tq[it] = __ldg(reinterpret_cast<const float4*>(g_q + (int64_t)rq_i[it] * ld_q + lc));
          }
          if ((FEAT & F_RES) && g_res) {
#pragma unroll
            for (int it = 0; it < 4; ++it) T.tr[it] = __ldg(reinterpret_cast<const float4*>(g_res + (int64_t)rr_i[it] * ld_r + lc));
          }
          if ((FEAT & F_MASK) && g_mask) {   // ReLU-backward gate rows: indexed by the GEMM row (rows beyond M read row 0)
#pragma unroll
            for (int it = 0; it < 4; ++it) {
              const int grow = m0 + quad * 32 + it * 8 + sub;
              T.tm[it] = __ldg(reinterpret_cast<const uint2*>(g_mask + (int64_t)(grow < M ? grow : 0) * ld_mask + lc));
            }
          }
        };
        auto process = [&](int blk, const Tabs& T) {
          const int c0 = part * (BN / kEpiParts) + blk * 16;
          if (n0 + c0 >= N) return;                    // warp-uniform
          const int col = n0 + c0 + cq;                // global column of this lane's 4 values
          const int lc = col - n_begin;                // column inside the group (table / output column)
          const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
          const float4 b4 = T.b4, w4 = T.w4;
          if (pass == 0) {
            if (!waited) {                             // first block of the tile: wait for the accumulator
              mbar_wait(tmem_full + acc, acc_phase);
              tcgen05_fence_after();
              waited = true;
            }
            uint32_t raw[16];
            tmem_ld16(t_row + c0, raw);
            __syncwarp();                              // the patch's previous readers are done
#pragma unroll
            for (int j = 0; j < 4; ++j)
              sts_f4(patch + (uint32_t)(lane * 4 + (j ^ ((lane >> 1) & 3))) * 16u,
                     make_float4(__uint_as_float(raw[4 * j]), __uint_as_float(raw[4 * j + 1]), __uint_as_float(raw[4 * j + 2]),
                                 __uint_as_float(raw[4 * j + 3])));
            __syncwarp();
          }
#pragma unroll
          for (int it = 0; it < 4; ++it) {
            const int r = it * 8 + sub;
            float4 v;
            if (pass == 0) {
              v = lds_f4(patch + (uint32_t)(r * 4 + ((lane & 3) ^ ((r >> 1) & 3))) * 16u);
              if (split_mode) {                        // partial sum -> scratch tile (fp32 reduction in L2)
                if (m0 + quad * 32 + r < M) {
                  float* dst = ws_rows + (size_t)(it * 8) * BN + c0 + cq;
                  asm volatile("red.relaxed.gpu.global.add.v4.f32 [%0], {%1, %2, %3, %4};\n" ::"l"(dst), "f"(v.x), "f"(v.y),
                               "f"(v.z), "f"(v.w)
                               : "memory");
                }
                continue;
              }
            } else {                                   // complete sum from the scratch tile; leave it zeroed
              float4* src = reinterpret_cast<float4*>(ws_rows + (size_t)(it * 8) * BN + c0 + cq);
              if (m0 + quad * 32 + r < M) { v = __ldcg(src); __stcg(src, z4); } else v = z4;
            }
            const int ro = ro_i[it];
            const bool store_ok = ro >= 0;             // also false for rows beyond M
            v.x += b4.x; v.y += b4.y; v.z += b4.z; v.w += b4.w;
            if ((FEAT & F_RAW) && g_raw && store_ok)
              *reinterpret_cast<uint2*>(g_raw + (int64_t)ro * ld_raw + lc) = make_uint2(pack_bf16(v.x, v.y), pack_bf16(v.z, v.w));
            if ((FEAT & F_P) && g_p) {
              float4 a = T.tp[it];
              if ((FEAT & F_Q) && g_q) { a.x += T.tq[it].x; a.y += T.tq[it].y; a.z += T.tq[it].z; a.w += T.tq[it].w; }
              v.x *= a.x; v.y *= a.y; v.z *= a.z; v.w *= a.w;
            }
            if (relu) { v.x = fmaxf(v.x, 0.f); v.y = fmaxf(v.y, 0.f); v.z = fmaxf(v.z, 0.f); v.w = fmaxf(v.w, 0.f); }
            if ((FEAT & F_RES) && g_res) { v.x += T.tr[it].x; v.y += T.tr[it].y; v.z += T.tr[it].z; v.w += T.tr[it].w; }
            if (FEAT & F_DOT) dot[it] += v.x * w4.x + v.y * w4.y + v.z * w4.z + v.w * w4.w;
            if ((FEAT & F_MASK) && g_mask) {            // bf16 > 0  <=>  sign bit clear and magnitude bits set
              const uint2 mk = T.tm[it];
              if ((mk.x & 0x8000u) || !(mk.x & 0x7fffu)) v.x = 0.f;
              if ((mk.x & 0x80000000u) || !(mk.x & 0x7fff0000u)) v.y = 0.f;
              if ((mk.y & 0x8000u) || !(mk.y & 0x7fffu)) v.z = 0.f;
              if ((mk.y & 0x80000000u) || !(mk.y & 0x7fff0000u)) v.w = 0.f;
            }
            if (store_ok) {
              if ((FEAT & F_OB) && g_ob)
                *reinterpret_cast<uint2*>(g_ob + (int64_t)ro * ld_ob + lc) = make_uint2(pack_bf16(v.x, v.y), pack_bf16(v.z, v.w));
              if ((FEAT & F_LO) && g_lo) {   // residual of the bf16 rounding: v ~= hi + lo to ~16 mantissa bits
                const float lx = v.x - __bfloat162float(__float2bfloat16(v.x)), ly = v.y - __bfloat162float(__float2bfloat16(v.y));
                const float lz = v.z - __bfloat162float(__float2bfloat16(v.z)), lw = v.w - __bfloat162float(__float2bfloat16(v.w));
                *reinterpret_cast<uint2*>(g_lo + (int64_t)ro * ld_lo + lc) = make_uint2(pack_bf16(lx, ly), pack_bf16(lz, lw));
              }
              if ((FEAT & F_OF) && g_of) {
                float* dst = g_of + (int64_t)ro * ld_of + lc;
                if (col + 4 <= N && (ld_of & 3) == 0) {
                  *reinterpret_cast<float4*>(dst) = v;
                } else {
                  if (col < N) dst[0] = v.x;
                  if (col + 1 < N) dst[1] = v.y;
                  if (col + 2 < N) dst[2] = v.z;
                  if (col + 3 < N) dst[3] = v.w;
                }
              }
            }
          }
        };
        if (kBlocks == 1 || !kEpiPrefetch) {
          Tabs T0;
#pragma unroll 1
          for (int blk = 0; blk < kBlocks; ++blk) {
            load_tabs(blk, T0);
            process(blk, T0);
          }
        } else {
          Tabs T0, T1;
          load_tabs(0, T0);
#pragma unroll 1
          for (int blk = 0; blk < kBlocks; blk += 2) {
            load_tabs(blk + 1, T1);
            process(blk, T0);
            if (blk + 2 < kBlocks) load_tabs(blk + 2, T0);
            process(blk + 1, T1);
          }
        }
        if (pass == 0) {
          if (!waited) {                               // this warp's columns lie beyond N: still consume the phase
            mbar_wait(tmem_full + acc, acc_phase);
            tcgen05_fence_after();
          }
          // accumulator drained: hand the TMEM buffer back to the MMA warp
          tcgen05_fence_before();
          __syncwarp();
          if (lane == 0) { if (TWO) mbar_arrive_leader(tmem_empty + acc); else mbar_arrive(tmem_empty + acc); }
          if (split_mode) {                            // threadfence-reduction: who adds the last partial owns the epilogue
            __threadfence();
            asm volatile("bar.sync 1, %0;\n" ::"n"(kEpiWarps * 32) : "memory");
            if (ew == 0 && lane == 0) {
              const int prev = atomicAdd(sk_cnt + tile, 1);
              const int last = (prev == splits - 1) ? 1 : 0;
              if (last) sk_cnt[tile] = 0;              // self-resetting: the next launch finds zeroed counters
              *sk_flag = last;
            }
            asm volatile("bar.sync 1, %0;\n" ::"n"(kEpiWarps * 32) : "memory");
            is_last = (*sk_flag != 0);
            __threadfence();
          }
        }
        if (do_math && (FEAT & F_DOT) && G.dot_out) {   // four partials per N tile (one per column quarter), summed by the consumer
#pragma unroll
          for (int it = 0; it < 4; ++it) {
            float d = dot[it];
            d += __shfl_xor_sync(0xffffffffu, d, 1);
            d += __shfl_xor_sync(0xffffffffu, d, 2);
            if ((lane & 3) == 0 && ro_i[it] >= 0) {       // ABI: four partial slots per N tile; unused ones are written as 0
              float* slot = G.dot_out + (int64_t)ro_i[it] * G.ld_dot + ((n0 - n_begin) / BN) * 4 + part * (4 / kEpiParts);
              slot[0] = d;
#pragma unroll
              for (int z = 1; z < 4 / kEpiParts; ++z) slot[z] = 0.f;
            }
          }
        }
      }
      if (split_mode) asm volatile("bar.sync 1, %0;\n" ::"n"(kEpiWarps * 32) : "memory");   // sk_flag is rewritten by the next item
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  }

  tcgen05_fence_before();
  if (TWO) cluster_sync_all(); else __syncthreads();   // pair: neither CTA leaves while the other may still read its smem / signal it
  if (warp == 2) {
    tcgen05_fence_after();
    if (TWO) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;\n" ::"r"(tmem_base), "r"(2 * BN));
    else asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" ::"r"(tmem_base), "r"(2 * BN));
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

// Descriptor cache: a tensor map is a pure function of (address, rows, cols, pitch, box rows), and a serving / training loop
// launches the same few dozen operand shapes at the same workspace addresses step after step -- cuTensorMapEncodeTiled (two
// driver calls per GEMM launch, ~60 per step on the plain-launch path) is paid once per distinct operand.
struct TmapKey {
  const void* ptr; int64_t rows, cols, ld; int box_rows;
  bool operator==(const TmapKey& o) const { return ptr == o.ptr && rows == o.rows && cols == o.cols && ld == o.ld && box_rows == o.box_rows; }
};
struct TmapSlot { TmapKey key; CUtensorMap map; bool used; };
constexpr int kTmapSlots = 1024;   // direct-mapped
static thread_local TmapSlot* g_tmaps = nullptr;
static thread_local long long g_tmap_stats[2] = {0, 0};   // hits, encodes

static int make_tmap_uncached(CUtensorMap* map, const void* ptr, int64_t rows, int64_t cols, int64_t ld, int box_rows);

static int make_tmap(CUtensorMap* map, const void* ptr, int64_t rows, int64_t cols, int64_t ld, int box_rows) {
  if (!g_tmaps) g_tmaps = new TmapSlot[kTmapSlots]();
  const TmapKey key{ptr, rows, cols, ld, box_rows};
  uint64_t h = reinterpret_cast<uintptr_t>(ptr) >> 4;
  h = (h ^ (uint64_t)rows * 0x9E3779B97F4A7C15ull ^ (uint64_t)cols * 0xC2B2AE3D27D4EB4Full ^ (uint64_t)ld * 0x165667B19E3779F9ull ^ (uint64_t)box_rows) * 0xD6E8FEB86659FD93ull;
  TmapSlot& slot = g_tmaps[(h >> 32) % kTmapSlots];
  if (slot.used && slot.key == key) { *map = slot.map; g_tmap_stats[0]++; return SKG_OK; }
  const int rc = make_tmap_uncached(map, ptr, rows, cols, ld, box_rows);
  if (rc) return rc;
  slot.key = key; slot.map = *map; slot.used = true; g_tmap_stats[1]++;
  return SKG_OK;
}

// 2-D bf16 row-major [rows, cols] (cols contiguous, row pitch ld elements), box = [box_rows, 64 cols], 128B swizzle
static int make_tmap_uncached(CUtensorMap* map, const void* ptr, int64_t rows, int64_t cols, int64_t ld, int box_rows) {
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

struct SplitK {
  int splits;
  float* ws;
  int* cnt;
  int a_k_wrap = 0;
};

template <int BN, uint32_t FEAT, bool SK, bool TWO = false>
static int launch_gemm(const CUtensorMap& tmA, const CUtensorMap& tmB, int M, int N, int K, const EpiGroups& epi,
                       const SplitK& sk, cudaStream_t stream) {
  const int a_k_wrap = sk.a_k_wrap;
  using S = GemmSmem<BN, (FEAT & (F_P | F_Q)) != 0 && FEAT != F_ALL, TWO>;
  auto kern = gemm_bf16_kernel<BN, FEAT, SK, TWO>;
  static bool attr_set = false;
  if (!attr_set) {
    SKG_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, S::kTotal));
    attr_set = true;
  }
  const int items = ceil_div(M, TWO ? 2 * BM : BM) * ceil_div(N, BN) * sk.splits;
  int grid = items < num_sms() ? items : num_sms();
  if (TWO) { const int pairs = num_sms() / 2; grid = 2 * (items < pairs ? items : pairs); }   // one cluster of two CTAs per pair tile
  static const bool pdl = [] { const char* e = getenv("SKG_GEMM_PDL"); return !(e && e[0] == '0'); }();
  if (pdl || TWO) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(kGemmThreads); cfg.dynamicSmemBytes = S::kTotal; cfg.stream = stream;
    cudaLaunchAttribute at[2];
    int na = 0;
    if (pdl) {
      at[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
      at[na].val.programmaticStreamSerializationAllowed = 1;
      ++na;
    }
    if (TWO) {
      at[na].id = cudaLaunchAttributeClusterDimension;
      at[na].val.clusterDim.x = 2; at[na].val.clusterDim.y = 1; at[na].val.clusterDim.z = 1;
      ++na;
    }
    cfg.attrs = at; cfg.numAttrs = na;
    SKG_CUDA(cudaLaunchKernelEx(&cfg, kern, tmA, tmB, M, N, K, epi, sk.splits, sk.ws, sk.cnt, a_k_wrap));
  } else {
    kern<<<grid, kGemmThreads, S::kTotal, stream>>>(tmA, tmB, M, N, K, epi, sk.splits, sk.ws, sk.cnt, a_k_wrap);
  }
  return check_launch("gemm_bf16_kernel");
}

}  // namespace skg

using namespace skg;

extern "C" void skg_tensor_map_cache_stats(long long* hits, long long* encodes) {
  if (hits) *hits = g_tmap_stats[0];
  if (encodes) *encodes = g_tmap_stats[1];
}

constexpr int kSplitKCounters = 1024;   // ints at the head of the split-K scratch

// split-K pays when the tiles alone leave most SMs idle (node-row GEMMs: M = humans / boxes of the batch)
static int choose_splits(int M, int N, int K, int block_n, int64_t ws_bytes) {
  const int tiles = ceil_div(M, BM) * ceil_div(N, block_n);
  const int num_k = ceil_div(K, BK);
  // the fp32 reduction through L2 costs ~10 us: it only pays for long-K problems (box_head: K = 12544)
  if (tiles * 2 > num_sms() || tiles > kSplitKCounters || num_k < 64) return 1;
  if (ws_bytes < (int64_t)kSplitKCounters * 4 + (int64_t)tiles * BM * block_n * 4) return 1;
  static const int min_kb = [] { const char* e = getenv("SKG_SPLITK_MIN_KB"); const int v = e ? atoi(e) : 16; return v > 0 ? v : 16; }();
  int splits = num_sms() / tiles;
  if (splits > num_k / min_kb) splits = num_k / min_kb;   // at least 16 k-blocks (K = 1024) per split (knob: experiments)
  return splits < 1 ? 1 : splits;
}

extern "C" int64_t skg_gemm_splitk_scratch_bytes(int max_tiles_m, int max_n) {
  // enough for any GEMM the split-K heuristic accepts: tiles * 2 <= SMs, 128 x 256 fp32 scratch tile each
  (void)max_tiles_m; (void)max_n;
  return (int64_t)kSplitKCounters * 4 + (int64_t)(num_sms() / 2 + 1) * BM * 256 * 4;
}

extern "C" int skg_gemm_bf16(const void* A, int64_t lda, const void* B, int64_t ldb, int M, int N, int K,
                             const skg_epilogue_group* groups, int num_groups, int block_n, cudaStream_t stream) {
  return skg_gemm_bf16_splitk(A, lda, B, ldb, M, N, K, groups, num_groups, block_n, nullptr, 0, stream);
}

extern "C" int skg_gemm_bf16_splitk(const void* A, int64_t lda, const void* B, int64_t ldb, int M, int N, int K,
                                    const skg_epilogue_group* groups, int num_groups, int block_n, void* scratch,
                                    int64_t scratch_bytes, cudaStream_t stream) {
  return skg_gemm_bf16_wrapk(A, lda, 0, B, ldb, M, N, K, groups, num_groups, block_n, scratch, scratch_bytes, stream);
}

extern "C" int skg_gemm_bf16_wrapk(const void* A, int64_t lda, int a_k_wrap, const void* B, int64_t ldb, int M, int N, int K,
                                   const skg_epilogue_group* groups, int num_groups, int block_n, void* scratch,
                                   int64_t scratch_bytes, cudaStream_t stream) {
  SKG_REQUIRE(M >= 0 && N > 0 && K > 0, "bad GEMM shape M=%d N=%d K=%d", M, N, K);
  SKG_REQUIRE(a_k_wrap == 0 || (a_k_wrap % BK == 0 && a_k_wrap > 0 && a_k_wrap <= K), "a_k_wrap must be a multiple of %d within K", BK);
  const int KA = a_k_wrap > 0 ? a_k_wrap : K;      // columns A really has
  if (M == 0) return SKG_OK;
  SKG_REQUIRE(A && B && groups, "null pointer argument");
  SKG_REQUIRE(num_groups >= 1 && num_groups <= kMaxGroups, "num_groups must be in [1, %d]", kMaxGroups);
  SKG_REQUIRE((lda % 8) == 0 && (ldb % 8) == 0 && lda >= KA && ldb >= K, "lda/ldb must be >= K and multiples of 8 elements");
  SKG_REQUIRE(((reinterpret_cast<uintptr_t>(A) | reinterpret_cast<uintptr_t>(B)) & 15u) == 0, "A and B must be 16-byte aligned");
  if (block_n == 0) {
    // widest tile that still spreads the problem over the SMs: node-row GEMMs (M = humans / boxes of the batch,
    // one or two M tiles) are bound by how fast ONE SM can stream its weight slab, so they get narrow N tiles
    const int mt = ceil_div(M, BM);
    block_n = (N % 256 == 0) ? 256 : 128;
    bool has_dot = false;   // the row-dot partial layout depends on block_n: never narrow those launches
    for (int i = 0; i < num_groups; ++i) has_dot |= (groups[i].dot_out != nullptr);
    if (!has_dot) {
    if (block_n == 256 && mt * (N / 256) < num_sms() / 2) block_n = 128;
    // long-K launches with a split-K scratch (box_head fc6: K = 12544) fill the SMs through K splits instead
    const bool split_k = scratch != nullptr && choose_splits(M, N, K, block_n, scratch_bytes) > 1;
    if (block_n == 128 && !split_k && (N % 64 == 0 || N < 128) && mt * ceil_div(N, 128) < num_sms() / 2) block_n = 64;
    }
  }
  SKG_REQUIRE(block_n == 64 || block_n == 128 || block_n == 256, "block_n must be 64, 128 or 256");
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
    SKG_REQUIRE(!g.mask_bf16 || g.ld_mask % 4 == 0, "group %d: ld_mask must be a multiple of 4", i);
    epi.g[i] = g;
  }
  // bias / tables are read in 32-column chunks up to the padded width: callers pad them to block_n multiples
  SplitK sk{1, nullptr, nullptr};
  sk.a_k_wrap = a_k_wrap;
  if (scratch) {
    SKG_REQUIRE((reinterpret_cast<uintptr_t>(scratch) & 15u) == 0, "split-K scratch must be 16-byte aligned");
    sk.splits = block_n == 64 ? 1 : choose_splits(M, N, K, block_n, scratch_bytes);
    if (sk.splits > 1) {
      sk.cnt = reinterpret_cast<int*>(scratch);
      sk.ws = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(scratch) + kSplitKCounters * 4);
    }
  }
  // CTA pairs (cta_group::2, 256 x 256 pair tiles) for the wide cell-row problems: enough pair tiles to fill the 74 SM pairs
  // and no split-K; SKG_GEMM_PAIR=0 keeps every launch on single CTAs (A/B measurements)
  // Measured on B200 (tools/gemm_bench.py --epi, 4096 x 4096 x 1024, flushed single launches): plain epilogue 37.0 us as
  // pairs vs 37.3-38.5 us on single CTAs; the gather epilogue (fused fc_2) 48.8 vs 47.2 us -- it is bound by its own
  // instruction path, not by the operand stream, and loses L1 to the deeper pipeline -- so it stays on single CTAs
  // (SKG_GEMM_PAIR=2 forces pairs there too).
  static const int pair_mode = [] { const char* e = getenv("SKG_GEMM_PAIR"); return e ? atoi(e) : 1; }();
  bool gather = false;
  for (int i = 0; i < num_groups; ++i) gather |= (groups[i].mul_p != nullptr);
  const bool use_pair = pair_mode > 0 && (pair_mode > 1 || !gather) && block_n == 256 && sk.splits == 1 && a_k_wrap == 0 && N % 256 == 0 &&
                        (int64_t)ceil_div(M, 2 * BM) * (N / 256) >= num_sms() / 2;
  CUtensorMap tmA, tmB;
  int rc = make_tmap(&tmA, A, M, KA, lda, BM);
  if (rc) return rc;
  rc = make_tmap(&tmB, B, N, K, ldb, use_pair ? block_n / 2 : block_n);
  if (rc) return rc;
  // pick the leanest compiled epilogue whose feature set covers what the groups ask for
  uint32_t need = 0;
  for (int i = 0; i < num_groups; ++i) {
    const skg_epilogue_group& g = groups[i];
    need |= (g.mul_p ? F_P : 0) | (g.mul_q ? F_Q : 0) | (g.raw_bf16 ? F_RAW : 0) | (g.residual ? F_RES : 0) |
            ((g.dot_w || g.dot_out) ? F_DOT : 0) | (g.out_bf16 ? F_OB : 0) | (g.out_f32 ? F_OF : 0) | (g.lo_bf16 ? F_LO : 0) |
            (g.mask_bf16 ? F_MASK : 0);
  }
#define SKG_GEMM_CASE64(F)                                                                        \
  if (block_n == 64 && (need & ~(uint32_t)(F)) == 0) {                                            \
    sk.splits = 1;                                                                                \
    return launch_gemm<64, (F), false>(tmA, tmB, M, N, K, epi, sk, stream);                       \
  }
  SKG_GEMM_CASE64(F_OB)
  SKG_GEMM_CASE64(F_OF)
  SKG_GEMM_CASE64(F_OB | F_OF)
  SKG_GEMM_CASE64(F_RES | F_OF)
  SKG_GEMM_CASE64(F_MASK | F_OB)
  SKG_GEMM_CASE64(F_RES | F_MASK | F_OB)
  SKG_GEMM_CASE64(F_ALL)
#undef SKG_GEMM_CASE64
#define SKG_GEMM_CASE(F, SKV, PAIR)                                                           \
  if ((need & ~(uint32_t)(F)) == 0) {                                                      \
    if (SKV && sk.splits > 1) {                                                                   \
      if (block_n == 256) return launch_gemm<256, (F), SKV>(tmA, tmB, M, N, K, epi, sk, stream);  \
      return launch_gemm<128, (F), SKV>(tmA, tmB, M, N, K, epi, sk, stream);                      \
    }                                                                                             \
    sk.splits = 1;                                                                                \
    if (block_n == 256) {                                                                         \
      if (PAIR && use_pair) return launch_gemm<256, (F), false, PAIR>(tmA, tmB, M, N, K, epi, sk, stream); \
      if (use_pair) { rc = make_tmap(&tmB, B, N, K, ldb, block_n); if (rc) return rc; }           \
      return launch_gemm<256, (F), false>(tmA, tmB, M, N, K, epi, sk, stream);                    \
    }                                                                                             \
    return launch_gemm<128, (F), false>(tmA, tmB, M, N, K, epi, sk, stream);                      \
  }
  SKG_GEMM_CASE(F_OB, true, true)                   // bias (+ReLU) -> bf16: spatial_head, box_head
  SKG_GEMM_CASE(F_OF, true, false)                  // -> fp32: per-node projections
  SKG_GEMM_CASE(F_OB | F_OF, true, false)           // fc_head / fc_tail
  SKG_GEMM_CASE(F_P | F_Q | F_RAW | F_OB, false, true)   // fused fc_2 of the MBFs (Hadamard with gathered tables)
  SKG_GEMM_CASE(F_DOT, false, true)                 // attention weights -> adjacency row-dot
  SKG_GEMM_CASE(F_RES | F_OF, true, false)          // message fc_3 + residual, split-bf16 predictor second pass
  SKG_GEMM_CASE(F_OB | F_LO, false, true)           // final pair features with rounding residual
  SKG_GEMM_CASE(F_MASK | F_OB, false, true)         // training: dX = (G W) gated by the ReLU mask of the layer's input
  SKG_GEMM_CASE(F_RES | F_MASK | F_OB, false, true) // training: dX + the residual path's gradient, gated
  SKG_GEMM_CASE(F_DOT | F_OB, false, true)          // training forward: adjacency row-dot that also keeps the weights
  SKG_GEMM_CASE(F_ALL, true, true)
#undef SKG_GEMM_CASE
  set_error("no epilogue variant");
  return SKG_ERR_UNSUPPORTED;
}
