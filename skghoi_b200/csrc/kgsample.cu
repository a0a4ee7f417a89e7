// Negative sampler of the knowledge-graph trainer on the device (reference OpenKE C++ sampler):
//   sampling() / getBatch()      OpenKE/openke/base/Base.cpp:78-197
//   corrupt_head / corrupt_tail  OpenKE/openke/base/Corrupt.h:7-83   (filtered: uniform over the entities that do NOT
//                                                                      complete a training triple, by binary search)
//   randd / rand_max             OpenKE/openke/base/Random.h:18-28   (x <- x * 25214903917 + 11 mod 2^64 per worker thread)
//
// The reference cuts a batch into `workThreads` contiguous slices, one pthread each, every thread drawing from its own
// linear congruential stream.  A batch item consumes a FIXED number of draws (1 to pick the positive, 2 per negative in
// mode 0: the head/tail coin and the corrupted entity; 1 per negative in modes -1 / +1), so item j of thread `id` starts
// at stream position (j - lef) * draws: one GPU thread per batch TRIPLE jumps to its item (and on to its negative) in O(log) steps (an LCG composes as
// x -> a^k x + c (a^k - 1)/(a - 1)) and reproduces the reference's output BIT FOR BIT for any thread count, while the
// per-thread seeds are advanced by the whole slice so that consecutive calls continue the same streams.
#include "common.cuh"
#include "skghoi_b200.h"

namespace skg {

constexpr unsigned long long kLcgA = 25214903917ull, kLcgC = 11ull;

struct Lcg {
  unsigned long long s;
  __device__ __forceinline__ unsigned long long next() { s = s * kLcgA + kLcgC; return s; }
  __device__ __forceinline__ int64_t below(int64_t x) { return (int64_t)(next() % (unsigned long long)x); }   // rand_max
};

__host__ __device__ inline unsigned long long lcg_jump(unsigned long long state, unsigned long long k) {
  unsigned long long acc_m = 1, acc_p = 0, cur_m = kLcgA, cur_p = kLcgC;
  while (k) {
    if (k & 1) { acc_m *= cur_m; acc_p = acc_p * cur_m + cur_p; }
    cur_p = (cur_m + 1) * cur_p;
    cur_m *= cur_m;
    k >>= 1;
  }
  return acc_m * state + acc_p;
}

struct KgIndex {
  const int64_t* head;      // [3, T] rows h | r | t, sorted by (h, r, t)   (trainHead == trainList after import)
  const int64_t* tail;      // [3, T] rows h | r | t, sorted by (t, r, h)   (trainTail)
  const int64_t* lef_head;  // [E]
  const int64_t* rig_head;
  const int64_t* lef_tail;
  const int64_t* rig_tail;
  const float* left_mean;   // [R] or null (bern off: prob = 500)
  const float* right_mean;
  int64_t T, E;
};

// Corrupt.h:7-83 with the roles abstracted: `rel` / `other` are the relation and the to-be-replaced entity columns of the
// ordering that is sorted by the FIXED entity e; returns the replacement.
__device__ __forceinline__ int64_t corrupt(const int64_t* __restrict__ rel, const int64_t* __restrict__ other,
                                           const int64_t* __restrict__ lef_arr, const int64_t* __restrict__ rig_arr, int64_t e,
                                           int64_t r, int64_t ent_total, int64_t keep, Lcg& g) {
  int64_t lef = lef_arr[e] - 1, rig = rig_arr[e], mid;
  while (lef + 1 < rig) {
    mid = (lef + rig) >> 1;
    if (rel[mid] >= r) rig = mid; else lef = mid;
  }
  const int64_t ll = rig;
  lef = lef_arr[e]; rig = rig_arr[e] + 1;
  while (lef + 1 < rig) {
    mid = (lef + rig) >> 1;
    if (rel[mid] <= r) lef = mid; else rig = mid;
  }
  const int64_t rr = lef;
  const int64_t free_cnt = ent_total - (rr - ll + 1);
  if (free_cnt <= 0) { g.next(); return keep; }   // (e, r) is linked to EVERY entity: the reference divides by zero here
  const int64_t tmp = g.below(free_cnt);
  if (tmp < other[ll]) return tmp;
  if (tmp > other[rr] - rr + ll - 1) return tmp + rr - ll + 1;
  lef = ll; rig = rr + 1;
  while (lef + 1 < rig) {
    mid = (lef + rig) >> 1;
    if (other[mid] - mid + ll - 1 < tmp) lef = mid; else rig = mid;
  }
  return tmp + lef - ll + 1;
}

__global__ void __launch_bounds__(128) kg_sample_kernel(KgIndex ix, const unsigned long long* __restrict__ seeds_in,
                                                        unsigned long long* __restrict__ seeds_out, int work_threads,
                                                        int64_t batch_size, int neg_rate, int mode,
                                                        int64_t* __restrict__ batch_h, int64_t* __restrict__ batch_t,
                                                        int64_t* __restrict__ batch_r, float* __restrict__ batch_y) {
  // One thread per TRIPLE of the batch: item b = j % batch_size, k = j / batch_size (0 = the positive, k >= 1 = negative
  // k - 1).  A negative needs exactly `dpn` draws (the head/tail coin in mode 0 + the corrupted entity), so it jumps straight
  // to its own position of the worker's stream and re-derives the positive with one extra draw: the neg_rate corruptions of
  // an item (three dependent binary searches each) run in parallel instead of back to back in one thread.
  const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t per = (batch_size % work_threads == 0) ? batch_size / work_threads : batch_size / work_threads + 1;
  const unsigned long long dpn = (mode == 0) ? 2ull : 1ull;
  const unsigned long long draws = 1ull + (unsigned long long)neg_rate * dpn;
  if (j < work_threads) {                       // the seeds after the call: every worker advanced by its whole slice
    int64_t lef = j * per, rig = (j + 1) * per;
    if (rig > batch_size) rig = batch_size;
    const int64_t items = rig > lef ? rig - lef : 0;
    seeds_out[j] = lcg_jump(seeds_in[j], (unsigned long long)items * draws);
  }
  if (j >= batch_size * (1 + (int64_t)neg_rate)) return;
  const int64_t k = j / batch_size, b = j - k * batch_size;
  const int64_t id = b / per, lef0 = id * per;
  Lcg g{lcg_jump(seeds_in[id], (unsigned long long)(b - lef0) * draws)};
  const int64_t T = ix.T;
  const int64_t i = g.below(T);
  const int64_t h = ix.head[i], r = ix.head[T + i], t = ix.head[2 * T + i];
  if (k == 0) {
    batch_h[b] = h; batch_t[b] = t; batch_r[b] = r; batch_y[b] = 1.f;
    return;
  }
  g.s = lcg_jump(g.s, (unsigned long long)(k - 1) * dpn);
  int64_t nh = h, nt = t;
  bool corrupt_tail_entity;                   // true: keep the head, draw a new tail (corrupt_head in the reference's naming)
  if (mode == 0) {
    float prob = 500.f;
    if (ix.left_mean) prob = __fdiv_rn(__fmul_rn(1000.f, ix.right_mean[r]), __fadd_rn(ix.right_mean[r], ix.left_mean[r]));
    corrupt_tail_entity = (float)(g.next() % 1000ull) < prob;
  } else {
    corrupt_tail_entity = (mode != -1);
  }
  if (corrupt_tail_entity)
    nt = corrupt(ix.head + T, ix.head + 2 * T, ix.lef_head, ix.rig_head, h, r, ix.E, t, g);
  else
    nh = corrupt(ix.tail + T, ix.tail, ix.lef_tail, ix.rig_tail, t, r, ix.E, h, g);
  batch_h[j] = nh; batch_t[j] = nt; batch_r[j] = r; batch_y[j] = -1.f;
}

}  // namespace skg

using namespace skg;

extern "C" int skg_kg_sample(const int64_t* head_hrt, const int64_t* tail_hrt, const int64_t* lef_head,
                             const int64_t* rig_head, const int64_t* lef_tail, const int64_t* rig_tail,
                             const float* left_mean, const float* right_mean, int64_t train_total, int64_t ent_total,
                             const uint64_t* seeds_in, uint64_t* seeds_out, int work_threads, int64_t batch_size, int neg_rate,
                             int mode, int64_t* batch_h, int64_t* batch_t, int64_t* batch_r, float* batch_y,
                             cudaStream_t stream) {
  SKG_REQUIRE(head_hrt && tail_hrt && lef_head && rig_head && lef_tail && rig_tail && seeds_in && seeds_out, "null pointer argument");
  SKG_REQUIRE(batch_h && batch_t && batch_r && batch_y, "null output pointer");
  SKG_REQUIRE(train_total > 0 && ent_total > 1 && work_threads >= 1 && batch_size >= 0 && neg_rate >= 0, "bad sizes");
  SKG_REQUIRE(mode == 0 || mode == -1 || mode == 1, "mode must be 0 (head or tail), -1 (head batch) or 1 (tail batch)");
  SKG_REQUIRE((left_mean == nullptr) == (right_mean == nullptr), "left_mean and right_mean come together (bern) or not at all");
  SKG_REQUIRE(seeds_in != seeds_out, "seeds_out must not alias seeds_in (every item reads the seeds of the call's start)");
  KgIndex ix{head_hrt, tail_hrt, lef_head, rig_head, lef_tail, rig_tail, left_mean, right_mean, train_total, ent_total};
  const int64_t triples = batch_size * (1 + (int64_t)neg_rate);
  const int64_t threads = triples > work_threads ? triples : work_threads;
  kg_sample_kernel<<<(int)ceil_div64(threads, 128), 128, 0, stream>>>(ix, (const unsigned long long*)seeds_in,
                                                                    (unsigned long long*)seeds_out, work_threads, batch_size,
                                                                    neg_rate, mode, batch_h, batch_t, batch_r, batch_y);
  return check_launch("kg_sample_kernel");
}
