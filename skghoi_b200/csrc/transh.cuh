// Shared between the TransH forward (transh_fwd.cu) and backward (transh.cu) kernels.
#pragma once
#include "common.cuh"

namespace skg {

struct TripleIndex {
  // explicit mode: head/relation/tail arrays.  implicit (grid) mode (head.py:570-572):
  //   head = head_const, relation = t % rel_tot, tail = cell_tail ? cell_tail[t / rel_tot] : t / rel_tot
  const int64_t* head;
  const int64_t* relation;
  const int64_t* tail;
  const int64_t* cell_tail;
  int64_t head_const;
  int rel_tot;
  int implicit;
  __device__ __forceinline__ void get(int64_t t, int64_t& h, int64_t& r, int64_t& tl) const {
    if (implicit) {
      const int64_t c = t / rel_tot;
      h = head_const;
      r = t - c * rel_tot;
      tl = cell_tail ? __ldg(cell_tail + c) : c;
    } else {
      h = __ldg(head + t); r = __ldg(relation + t); tl = __ldg(tail + t);
    }
  }
};

}  // namespace skg
