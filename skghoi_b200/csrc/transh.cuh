// Shared between the TransH forward (transh_fwd.cu), backward (transh.cu) and fused KG-training (kgtrain.cu) kernels.
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

struct TranshTriple {   // per-lane slice of one triple's intermediate state
  float inv_w, a_h, a_t, inv_h, inv_t, inv_r, nrm_s;
};

// Forward math for one triple held EPL-elements-per-lane.  Returns the score (all lanes).
template <int EPL>
__device__ __forceinline__ float transh_triple_fwd(const float (&h)[EPL], const float (&t)[EPL], const float (&r)[EPL],
                                                   const float (&w)[EPL], int p_norm, int norm_flag, float (&s)[EPL],
                                                   TranshTriple& st) {
  float ww = 0.f, hw = 0.f, tw = 0.f, rr = 0.f;
#pragma unroll
  for (int e = 0; e < EPL; ++e) { ww += w[e] * w[e]; hw += h[e] * w[e]; tw += t[e] * w[e]; rr += r[e] * r[e]; }
  ww = warp_sum(ww); hw = warp_sum(hw); tw = warp_sum(tw); rr = warp_sum(rr);
  const float inv_w = 1.f / fmaxf(sqrtf(ww), 1e-12f);      // F.normalize(norm, 2, -1)  TransH.py:76
  const float a_h = hw * inv_w, a_t = tw * inv_w;            // sum(e * w_hat)           TransH.py:85
  float hp[EPL], tp[EPL];
  float hh = 0.f, tt = 0.f;
#pragma unroll
  for (int e = 0; e < EPL; ++e) {
    const float wn = w[e] * inv_w;
    hp[e] = h[e] - a_h * wn; tp[e] = t[e] - a_t * wn;
    hh += hp[e] * hp[e]; tt += tp[e] * tp[e];
  }
  float inv_h = 1.f, inv_t = 1.f, inv_r = 1.f;
  if (norm_flag) {                                           // TransH.py:57-60
    hh = warp_sum(hh); tt = warp_sum(tt);
    inv_h = 1.f / fmaxf(sqrtf(hh), 1e-12f); inv_t = 1.f / fmaxf(sqrtf(tt), 1e-12f);
    inv_r = 1.f / fmaxf(sqrtf(rr), 1e-12f);
  }
  float acc = 0.f;
#pragma unroll
  for (int e = 0; e < EPL; ++e) {
    s[e] = (hp[e] * inv_h + r[e] * inv_r) - tp[e] * inv_t;   // (h + r) - t   TransH.py:68
    acc += (p_norm == 1) ? fabsf(s[e]) : s[e] * s[e];
  }
  acc = warp_sum(acc);
  const float score = (p_norm == 1) ? acc : sqrtf(acc);      // torch.norm(score, p, -1)  TransH.py:70
  st.inv_w = inv_w; st.a_h = a_h; st.a_t = a_t; st.inv_h = inv_h; st.inv_t = inv_t; st.inv_r = inv_r; st.nrm_s = score;
  return score;
}

// Gradient of one triple's score w.r.t. its four table rows, given g = d loss / d score and the forward state `st`
// (s = the (h + r) - t vector of transh_triple_fwd).  All lanes of the warp take part.
template <int EPL>
__device__ __forceinline__ void transh_triple_bwd(const float (&h)[EPL], const float (&tl)[EPL], const float (&r)[EPL],
                                                  const float (&w)[EPL], const float (&s)[EPL], const TranshTriple& st, float g,
                                                  int p_norm, int norm_flag, float (&g_h)[EPL], float (&g_t)[EPL],
                                                  float (&g_rr)[EPL], float (&g_w)[EPL]) {
    // g_s = g * d||s||_p / ds   (zero sub-gradient at s == 0, as torch.norm)
    float gs[EPL];
#pragma unroll
    for (int e = 0; e < EPL; ++e) {
      if (p_norm == 1) gs[e] = g * ((s[e] > 0.f) - (s[e] < 0.f));
      else gs[e] = st.nrm_s > 0.f ? g * s[e] / st.nrm_s : 0.f;
    }
    // unit vectors u = hp*inv_h, q = r*inv_r, v = tp*inv_t ;  s = u + q - v
    float wn[EPL], hp[EPL], tp[EPL];
    float ugs = 0.f, vgs = 0.f, qgs = 0.f;
#pragma unroll
    for (int e = 0; e < EPL; ++e) {
      wn[e] = w[e] * st.inv_w;
      hp[e] = h[e] - st.a_h * wn[e]; tp[e] = tl[e] - st.a_t * wn[e];
      ugs += hp[e] * st.inv_h * gs[e]; vgs += tp[e] * st.inv_t * gs[e]; qgs += r[e] * st.inv_r * gs[e];
    }
    float g_hp[EPL], g_tp[EPL], g_r_[EPL];
    if (norm_flag) {
      ugs = warp_sum(ugs); vgs = warp_sum(vgs); qgs = warp_sum(qgs);
#pragma unroll
      for (int e = 0; e < EPL; ++e) {
        g_hp[e] = (gs[e] - hp[e] * st.inv_h * ugs) * st.inv_h;
        g_tp[e] = -(gs[e] - tp[e] * st.inv_t * vgs) * st.inv_t;
        g_r_[e] = (gs[e] - r[e] * st.inv_r * qgs) * st.inv_r;
      }
    } else {
#pragma unroll
      for (int e = 0; e < EPL; ++e) { g_hp[e] = gs[e]; g_tp[e] = -gs[e]; g_r_[e] = gs[e]; }
    }
    // hp = h - (h.wn) wn  ->  g_h = g_hp - (g_hp.wn) wn ;  g_wn += -(a_h g_hp + (g_hp.wn) h)   (same for t)
    float bh = 0.f, bt = 0.f;
#pragma unroll
    for (int e = 0; e < EPL; ++e) { bh += g_hp[e] * wn[e]; bt += g_tp[e] * wn[e]; }
    bh = warp_sum(bh); bt = warp_sum(bt);
    float g_wn[EPL];
    float wg = 0.f;
#pragma unroll
    for (int e = 0; e < EPL; ++e) {
      g_wn[e] = -(st.a_h * g_hp[e] + bh * h[e]) - (st.a_t * g_tp[e] + bt * tl[e]);
      wg += wn[e] * g_wn[e];
    }
    wg = warp_sum(wg);
#pragma unroll
  for (int e = 0; e < EPL; ++e) {
    g_h[e] = g_hp[e] - bh * wn[e];
    g_t[e] = g_tp[e] - bt * wn[e];
    g_rr[e] = g_r_[e];
    g_w[e] = (g_wn[e] - wn[e] * wg) * st.inv_w;   // w_hat = w / ||w||
  }
}

}  // namespace skg
