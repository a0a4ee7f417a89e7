// Element-wise binary focal loss and its derivative (reference ops.py:159-211).
//   loss = |1 - y - alpha| * (|y - x| + eps)^gamma * BCE(x, y)
// BCE follows torch.nn.functional.binary_cross_entropy: log terms clamped at -100, backward
// (x - y) / max((1 - x) x, 1e-12).
#include "common.cuh"
#include "skghoi_b200.h"

namespace skg {

__global__ void focal_loss_kernel(const float* __restrict__ x, const float* __restrict__ y, int64_t n, float alpha,
                                  float gamma, float eps, float* __restrict__ loss, const float* __restrict__ gout,
                                  float* __restrict__ dx) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float xv = x[i], yv = y[i];
    const float a = fabsf(1.f - yv - alpha);
    const float u = fabsf(yv - xv) + eps;
    const float bce = -(yv * fmaxf(logf(xv), -100.f) + (1.f - yv) * fmaxf(logf(1.f - xv), -100.f));
    const float ug = powf(u, gamma);
    if (loss) loss[i] = a * ug * bce;
    if (dx) {
      const float sgn = (xv > yv) ? 1.f : ((xv < yv) ? -1.f : 0.f);       // d|y-x|/dx
      const float dbce = (xv - yv) / fmaxf((1.f - xv) * xv, 1e-12f);
      const float dug = gamma * powf(u, gamma - 1.f) * sgn;
      dx[i] = gout[i] * a * (dug * bce + ug * dbce);
    }
  }
}

}  // namespace skg

using namespace skg;

extern "C" int skg_focal_loss(const float* x, const float* y, int64_t n, float alpha, float gamma, float eps,
                              float* loss, const float* grad_out, float* dx, cudaStream_t stream) {
  if (n == 0) return SKG_OK;
  SKG_REQUIRE(x && y && (loss || dx), "null pointer argument");
  SKG_REQUIRE(!dx || grad_out, "dx requires grad_out");
  const int grid = (int)(ceil_div64(n, 256) < (int64_t)num_sms() * 8 ? ceil_div64(n, 256) : (int64_t)num_sms() * 8);
  focal_loss_kernel<<<grid, 256, 0, stream>>>(x, y, n, alpha, gamma, eps, loss, grad_out, dx);
  return check_launch("focal_loss_kernel");
}
