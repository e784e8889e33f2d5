// adam_math.cuh -- the per-element Adam update shared by optim.cu and exchange.cu (torch.optim.Adam as
// the reference configures it, scene/gaussian_model.py:198; see optim.cu for the derivation).
#pragma once

namespace egs {

__device__ __forceinline__ void adam_elem(float& p, float g, float& m, float& v, float omb1,
                                          float b2, float omb2, float eps, float neg_step,
                                          float bc2_sqrt) {
  m = m + (g - m) * omb1;
  v = v * b2 + omb2 * g * g;
  const float denom = sqrtf(v) / bc2_sqrt + eps;
  p = p + neg_step * (m / denom);
}

}  // namespace egs
