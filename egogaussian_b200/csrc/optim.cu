// optim.cu -- fused Adam over all Gaussian parameter groups in one launch, and the rigid SE(3)
// object-pose transform EgoGaussian applies to object Gaussians.
//
// Adam follows torch.optim.Adam as the reference configures it (scene/gaussian_model.py:198:
// eps = 1e-15, default betas, bias correction, no weight decay, no amsgrad), i.e. per element
//   m <- m + (g - m) * (1 - b1);  v <- v * b2 + (1 - b2) * g * g
//   p <- p - (lr / (1 - b1^t)) * m / (sqrt(v) / sqrt(1 - b2^t) + eps)
// The reference runs this as ~6 foreach kernels per step over 28 B per parameter float
// (SURVEY.md 8a A-23); here it is one pass: read p, g, m, v and write p, m, v (28 B/float).
//
// SE(3): scene/gaussian_model.py:939-986 moves rows with is_object == k by x' = R x + t
// (utils/geometry_utils.py:188-193) and rotates their covariance factor (gaussian_model.py:46-63),
// then un-moves them after backward (:1037-1060). The fused step keeps parameters canonical and
// pulls the gradient back instead: g = R^T g' (SURVEY.md 8a A-24).
#include "common.cuh"
#include "adam_math.cuh"
#include <math.h>

namespace egs {

struct AdamGroupDev {
  float* p;
  const float* g;
  float* m;
  float* v;
  long long n;
  long long first_chunk;  // index of this group's first 4-float chunk in the virtual concatenation
  float neg_step_size;    // -(lr / (1 - b1^t))
  float inv_bc2_sqrt;     // 1 / sqrt(1 - b2^t)   (applied as a division, see below)
  float bc2_sqrt;
  int vec_ok;             // all four pointers 16-byte aligned
};

struct AdamLaunch {
  AdamGroupDev g[EGS_ADAM_MAX_GROUPS];
  int n_groups;
  long long total_chunks;
  float one_minus_b1, b2, one_minus_b2, eps;
};

__global__ void __launch_bounds__(256) adam_multi_kernel(const AdamLaunch L) {
  for (long long c = (long long)blockIdx.x * blockDim.x + threadIdx.x; c < L.total_chunks;
       c += (long long)gridDim.x * blockDim.x) {
    int gi = 0;
#pragma unroll 1
    for (int k = 1; k < L.n_groups; ++k)
      if (c >= L.g[k].first_chunk) gi = k;
    const AdamGroupDev& G = L.g[gi];
    const long long e = (c - G.first_chunk) * 4;
    if (G.vec_ok && e + 4 <= G.n) {
      float4 p = *reinterpret_cast<float4*>(G.p + e);
      const float4 g = *reinterpret_cast<const float4*>(G.g + e);
      float4 m = *reinterpret_cast<float4*>(G.m + e);
      float4 v = *reinterpret_cast<float4*>(G.v + e);
      adam_elem(p.x, g.x, m.x, v.x, L.one_minus_b1, L.b2, L.one_minus_b2, L.eps, G.neg_step_size, G.bc2_sqrt);
      adam_elem(p.y, g.y, m.y, v.y, L.one_minus_b1, L.b2, L.one_minus_b2, L.eps, G.neg_step_size, G.bc2_sqrt);
      adam_elem(p.z, g.z, m.z, v.z, L.one_minus_b1, L.b2, L.one_minus_b2, L.eps, G.neg_step_size, G.bc2_sqrt);
      adam_elem(p.w, g.w, m.w, v.w, L.one_minus_b1, L.b2, L.one_minus_b2, L.eps, G.neg_step_size, G.bc2_sqrt);
      *reinterpret_cast<float4*>(G.p + e) = p;
      *reinterpret_cast<float4*>(G.m + e) = m;
      *reinterpret_cast<float4*>(G.v + e) = v;
    } else {
      for (long long i = e; i < e + 4 && i < G.n; ++i) {
        float p = G.p[i], m = G.m[i], v = G.v[i];
        adam_elem(p, G.g[i], m, v, L.one_minus_b1, L.b2, L.one_minus_b2, L.eps, G.neg_step_size, G.bc2_sqrt);
        G.p[i] = p; G.m[i] = m; G.v[i] = v;
      }
    }
  }
}

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

static int adam_launch(const EgsAdamGroup* groups, int n_groups, double beta1, double beta2, double eps,
                       cudaStream_t stream) {
  if (n_groups < 0 || n_groups > EGS_ADAM_MAX_GROUPS || (n_groups > 0 && !groups)) {
    set_error("egs_adam_step: n_groups must be in [0, %d]", EGS_ADAM_MAX_GROUPS);
    return EGS_ERR_INVALID_ARGUMENT;
  }
  AdamLaunch L;
  L.n_groups = 0;
  long long chunks = 0;
  for (int k = 0; k < n_groups; ++k) {
    const EgsAdamGroup& s = groups[k];
    if (s.n == 0) continue;
    if (!s.param || !s.grad || !s.exp_avg || !s.exp_avg_sq || s.n < 0 || s.step < 1) {
      set_error("egs_adam_step: group %d has a NULL pointer, negative n or step < 1", k);
      return EGS_ERR_INVALID_ARGUMENT;
    }
    AdamGroupDev& d = L.g[L.n_groups++];
    d.p = s.param; d.g = s.grad; d.m = s.exp_avg; d.v = s.exp_avg_sq; d.n = s.n;
    d.first_chunk = chunks;
    chunks += (s.n + 3) / 4;
    // scalar prologue in double like torch's python-side bias corrections
    const double bc1 = 1.0 - pow(beta1, (double)s.step);
    const double bc2 = 1.0 - pow(beta2, (double)s.step);
    d.neg_step_size = (float)(-(s.lr / bc1));
    d.bc2_sqrt = (float)sqrt(bc2);
    d.inv_bc2_sqrt = (float)(1.0 / sqrt(bc2));
    d.vec_ok = aligned16(s.param) && aligned16(s.grad) && aligned16(s.exp_avg) && aligned16(s.exp_avg_sq);
  }
  if (L.n_groups == 0) return EGS_OK;
  L.total_chunks = chunks;
  L.one_minus_b1 = (float)(1.0 - beta1);
  L.b2 = (float)beta2;
  L.one_minus_b2 = (float)(1.0 - beta2);
  L.eps = (float)eps;
  long long blocks = (chunks + 255) / 256;
  const long long cap = 148LL * 16;
  if (blocks > cap) blocks = cap;
  StageTimer timer(stream, ST_ADAM);
  adam_multi_kernel<<<(unsigned)blocks, 256, 0, stream>>>(L);
  EGS_AFTER_LAUNCH(stream, false);
  return EGS_OK;
}

__global__ void __launch_bounds__(256)
se3_apply_kernel(int P, const float* __restrict__ xyz, const int* __restrict__ is_object, int which,
                 const float* __restrict__ Rt, float* __restrict__ xyz_out,
                 const float* __restrict__ cov_in, float* __restrict__ cov_out) {
  __shared__ float s[12];
  if (threadIdx.x < 12) s[threadIdx.x] = Rt[threadIdx.x];
  __syncthreads();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  const bool obj = is_object[i] == which;
  const float x = xyz[3 * i], y = xyz[3 * i + 1], z = xyz[3 * i + 2];
  if (obj) {
    xyz_out[3 * i + 0] = s[0] * x + s[1] * y + s[2] * z + s[9];
    xyz_out[3 * i + 1] = s[3] * x + s[4] * y + s[5] * z + s[10];
    xyz_out[3 * i + 2] = s[6] * x + s[7] * y + s[8] * z + s[11];
  } else {
    xyz_out[3 * i + 0] = x; xyz_out[3 * i + 1] = y; xyz_out[3 * i + 2] = z;
  }
  if (cov_in != nullptr) {
    float c[6];
#pragma unroll
    for (int k = 0; k < 6; ++k) c[k] = cov_in[6 * i + k];
    if (obj) {
      // Sigma' = R Sigma R^T with Sigma = [c0 c1 c2; c1 c3 c4; c2 c4 c5]
      float A[3][3];  // A = R * Sigma
#pragma unroll
      for (int r = 0; r < 3; ++r) {
        A[r][0] = s[3 * r] * c[0] + s[3 * r + 1] * c[1] + s[3 * r + 2] * c[2];
        A[r][1] = s[3 * r] * c[1] + s[3 * r + 1] * c[3] + s[3 * r + 2] * c[4];
        A[r][2] = s[3 * r] * c[2] + s[3 * r + 1] * c[4] + s[3 * r + 2] * c[5];
      }
      auto rr = [&](int r, int q) { return A[r][0] * s[3 * q] + A[r][1] * s[3 * q + 1] + A[r][2] * s[3 * q + 2]; };
      c[0] = rr(0, 0); c[1] = rr(0, 1); c[2] = rr(0, 2); c[3] = rr(1, 1); c[4] = rr(1, 2); c[5] = rr(2, 2);
    }
#pragma unroll
    for (int k = 0; k < 6; ++k) cov_out[6 * i + k] = c[k];
  }
}

__global__ void __launch_bounds__(256)
adam_se3_kernel(int P, float* __restrict__ xyz, const float* __restrict__ grad, float* __restrict__ m,
                float* __restrict__ v, const int* __restrict__ is_object, int which,
                const float* __restrict__ Rt, float omb1, float b2, float omb2, float eps,
                float neg_step, float bc2_sqrt) {
  __shared__ float s[9];
  if (threadIdx.x < 9) s[threadIdx.x] = Rt[threadIdx.x];
  __syncthreads();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  float g[3] = {grad[3 * i], grad[3 * i + 1], grad[3 * i + 2]};
  if (is_object[i] == which) {
    // R^T g
    const float gx = s[0] * g[0] + s[3] * g[1] + s[6] * g[2];
    const float gy = s[1] * g[0] + s[4] * g[1] + s[7] * g[2];
    const float gz = s[2] * g[0] + s[5] * g[1] + s[8] * g[2];
    g[0] = gx; g[1] = gy; g[2] = gz;
  }
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    float p = xyz[3 * i + k], mm = m[3 * i + k], vv = v[3 * i + k];
    adam_elem(p, g[k], mm, vv, omb1, b2, omb2, eps, neg_step, bc2_sqrt);
    xyz[3 * i + k] = p; m[3 * i + k] = mm; v[3 * i + k] = vv;
  }
}

}  // namespace egs

extern "C" int egs_adam_step(const EgsAdamGroup* groups, int n_groups, double beta1, double beta2,
                             double eps, void* stream) {
  return egs::adam_launch(groups, n_groups, beta1, beta2, eps, static_cast<cudaStream_t>(stream));
}

extern "C" int egs_se3_apply(int P, const float* xyz, const int* is_object, int which_object,
                             const float* Rt, float* xyz_out, const float* cov_in, float* cov_out,
                             void* stream) {
  if (P < 0 || (P > 0 && (!xyz || !is_object || !Rt || !xyz_out)) || (cov_in && !cov_out)) {
    egs::set_error("egs_se3_apply: NULL argument");
    return EGS_ERR_INVALID_ARGUMENT;
  }
  if (P == 0) return EGS_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  egs::StageTimer timer(st, egs::ST_SE3);
  egs::se3_apply_kernel<<<(P + 255) / 256, 256, 0, st>>>(P, xyz, is_object, which_object, Rt, xyz_out,
                                                         cov_in, cov_out);
  EGS_AFTER_LAUNCH(st, false);
  return EGS_OK;
}

extern "C" int egs_adam_se3_step(int P, float* xyz, const float* grad_moved, float* exp_avg,
                                 float* exp_avg_sq, const int* is_object, int which_object,
                                 const float* Rt, double lr, int step, double beta1, double beta2,
                                 double eps, void* stream) {
  if (P < 0 || step < 1 || (P > 0 && (!xyz || !grad_moved || !exp_avg || !exp_avg_sq || !is_object || !Rt))) {
    egs::set_error("egs_adam_se3_step: NULL argument or step < 1");
    return EGS_ERR_INVALID_ARGUMENT;
  }
  if (P == 0) return EGS_OK;
  const double bc1 = 1.0 - pow(beta1, (double)step);
  const double bc2 = 1.0 - pow(beta2, (double)step);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  egs::StageTimer timer(st, egs::ST_ADAM);
  egs::adam_se3_kernel<<<(P + 255) / 256, 256, 0, st>>>(
      P, xyz, grad_moved, exp_avg, exp_avg_sq, is_object, which_object, Rt,
      (float)(1.0 - beta1), (float)beta2, (float)(1.0 - beta2), (float)eps,
      (float)(-(lr / bc1)), (float)sqrt(bc2));
  EGS_AFTER_LAUNCH(st, false);
  return EGS_OK;
}
