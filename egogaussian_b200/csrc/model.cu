// model.cu -- the GaussianModel activations between the optimiser's raw parameters and the
// rasterizer inputs, forward and backward, as three small HBM-bound kernels (SURVEY.md 8a rows
// A-21/A-22): covariance + opacity per Gaussian, and the dc/rest feature concatenation.
//
// Replaces, per iteration, in the reference (compute_cov3D_python is forced on, train.py:49):
//   get_scaling / get_opacity / get_features          scene/gaussian_model.py:126-153
//   get_covariance / get_rotated_covariance           scene/gaussian_model.py:29-63,167-171
//   build_rotation / build_scaling_rotation / strip_symmetric   utils/general_utils.py:110-158
// i.e. ~25 torch kernels and their autograd graph, by one launch each way (plus the concat).
#include "common.cuh"
#include "model_math.cuh"

namespace egs {

constexpr int MODEL_THREADS = 256;

using ModelKernelArgs = RawModel;   // model_math.cuh: shared with the raw mode of the preprocess kernels

__global__ void __launch_bounds__(MODEL_THREADS)
model_forward_kernel(ModelKernelArgs a, float* __restrict__ cov3D, float* __restrict__ opacity) {
  __shared__ float s_rot[18];
  stage_rotations(a, s_rot);
  __syncthreads();
  const int i = blockIdx.x * MODEL_THREADS + threadIdx.x;
  if (i >= a.P) return;
  if (opacity != nullptr) opacity[i] = sigmoid_fwd(a.opacity_raw[i]);
  if (cov3D == nullptr) return;
  float c[6];
  raw_covariance(a, i, s_rot, c);
  float2* out = reinterpret_cast<float2*>(cov3D + 6 * (size_t)i);
  out[0] = make_float2(c[0], c[1]);
  out[1] = make_float2(c[2], c[3]);
  out[2] = make_float2(c[4], c[5]);
}

__global__ void __launch_bounds__(MODEL_THREADS)
model_backward_kernel(ModelKernelArgs a, const float* __restrict__ dL_dcov3D,
                      const float* __restrict__ dL_dopacity, float* __restrict__ d_scaling_raw,
                      float* __restrict__ d_rotation_raw, float* __restrict__ d_opacity_raw,
                      float* __restrict__ d_train_R) {
  __shared__ float s_rot[18];
  __shared__ float s_dT[MODEL_THREADS / 32][9];
  stage_rotations(a, s_rot);
  __syncthreads();
  const int i = blockIdx.x * MODEL_THREADS + threadIdx.x;
  const bool live = i < a.P;
  float dT[9];
#pragma unroll
  for (int k = 0; k < 9; ++k) dT[k] = 0.f;

  if (live && d_opacity_raw != nullptr) {
    const float y = sigmoid_fwd(a.opacity_raw[i]);
    d_opacity_raw[i] = sigmoid_bwd(y, dL_dopacity[i]);
  }
  if (live && dL_dcov3D != nullptr) {
    const float2* g2 = reinterpret_cast<const float2*>(dL_dcov3D + 6 * (size_t)i);
    const float2 g01 = g2[0], g23 = g2[1], g45 = g2[2];
    const float d6[6] = {g01.x, g01.y, g23.x, g23.y, g45.x, g45.y};
    float ds[3], dq[4];
    raw_covariance_backward(a, i, s_rot, d6, ds, dq, dT);
    d_scaling_raw[3 * i] = ds[0];
    d_scaling_raw[3 * i + 1] = ds[1];
    d_scaling_raw[3 * i + 2] = ds[2];
    reinterpret_cast<float4*>(d_rotation_raw)[i] = make_float4(dq[0], dq[1], dq[2], dq[3]);
  }

  // block-reduce the 3x3 pose gradient (every thread takes part)
  if (d_train_R != nullptr && a.train_R != nullptr) reduce_pose_gradient(dT, s_dT, d_train_R);
}

// get_features: cat(_features_dc [P,1,3], _features_rest [P,M-1,3]) along dim 1, and its backward
// (the split). One CTA moves FEAT_ROWS whole rows; both sides are contiguous runs, so the scalar
// accesses coalesce.
constexpr int FEAT_ROWS = 64;

template <bool SPLIT>
__global__ void __launch_bounds__(MODEL_THREADS)
features_kernel(int P, int M, float* __restrict__ dc, float* __restrict__ rest, float* __restrict__ full) {
  const int row0 = blockIdx.x * FEAT_ROWS;
  const int rows = min(FEAT_ROWS, P - row0);
  const int width = 3 * M, wrest = width - 3;
  float* f = full + (size_t)row0 * width;
  float* d = dc + (size_t)row0 * 3;
  float* r = rest + (size_t)row0 * wrest;
  for (int t = threadIdx.x; t < rows * width; t += MODEL_THREADS) {
    const int gi = t / width, k = t - gi * width;
    float* part = k < 3 ? d + gi * 3 + k : r + gi * wrest + (k - 3);
    if (SPLIT) *part = f[t];
    else f[t] = *part;
  }
}

// Per-iteration densification bookkeeping (scene/gaussian_model.py:735-740 and
// trainers/train_static_bg.py:128), for the rows with radii > 0:
//   max_radii2D = max(max_radii2D, radii); xyz_gradient_accum += |dL/dmean2D.xy|; denom += 1.
// The reference does this with boolean-mask indexing, i.e. several nonzero() host syncs per
// iteration; here it is one pass without any.
__global__ void __launch_bounds__(MODEL_THREADS)
densify_stats_kernel(int P, const int* __restrict__ radii, const float* __restrict__ dL_dmeans2D,
                     float* __restrict__ max_radii2D, float* __restrict__ grad_accum, float* __restrict__ denom) {
  const int i = blockIdx.x * MODEL_THREADS + threadIdx.x;
  if (i >= P) return;
  const int r = radii[i];
  if (r <= 0) return;
  max_radii2D[i] = fmaxf(max_radii2D[i], (float)r);
  const float gx = dL_dmeans2D[3 * i], gy = dL_dmeans2D[3 * i + 1];
  grad_accum[i] += sqrtf(gx * gx + gy * gy);
  denom[i] += 1.f;
}

static int check_model(const EgsModelArgs* m, const char* who) {
  if (m == nullptr || m->P < 0) {
    set_error("%s: NULL args or negative P", who);
    return EGS_ERR_INVALID_ARGUMENT;
  }
  if (m->row0_copies < 0) {
    set_error("%s: row0_copies must be >= 0", who);
    return EGS_ERR_INVALID_ARGUMENT;
  }
  return EGS_OK;
}

static ModelKernelArgs kernel_args(const EgsModelArgs* m) { return raw_model_from(m); }

static bool aligned_to(const void* p, uintptr_t a) { return (reinterpret_cast<uintptr_t>(p) & (a - 1)) == 0; }

}  // namespace egs

extern "C" int egs_model_forward(const EgsModelArgs* m, float* cov3D, float* opacity, float* shs, void* stream) {
  using namespace egs;
  if (int rc = check_model(m, "egs_model_forward")) return rc;
  const int P = m->P, M = m->sh_coeffs;
  if (P == 0) return EGS_OK;
  if ((cov3D && (!m->scaling_raw || !m->rotation_raw)) || (opacity && !m->opacity_raw) ||
      (shs && (M < 1 || !m->features_dc || (M > 1 && !m->features_rest)))) {
    set_error("egs_model_forward: a requested output lacks its raw input (or sh_coeffs < 1)");
    return EGS_ERR_INVALID_ARGUMENT;
  }
  if (cov3D && (!aligned_to(m->rotation_raw, 16) || !aligned_to(cov3D, 8))) {
    set_error("egs_model_forward: rotation_raw must be 16-byte and cov3D 8-byte aligned");
    return EGS_ERR_INVALID_ARGUMENT;
  }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  StageTimer timer(st, ST_MODEL_FWD);
  if (cov3D || opacity) {
    model_forward_kernel<<<(P + MODEL_THREADS - 1) / MODEL_THREADS, MODEL_THREADS, 0, st>>>(
        kernel_args(m), cov3D, opacity);
    EGS_AFTER_LAUNCH(st, false);
  }
  if (shs) {
    features_kernel<false><<<(P + FEAT_ROWS - 1) / FEAT_ROWS, MODEL_THREADS, 0, st>>>(
        P, M, const_cast<float*>(m->features_dc), const_cast<float*>(m->features_rest), shs);
    EGS_AFTER_LAUNCH(st, false);
  }
  return EGS_OK;
}

extern "C" int egs_model_backward(const EgsModelArgs* m, const float* dL_dcov3D, const float* dL_dopacity,
                                  const float* dL_dshs, float* d_scaling_raw, float* d_rotation_raw,
                                  float* d_opacity_raw, float* d_features_dc, float* d_features_rest,
                                  float* d_train_R, void* stream) {
  using namespace egs;
  if (int rc = check_model(m, "egs_model_backward")) return rc;
  const int P = m->P, M = m->sh_coeffs;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (d_train_R) EGS_CUDA_TRY(cudaMemsetAsync(d_train_R, 0, 9 * sizeof(float), st));
  if (P == 0) return EGS_OK;
  if ((dL_dcov3D && (!m->scaling_raw || !m->rotation_raw || !d_scaling_raw || !d_rotation_raw)) ||
      (dL_dopacity && (!m->opacity_raw || !d_opacity_raw)) ||
      (dL_dshs && (M < 1 || !d_features_dc || (M > 1 && !d_features_rest)))) {
    set_error("egs_model_backward: an upstream gradient lacks its raw input or its output");
    return EGS_ERR_INVALID_ARGUMENT;
  }
  if (dL_dcov3D && (!aligned_to(m->rotation_raw, 16) || !aligned_to(d_rotation_raw, 16) || !aligned_to(dL_dcov3D, 8))) {
    set_error("egs_model_backward: rotation tensors must be 16-byte and dL_dcov3D 8-byte aligned");
    return EGS_ERR_INVALID_ARGUMENT;
  }
  StageTimer timer(st, ST_MODEL_BWD);
  if (dL_dcov3D || dL_dopacity) {
    model_backward_kernel<<<(P + MODEL_THREADS - 1) / MODEL_THREADS, MODEL_THREADS, 0, st>>>(
        kernel_args(m), dL_dcov3D, dL_dopacity, d_scaling_raw, d_rotation_raw,
        dL_dopacity ? d_opacity_raw : nullptr, dL_dcov3D ? d_train_R : nullptr);
    EGS_AFTER_LAUNCH(st, false);
  }
  if (dL_dshs) {
    features_kernel<true><<<(P + FEAT_ROWS - 1) / FEAT_ROWS, MODEL_THREADS, 0, st>>>(
        P, M, d_features_dc, d_features_rest, const_cast<float*>(dL_dshs));
    EGS_AFTER_LAUNCH(st, false);
  }
  return EGS_OK;
}

extern "C" int egs_densify_stats(int P, const int* radii, const float* dL_dmeans2D, float* max_radii2D,
                                 float* xyz_gradient_accum, float* denom, void* stream) {
  using namespace egs;
  if (P < 0 || (P > 0 && (!radii || !dL_dmeans2D || !max_radii2D || !xyz_gradient_accum || !denom))) {
    set_error("egs_densify_stats: NULL argument or negative P");
    return EGS_ERR_INVALID_ARGUMENT;
  }
  if (P == 0) return EGS_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  densify_stats_kernel<<<(P + MODEL_THREADS - 1) / MODEL_THREADS, MODEL_THREADS, 0, st>>>(
      P, radii, dL_dmeans2D, max_radii2D, xyz_gradient_accum, denom);
  EGS_AFTER_LAUNCH(st, false);
  return EGS_OK;
}
