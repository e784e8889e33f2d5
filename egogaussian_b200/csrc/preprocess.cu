// preprocess.cu -- per-Gaussian forward stage: near culling, projection, EWA covariance,
// conic + radius, tile rectangle, SH -> RGB. One thread per Gaussian.
//
// Semantics follow the reference's preprocessCUDA (DGR/cuda_rasterizer/forward.cu:155-256 with
// helpers :20-152 and auxiliary.h:41-76,139-164) as restated in SURVEY.md App. A.1-A.4. Radii and
// sort keys must be bit-exact, so the covariance chain keeps the reference's expression shapes
// (column-major products summed left to right, double-precision ndc->pixel) and the file is
// compiled with nvcc's default -fmad=true and no fast-math, like the reference.
#include "common.cuh"
#include "model_math.cuh"

namespace egs {

// Sigma = (S R)^T (S R) from scale (already multiplied by the modifier inside) and the raw,
// un-normalised quaternion (w,x,y,z)  [forward.cu:118-152].
__device__ __forceinline__ void cov3d_from_scale_rot(const float3 s, float mod, const float4 q,
                                                     float* cov6) {
  Mat3 S = mat3_cols(1.f, 0.f, 0.f, 0.f, 1.f, 0.f, 0.f, 0.f, 1.f);
  S.c[0][0] = mod * s.x;
  S.c[1][1] = mod * s.y;
  S.c[2][2] = mod * s.z;
  const float r = q.x, x = q.y, y = q.z, z = q.w;
  const Mat3 R = mat3_cols(1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y),
                           2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x),
                           2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y));
  const Mat3 M = mul(S, R);
  const Mat3 Sigma = mul(transpose(M), M);
  cov6[0] = Sigma.c[0][0];
  cov6[1] = Sigma.c[0][1];
  cov6[2] = Sigma.c[0][2];
  cov6[3] = Sigma.c[1][1];
  cov6[4] = Sigma.c[1][2];
  cov6[5] = Sigma.c[2][2];
}

// 2D screen-space covariance (EWA, Zwicker et al. 2002 eq. 29/31) + 0.3 px low-pass
// [forward.cu:74-113].
__device__ __forceinline__ float3 cov2d_ewa(const float3& mean, float fx, float fy, float tanx,
                                            float tany, const float* cov6, const float* view) {
  float3 t = xform_point_4x3(mean, view);
  const float limx = 1.3f * tanx;
  const float limy = 1.3f * tany;
  const float txtz = t.x / t.z;
  const float tytz = t.y / t.z;
  t.x = min(limx, max(-limx, txtz)) * t.z;
  t.y = min(limy, max(-limy, tytz)) * t.z;

  const Mat3 J = mat3_cols(fx / t.z, 0.0f, -(fx * t.x) / (t.z * t.z),
                           0.0f, fy / t.z, -(fy * t.y) / (t.z * t.z),
                           0.f, 0.f, 0.f);
  const Mat3 Wm = mat3_cols(view[0], view[4], view[8], view[1], view[5], view[9], view[2], view[6],
                            view[10]);
  const Mat3 T = mul(Wm, J);
  const Mat3 V = mat3_cols(cov6[0], cov6[1], cov6[2], cov6[1], cov6[3], cov6[4], cov6[2], cov6[4],
                           cov6[5]);
  Mat3 cov = mul(mul(transpose(T), transpose(V)), T);
  cov.c[0][0] += 0.3f;
  cov.c[1][1] += 0.3f;
  return make_float3(cov.c[0][0], cov.c[0][1], cov.c[1][1]);
}

// SH degree 0..3 -> RGB, +0.5, clamp at 0 with per-channel clamp flags [forward.cu:20-71].
__device__ __forceinline__ float3 sh_to_rgb(int deg, int M, const float3 pos, const float3 cam,
                                            const float* sh, unsigned char& clamp_bits) {
  float3 dir = make_float3(pos.x - cam.x, pos.y - cam.y, pos.z - cam.z);
  const float len = sqrtf(dir.x * dir.x + dir.y * dir.y + dir.z * dir.z);
  dir.x = dir.x / len;
  dir.y = dir.y / len;
  dir.z = dir.z / len;
#define SH(k, ch) sh[3 * (k) + (ch)]
  float res[3];
#pragma unroll
  for (int ch = 0; ch < 3; ++ch) res[ch] = kSH0 * SH(0, ch);
  if (deg > 0) {
    const float x = dir.x, y = dir.y, z = dir.z;
#pragma unroll
    for (int ch = 0; ch < 3; ++ch)
      res[ch] = res[ch] - kSH1 * y * SH(1, ch) + kSH1 * z * SH(2, ch) - kSH1 * x * SH(3, ch);
    if (deg > 1) {
      const float xx = x * x, yy = y * y, zz = z * z;
      const float xy = x * y, yz = y * z, xz = x * z;
#pragma unroll
      for (int ch = 0; ch < 3; ++ch)
        res[ch] = res[ch] + kSH2[0] * xy * SH(4, ch) + kSH2[1] * yz * SH(5, ch) +
                  kSH2[2] * (2.0f * zz - xx - yy) * SH(6, ch) + kSH2[3] * xz * SH(7, ch) +
                  kSH2[4] * (xx - yy) * SH(8, ch);
      if (deg > 2) {
#pragma unroll
        for (int ch = 0; ch < 3; ++ch)
          res[ch] = res[ch] + kSH3[0] * y * (3.0f * xx - yy) * SH(9, ch) +
                    kSH3[1] * xy * z * SH(10, ch) + kSH3[2] * y * (4.0f * zz - xx - yy) * SH(11, ch) +
                    kSH3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * SH(12, ch) +
                    kSH3[4] * x * (4.0f * zz - xx - yy) * SH(13, ch) +
                    kSH3[5] * z * (xx - yy) * SH(14, ch) + kSH3[6] * x * (xx - 3.0f * yy) * SH(15, ch);
      }
    }
  }
#undef SH
  clamp_bits = 0;
#pragma unroll
  for (int ch = 0; ch < 3; ++ch) {
    res[ch] += 0.5f;
    if (res[ch] < 0.f) clamp_bits |= (1u << ch);
    res[ch] = fmaxf(res[ch], 0.0f);
  }
  return make_float3(res[0], res[1], res[2]);
}

// 128-thread blocks. With 16 coefficients per Gaussian the SH rows of a warp's 32 Gaussians (6 KB,
// contiguous) are staged through shared memory with float4 loads; a thread walking its own
// 192-byte row in global memory made the kernel L1-bound (ncu: l1tex 86 %, dram 29 %).
constexpr int PF_THREADS = 128;
constexpr int PF_SH_ROW = 48;
// pitch of a staged SH row: 52 floats (16-byte aligned rows, read with 16-byte shared loads) for the
// contiguous tensor, 49 (odd: conflict-free scalar reads) for the split dc / rest rows, whose pieces
// start at 12- and 180-byte offsets
constexpr int PF_SH_PITCH = 52;
constexpr int PF_SH_PITCH_SPLIT = 49;

// SH_MODE: 0 = every thread reads its own row from global memory, 1 = rows staged with float4
// loads from one contiguous [P,16,3] tensor, 2 = rows staged from the split features_dc /
// features_rest tensors of the raw-parameter mode (any M; get_features' concat never materialises).
// `raw` carries the optimiser's raw tensors: when raw.scaling_raw / raw.opacity_raw are set the
// covariance and the opacity are derived here (model_math.cuh) instead of being read.
// RAW is a template parameter so that the ordinary instantiations carry none of that code.
template <int SH_MODE, bool RAW>
__global__ void __launch_bounds__(PF_THREADS)
preprocess_forward_kernel(RawModel raw, int P, int D, int M, const float* __restrict__ means3D,
                          const float* __restrict__ scales, float scale_modifier,
                          const float* __restrict__ rotations, const float* __restrict__ opacities,
                          const float* __restrict__ shs, const float* __restrict__ cov3D_precomp,
                          const float* __restrict__ colors_precomp,
                          const float* __restrict__ viewmatrix, const float* __restrict__ projmatrix,
                          const float* __restrict__ campos, int W, int H, float tan_fovx,
                          float tan_fovy, float focal_x, float focal_y, int grid_x, int grid_y,
                          int* __restrict__ radii, float2* __restrict__ means2D,
                          float* __restrict__ depths, float4* __restrict__ conic_opacity,
                          float4* __restrict__ rgb_depth, unsigned char* __restrict__ clamped,
                          uint32_t* __restrict__ tiles_touched, int prefiltered,
                          uint32_t* __restrict__ prefilter_fault) {
  __shared__ float s_view[16], s_proj[16], s_cam[3];
  if (threadIdx.x < 16) {
    s_view[threadIdx.x] = viewmatrix[threadIdx.x];
    s_proj[threadIdx.x] = projmatrix[threadIdx.x];
  }
  if (threadIdx.x < 3) s_cam[threadIdx.x] = campos[threadIdx.x];
  __shared__ __align__(16) float s_sh[SH_MODE != 0 ? (PF_THREADS / 32) * 32 * PF_SH_PITCH : 4];
  __shared__ float s_rot[RAW ? 18 : 1];
  if (RAW) stage_rotations(raw, s_rot);
  __syncthreads();

  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  const float* sh_row = nullptr;

  // Issue this Gaussian's own loads before the SH rows are staged, so that they are in flight
  // together with the 6 KB the warp is about to pull in (ncu: the kernel waits on memory, not on
  // issue slots). g[] holds raw scale + quaternion, scale + quaternion, or the six covariance entries.
  float3 p_orig = make_float3(0.f, 0.f, 0.f);
  float g[7] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  float opacity_in = 0.f;
  if (idx < P) {
    p_orig = make_float3(means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2]);
    if (RAW || cov3D_precomp == nullptr) {
      const float* sc = RAW ? raw.scaling_raw : scales;
      const float4 q = reinterpret_cast<const float4*>(RAW ? raw.rotation_raw : rotations)[idx];
      g[0] = sc[3 * idx]; g[1] = sc[3 * idx + 1]; g[2] = sc[3 * idx + 2];
      g[3] = q.x; g[4] = q.y; g[5] = q.z; g[6] = q.w;
    } else {
#pragma unroll
      for (int k = 0; k < 6; ++k) g[k] = cov3D_precomp[6 * idx + k];
    }
    opacity_in = RAW ? raw.opacity_raw[idx] : opacities[idx];
  }
  if (SH_MODE == 2) {
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int first = blockIdx.x * blockDim.x + warp * 32;
    const int rows = min(32, P - first);
    float* tile = s_sh + warp * 32 * PF_SH_PITCH;
    const int pitch = M == 16 ? PF_SH_PITCH : PF_SH_PITCH_SPLIT;   // 16 coefficients: aligned rows, vector reads
    if (rows > 0) {
      const int wrest = 3 * M - 3;
      stage_rows_async(tile, pitch, 0, 3, raw.features_dc + (size_t)first * 3, rows * 3, lane);
      if (wrest > 0)
        stage_rows_async(tile, pitch, 3, wrest, raw.features_rest + (size_t)first * wrest, rows * wrest, lane);
    }
    stage_rows_wait();
    sh_row = tile + lane * pitch;
  } else if (SH_MODE == 1) {
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int first = blockIdx.x * blockDim.x + warp * 32;
    const int rows = min(32, P - first);
    float* tile = s_sh + warp * 32 * PF_SH_PITCH;
    if (rows > 0) {
      stage_rows_async16(tile, PF_SH_PITCH, PF_SH_ROW, shs + (size_t)first * PF_SH_ROW, rows * PF_SH_ROW, lane);
    }
    stage_rows_wait();
    sh_row = tile + lane * PF_SH_PITCH;
  } else if (shs != nullptr && idx < P) {
    sh_row = shs + (size_t)idx * M * 3;
  }
  if (idx >= P) return;

  int radius_out = 0;
  uint32_t tiles_out = 0;

  const float3 p_view = xform_point_4x3(p_orig, s_view);

  // near culling: z_view <= 0.2 (auxiliary.h:150); the x/y frustum test is disabled there too
  if (p_view.z <= 0.2f) {
    // the reference prints and traps here (auxiliary.h:154-161); forward reads the flag back with the
    // instance count and returns EGS_ERR_PREFILTERED
    if (prefiltered) *prefilter_fault = 1u;
  } else {
    const float4 p_hom = xform_point_4x4(p_orig, s_proj);
    const float p_w = 1.0f / (p_hom.w + 0.0000001f);
    const float3 p_proj = make_float3(p_hom.x * p_w, p_hom.y * p_w, p_hom.z * p_w);

    float cov6[6];
    if (RAW) {
      raw_covariance_from(raw, idx, g, make_float4(g[3], g[4], g[5], g[6]), s_rot, cov6);
    } else if (cov3D_precomp != nullptr) {
#pragma unroll
      for (int k = 0; k < 6; ++k) cov6[k] = g[k];
    } else {
      cov3d_from_scale_rot(make_float3(g[0], g[1], g[2]), scale_modifier, make_float4(g[3], g[4], g[5], g[6]), cov6);
    }

    const float3 cov = cov2d_ewa(p_orig, focal_x, focal_y, tan_fovx, tan_fovy, cov6, s_view);

    const float det = (cov.x * cov.z - cov.y * cov.y);
    if (det != 0.0f) {
      const float det_inv = 1.f / det;
      const float3 conic = make_float3(cov.z * det_inv, -cov.y * det_inv, cov.x * det_inv);

      const float mid = 0.5f * (cov.x + cov.z);
      const float lambda1 = mid + sqrtf(max(0.1f, mid * mid - det));
      const float lambda2 = mid - sqrtf(max(0.1f, mid * mid - det));
      const float my_radius = ceilf(3.f * sqrtf(max(lambda1, lambda2)));
      const float2 point_image = make_float2(ndc_to_pix(p_proj.x, W), ndc_to_pix(p_proj.y, H));
      uint2 lo, hi;
      tile_rect(point_image, (int)my_radius, grid_x, grid_y, lo, hi);
      const uint32_t n_tiles = (hi.x - lo.x) * (hi.y - lo.y);
      if (n_tiles != 0) {
        float3 rgb;
        unsigned char cb = 0;
        if (colors_precomp == nullptr) {
          if (SH_MODE == 1 || (SH_MODE == 2 && M == 16)) {
            // the staged row is 16-byte aligned: twelve 16-byte shared loads instead of 48 scalar ones
            // (the kernel waited on the load/store queue, ncu mio_throttle); the values then sit in
            // registers and the evaluation keeps the reference's expression shapes
            const float4* r4 = reinterpret_cast<const float4*>(__builtin_assume_aligned(sh_row, 16));
            float v[PF_SH_ROW];
#pragma unroll
            for (int c = 0; c < PF_SH_ROW / 4; ++c) {
              const float4 t = r4[c];
              v[4 * c] = t.x; v[4 * c + 1] = t.y; v[4 * c + 2] = t.z; v[4 * c + 3] = t.w;
            }
            rgb = sh_to_rgb(D, M, p_orig, make_float3(s_cam[0], s_cam[1], s_cam[2]), v, cb);
          } else {
            rgb = sh_to_rgb(D, M, p_orig, make_float3(s_cam[0], s_cam[1], s_cam[2]), sh_row, cb);
          }
          clamped[idx] = cb;
        } else {
          rgb = make_float3(colors_precomp[3 * idx], colors_precomp[3 * idx + 1],
                            colors_precomp[3 * idx + 2]);
        }
        depths[idx] = p_view.z;
        means2D[idx] = point_image;
        const float opacity = RAW ? sigmoid_fwd(opacity_in) : opacity_in;
        conic_opacity[idx] = make_float4(conic.x, conic.y, conic.z, opacity);
        rgb_depth[idx] = make_float4(rgb.x, rgb.y, rgb.z, p_view.z);
        radius_out = (int)my_radius;
        tiles_out = n_tiles;
      }
    }
  }
  radii[idx] = radius_out;
  tiles_touched[idx] = tiles_out;
}

__global__ void __launch_bounds__(256)
mark_visible_kernel(int P, const float* __restrict__ means3D, const float* __restrict__ view,
                    unsigned char* __restrict__ present) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= P) return;
  const float3 p = make_float3(means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2]);
  const float3 pv = xform_point_4x3(p, view);
  // the reference culls with `z <= 0.2` (auxiliary.h:150), so a NaN depth counts as visible
  present[idx] = (pv.z <= 0.2f) ? 0 : 1;
}

int launch_preprocess_forward(const EgsForwardArgs& a, const EgsModelArgs* model, const GeomPtrs& g,
                              cudaStream_t stream) {
  const int P = a.P;
  const float focal_y = a.height / (2.0f * a.tan_fovy);
  const float focal_x = a.width / (2.0f * a.tan_fovx);
  const int blocks = (P + PF_THREADS - 1) / PF_THREADS;
  StageTimer timer(stream, ST_PREPROCESS);
  const bool staged = a.shs != nullptr && a.sh_coeffs == 16 && (reinterpret_cast<uintptr_t>(a.shs) & 15u) == 0;
  RawModel raw{};
  if (model != nullptr) raw = raw_model_from(model);
  const bool split = model != nullptr && a.colors_precomp == nullptr;
  const int sh_coeffs = split ? model->sh_coeffs : a.sh_coeffs;
#define EGS_PF_ARGS                                                                                     \
  raw, P, a.sh_degree, sh_coeffs, a.means3D, a.scales, a.scale_modifier, a.rotations, a.opacities, a.shs,  \
      a.cov3D_precomp, a.colors_precomp, a.viewmatrix, a.projmatrix, a.campos, a.width, a.height,       \
      a.tan_fovx, a.tan_fovy, focal_x, focal_y, tiles_x(a.width), tiles_y(a.height), a.radii,           \
      g.means2D, g.depths, g.conic_opacity, g.rgb_depth, g.clamped, g.tiles_touched, a.prefiltered,     \
      g.prefilter_fault
  if (split)
    preprocess_forward_kernel<2, true><<<blocks, PF_THREADS, 0, stream>>>(EGS_PF_ARGS);
  else if (model != nullptr)
    preprocess_forward_kernel<0, true><<<blocks, PF_THREADS, 0, stream>>>(EGS_PF_ARGS);
  else if (staged)
    preprocess_forward_kernel<1, false><<<blocks, PF_THREADS, 0, stream>>>(EGS_PF_ARGS);
  else
    preprocess_forward_kernel<0, false><<<blocks, PF_THREADS, 0, stream>>>(EGS_PF_ARGS);
#undef EGS_PF_ARGS
  EGS_AFTER_LAUNCH(stream, a.debug);
  return EGS_OK;
}

int launch_mark_visible(int P, const float* means3D, const float* view, const float* proj,
                        unsigned char* present, cudaStream_t stream) {
  (void)proj;  // the reference computes the projection too but only tests view-space z
  mark_visible_kernel<<<(P + 255) / 256, 256, 0, stream>>>(P, means3D, view, present);
  EGS_AFTER_LAUNCH(stream, false);
  return EGS_OK;
}

}  // namespace egs
