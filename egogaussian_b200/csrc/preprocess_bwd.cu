// preprocess_bwd.cu -- per-Gaussian backward stage, one fused kernel.
//
// Replaces the reference's two kernels computeCov2DCUDA (DGR/cuda_rasterizer/backward.cu:144-274)
// and backward preprocessCUDA (:346-412, with the SH backward :20-139 and the scale/rotation
// backward :278-341), as restated in SURVEY.md App. A.8b / A.8c. Fusing them keeps dL/dcov3D in
// registers in the scale+rotation mode and sums the four parts of dL/dmean3D in registers in the
// reference's order (covariance part, projection part, depth part, SH view-direction part).
//
// The kernel also unpacks the 48-byte accumulator record written by the tile kernel into the
// caller-facing dL/dmeans2D, dL/dopacity and dL/dcolors tensors, and writes zeros for culled
// Gaussians, so none of the outputs needs a prior memset (the reference zero-fills ten tensors,
// 304 B per Gaussian at SH degree 3, rasterize_points.cu:158-167).
#include "common.cuh"
#include "model_math.cuh"

namespace egs {

__device__ __forceinline__ float3 xform_vec_4x3_T(const float3& p, const float* m) {
  float3 r;
  r.x = m[0] * p.x + m[1] * p.y + m[2] * p.z;
  r.y = m[4] * p.x + m[5] * p.y + m[6] * p.z;
  r.z = m[8] * p.x + m[9] * p.y + m[10] * p.z;
  return r;
}

// d(v/|v|)/dv applied to dv (auxiliary.h:107-117)
__device__ __forceinline__ float3 dnormalize(const float3 v, const float3 dv) {
  const float sum2 = v.x * v.x + v.y * v.y + v.z * v.z;
  const float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
  float3 r;
  r.x = ((+sum2 - v.x * v.x) * dv.x - v.y * v.x * dv.y - v.z * v.x * dv.z) * invsum32;
  r.y = (-v.x * v.y * dv.x + (sum2 - v.y * v.y) * dv.y - v.z * v.y * dv.z) * invsum32;
  r.z = (-v.x * v.z * dv.x - v.y * v.z * dv.y + (sum2 - v.z * v.z) * dv.z) * invsum32;
  return r;
}

struct F3 {
  float x, y, z;
};
__device__ __forceinline__ F3 operator*(float s, const F3& a) { return {s * a.x, s * a.y, s * a.z}; }
__device__ __forceinline__ F3 operator*(const F3& a, float s) { return {a.x * s, a.y * s, a.z * s}; }
__device__ __forceinline__ F3 operator+(const F3& a, const F3& b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ F3 operator-(const F3& a) { return {-a.x, -a.y, -a.z}; }
__device__ __forceinline__ float dot(const F3& a, const F3& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }

// SH backward: writes dL/dsh[0 .. M) (zeros above the active degree) and returns the part of
// dL/dmean caused by the view direction. All coefficients are read before any gradient is written, so
// `shp` and `dL_dsh_out` may be the same row (the staged kernel works in place in shared memory).
// Basis functions and their derivatives are those of backward.cu:20-139, grouped per coefficient.
// VEC: the row holds 16 coefficients and is 16-byte aligned (staged [P,16,3] rows, pitch 52): twelve
// 16-byte shared loads and stores instead of 48 scalar ones each way.
// dL_dsh_out == NULL (factored mode): the rows are not formed; the caller keeps the clamp-masked
// dL/dRGB (returned through dL_dRGB) and whoever needs dL/dsh multiplies it by the basis values.
template <bool VEC>
__device__ __forceinline__ float3 sh_backward(int deg, int M, const float3 pos, const float3 cam,
                                              const float* shp, unsigned char clamp_bits, F3& dL_dRGB,
                                              float* dL_dsh_out) {
  const float3 dir_orig = make_float3(pos.x - cam.x, pos.y - cam.y, pos.z - cam.z);
  const float len = sqrtf(dir_orig.x * dir_orig.x + dir_orig.y * dir_orig.y + dir_orig.z * dir_orig.z);
  const float x = dir_orig.x / len, y = dir_orig.y / len, z = dir_orig.z / len;
  dL_dRGB.x *= (clamp_bits & 1) ? 0.f : 1.f;
  dL_dRGB.y *= (clamp_bits & 2) ? 0.f : 1.f;
  dL_dRGB.z *= (clamp_bits & 4) ? 0.f : 1.f;

  const int n_act = (deg + 1) * (deg + 1);   // coefficients of the active degree
  float cf[48];
  if (VEC) {
    const float4* r4 = reinterpret_cast<const float4*>(__builtin_assume_aligned(shp, 16));
#pragma unroll
    for (int c = 0; c < 12; ++c) {
      const float4 t = r4[c];
      cf[4 * c] = t.x; cf[4 * c + 1] = t.y; cf[4 * c + 2] = t.z; cf[4 * c + 3] = t.w;
    }
  } else {
#pragma unroll
    for (int k = 0; k < 16; ++k)
      if (k < n_act) {
        cf[3 * k] = shp[3 * k]; cf[3 * k + 1] = shp[3 * k + 1]; cf[3 * k + 2] = shp[3 * k + 2];
      }
  }
  F3 dRGBdx = {0, 0, 0}, dRGBdy = {0, 0, 0}, dRGBdz = {0, 0, 0};
  float basis[16];
#pragma unroll
  for (int k = 0; k < 16; ++k) basis[k] = 0.f;
  sh_basis_values(deg, x, y, z, basis);
  // one coefficient: the derivatives (bx, by, bz) of its basis function with respect to the direction
  auto term = [&](int k, float bx, float by, float bz) {
    const F3 c = {cf[3 * k], cf[3 * k + 1], cf[3 * k + 2]};
    dRGBdx = dRGBdx + bx * c;
    dRGBdy = dRGBdy + by * c;
    dRGBdz = dRGBdz + bz * c;
  };
  if (deg > 0) {
    term(1, 0.f, -kSH1, 0.f);
    term(2, 0.f, 0.f, kSH1);
    term(3, -kSH1, 0.f, 0.f);
    if (deg > 1) {
      const float xx = x * x, yy = y * y, zz = z * z;
      const float xy = x * y, yz = y * z, xz = x * z;
      term(4, kSH2[0] * y, kSH2[0] * x, 0.f);
      term(5, 0.f, kSH2[1] * z, kSH2[1] * y);
      term(6, kSH2[2] * 2.f * -x, kSH2[2] * 2.f * -y, kSH2[2] * 2.f * 2.f * z);
      term(7, kSH2[3] * z, 0.f, kSH2[3] * x);
      term(8, kSH2[4] * 2.f * x, kSH2[4] * 2.f * -y, 0.f);
      if (deg > 2) {
        term(9, kSH3[0] * 3.f * 2.f * xy, kSH3[0] * 3.f * (xx - yy), 0.f);
        term(10, kSH3[1] * yz, kSH3[1] * xz, kSH3[1] * xy);
        term(11, kSH3[2] * -2.f * xy, kSH3[2] * (-3.f * yy + 4.f * zz - xx), kSH3[2] * 4.f * 2.f * yz);
        term(12, kSH3[3] * -3.f * 2.f * xz, kSH3[3] * -3.f * 2.f * yz, kSH3[3] * 3.f * (2.f * zz - xx - yy));
        term(13, kSH3[4] * (-3.f * xx + 4.f * zz - yy), kSH3[4] * -2.f * xy, kSH3[4] * 4.f * 2.f * xz);
        term(14, kSH3[5] * 2.f * xz, kSH3[5] * -2.f * yz, kSH3[5] * (xx - yy));
        term(15, kSH3[6] * 3.f * (xx - yy), kSH3[6] * -3.f * 2.f * xy, 0.f);
      }
    }
  }
  // dL/dsh[k][c] = basis_k * dL/dRGB[c]; basis_k = 0 above the active degree
  if (dL_dsh_out != nullptr) {
    const float g[3] = {dL_dRGB.x, dL_dRGB.y, dL_dRGB.z};
    if (VEC) {
      float4* o4 = reinterpret_cast<float4*>(__builtin_assume_aligned(dL_dsh_out, 16));
#pragma unroll
      for (int c = 0; c < 12; ++c)
        o4[c] = make_float4(basis[(4 * c) / 3] * g[(4 * c) % 3], basis[(4 * c + 1) / 3] * g[(4 * c + 1) % 3],
                            basis[(4 * c + 2) / 3] * g[(4 * c + 2) % 3], basis[(4 * c + 3) / 3] * g[(4 * c + 3) % 3]);
    } else {
#pragma unroll
      for (int k = 0; k < 16; ++k)
        if (k < M) {
          dL_dsh_out[3 * k] = basis[k] * g[0];
          dL_dsh_out[3 * k + 1] = basis[k] * g[1];
          dL_dsh_out[3 * k + 2] = basis[k] * g[2];
        }
      for (int k = 16; k < M; ++k) {   // storage beyond degree 3: no basis function, zero gradient
        dL_dsh_out[3 * k] = 0.f;
        dL_dsh_out[3 * k + 1] = 0.f;
        dL_dsh_out[3 * k + 2] = 0.f;
      }
    }
  }
  const float3 dL_ddir = make_float3(dot(dRGBdx, dL_dRGB), dot(dRGBdy, dL_dRGB), dot(dRGBdz, dL_dRGB));
  return dnormalize(dir_orig, dL_ddir);
}

// dL/dSigma (6 packed) -> dL/dscale, dL/dquaternion (no normalisation Jacobian; the reference
// has it commented out, backward.cu:281,340).
__device__ __forceinline__ void cov3d_backward(const float3 scale, float mod, const float4 rot,
                                               const float* dL_dcov6, float3& dL_dscale,
                                               float4& dL_drot) {
  const float r = rot.x, x = rot.y, y = rot.z, z = rot.w;
  const Mat3 R = mat3_cols(1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y),
                           2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x),
                           2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y));
  Mat3 S = mat3_cols(1.f, 0.f, 0.f, 0.f, 1.f, 0.f, 0.f, 0.f, 1.f);
  const float3 s = make_float3(mod * scale.x, mod * scale.y, mod * scale.z);
  S.c[0][0] = s.x;
  S.c[1][1] = s.y;
  S.c[2][2] = s.z;
  const Mat3 M = mul(S, R);
  const Mat3 dSigma = mat3_cols(dL_dcov6[0], 0.5f * dL_dcov6[1], 0.5f * dL_dcov6[2],
                                0.5f * dL_dcov6[1], dL_dcov6[3], 0.5f * dL_dcov6[4],
                                0.5f * dL_dcov6[2], 0.5f * dL_dcov6[4], dL_dcov6[5]);
  Mat3 M2 = M;
#pragma unroll
  for (int c = 0; c < 3; ++c)
#pragma unroll
    for (int rr = 0; rr < 3; ++rr) M2.c[c][rr] = M.c[c][rr] * 2.0f;
  const Mat3 dL_dM = mul(M2, dSigma);
  const Mat3 Rt = transpose(R);
  Mat3 dMt = transpose(dL_dM);
  dL_dscale.x = Rt.c[0][0] * dMt.c[0][0] + Rt.c[0][1] * dMt.c[0][1] + Rt.c[0][2] * dMt.c[0][2];
  dL_dscale.y = Rt.c[1][0] * dMt.c[1][0] + Rt.c[1][1] * dMt.c[1][1] + Rt.c[1][2] * dMt.c[1][2];
  dL_dscale.z = Rt.c[2][0] * dMt.c[2][0] + Rt.c[2][1] * dMt.c[2][1] + Rt.c[2][2] * dMt.c[2][2];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    dMt.c[0][k] *= s.x;
    dMt.c[1][k] *= s.y;
    dMt.c[2][k] *= s.z;
  }
  dL_drot.x = 2 * z * (dMt.c[0][1] - dMt.c[1][0]) + 2 * y * (dMt.c[2][0] - dMt.c[0][2]) +
              2 * x * (dMt.c[1][2] - dMt.c[2][1]);
  dL_drot.y = 2 * y * (dMt.c[1][0] + dMt.c[0][1]) + 2 * z * (dMt.c[2][0] + dMt.c[0][2]) +
              2 * r * (dMt.c[1][2] - dMt.c[2][1]) - 4 * x * (dMt.c[2][2] + dMt.c[1][1]);
  dL_drot.z = 2 * x * (dMt.c[1][0] + dMt.c[0][1]) + 2 * r * (dMt.c[2][0] - dMt.c[0][2]) +
              2 * z * (dMt.c[1][2] + dMt.c[2][1]) - 4 * y * (dMt.c[2][2] + dMt.c[0][0]);
  dL_drot.w = 2 * r * (dMt.c[0][1] - dMt.c[1][0]) + 2 * x * (dMt.c[2][0] + dMt.c[0][2]) +
              2 * y * (dMt.c[1][2] + dMt.c[2][1]) - 4 * z * (dMt.c[1][1] + dMt.c[0][0]);
}

// forward-side helper shared with preprocess.cu (same arithmetic, kept local to this TU)
__device__ __forceinline__ void cov3d_recompute(const float3 s, float mod, const float4 q, float* cov6) {
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

// Everything one Gaussian needs. `sh_in` / `sh_out` are this Gaussian's rows of SH coefficients
// and of their gradients; they may be the same shared-memory row (staged mode) because every
// coefficient is read before its gradient is written.
template <bool RAW, bool VEC>
__device__ __forceinline__ void
gaussian_backward(int idx, int D, int M, const float* __restrict__ means3D, const int* __restrict__ radii,
                  const float* sh_in, float* sh_out, bool has_sh,
                  const unsigned char* __restrict__ clamped, const float* __restrict__ scales,
                  const float* __restrict__ rotations, const float* __restrict__ cov3D_precomp,
                  const float* __restrict__ colors_precomp, float scale_modifier, const float* view,
                  const float* proj, const float* s_cam, float h_x, float h_y, float tan_fovx,
                  float tan_fovy, const float4* __restrict__ acc, float* __restrict__ dL_dmeans2D,
                  float* __restrict__ dL_dcolors, float* __restrict__ dL_dopacity,
                  float* __restrict__ dL_dmeans3D, float* __restrict__ dL_dcov3D,
                  float* __restrict__ dL_dscales, float* __restrict__ dL_drot,
                  // raw-parameter mode (egs_backward_raw): gradients go to the raw tensors
                  const RawModel& raw, const float* s_rot, const EgsModelGrads& rg, float* dT, bool colors_only,
                  // factored SH gradient: this Gaussian's clamp-masked dL/dRGB (zero when culled)
                  float3& drgb_out) {
  constexpr bool raw_geom = RAW;
  drgb_out = make_float3(0.f, 0.f, 0.f);
  if (colors_only) {   // EgsBackwardArgs.colors_only: unpack the colour accumulators, nothing else
    const bool seen = radii[idx] > 0;
    const float4 a2 = seen ? acc[3 * (size_t)idx + 2] : make_float4(0.f, 0.f, 0.f, 0.f);
    dL_dcolors[3 * idx + 0] = a2.x;
    dL_dcolors[3 * idx + 1] = a2.y;
    dL_dcolors[3 * idx + 2] = a2.z;
    return;
  }
  if (!(radii[idx] > 0)) {
    // culled: every gradient is zero
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      dL_dmeans2D[3 * idx + k] = 0.f;
      dL_dmeans3D[3 * idx + k] = 0.f;
    }
    if (RAW) rg.d_opacity_raw[idx] = 0.f;
    else dL_dopacity[idx] = 0.f;
    if (colors_precomp != nullptr) {
#pragma unroll
      for (int k = 0; k < 3; ++k) dL_dcolors[3 * idx + k] = 0.f;
    }
    if (raw_geom) {
#pragma unroll
      for (int k = 0; k < 3; ++k) rg.d_scaling_raw[3 * idx + k] = 0.f;
      reinterpret_cast<float4*>(rg.d_rotation_raw)[idx] = make_float4(0.f, 0.f, 0.f, 0.f);
    } else if (cov3D_precomp != nullptr) {
#pragma unroll
      for (int k = 0; k < 6; ++k) dL_dcov3D[6 * idx + k] = 0.f;
    }
    if (has_sh && sh_out != nullptr) {
      if (VEC) {
        float4* o4 = reinterpret_cast<float4*>(__builtin_assume_aligned(sh_out, 16));
#pragma unroll
        for (int c = 0; c < 12; ++c) o4[c] = make_float4(0.f, 0.f, 0.f, 0.f);
      } else {
        for (int k = 0; k < 3 * M; ++k) sh_out[k] = 0.f;
      }
    }
    if (scales != nullptr && !raw_geom) {
#pragma unroll
      for (int k = 0; k < 3; ++k) dL_dscales[3 * idx + k] = 0.f;
      reinterpret_cast<float4*>(dL_drot)[idx] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    return;
  }

  const float4 a0 = acc[3 * (size_t)idx + 0];  // dmean2D.x, dmean2D.y, dopacity, ddepth
  const float4 a1 = acc[3 * (size_t)idx + 1];  // dconic.x, .y, .w
  const float4 a2 = acc[3 * (size_t)idx + 2];  // dcolor rgb
  dL_dmeans2D[3 * idx + 0] = a0.x;
  dL_dmeans2D[3 * idx + 1] = a0.y;
  dL_dmeans2D[3 * idx + 2] = 0.f;
  if (RAW) rg.d_opacity_raw[idx] = sigmoid_bwd(sigmoid_fwd(raw.opacity_raw[idx]), a0.z);
  else dL_dopacity[idx] = a0.z;
  const float dL_ddepth = a0.w;

  const float3 mean = make_float3(means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2]);

  // ---- covariance chain (App. A.8b) --------------------------------------------------------
  float cov6[6];
  float3 scl = make_float3(0.f, 0.f, 0.f);
  float4 rot = make_float4(0.f, 0.f, 0.f, 0.f);
  if (raw_geom) {
    raw_covariance(raw, idx, s_rot, cov6);
  } else if (cov3D_precomp != nullptr) {
#pragma unroll
    for (int k = 0; k < 6; ++k) cov6[k] = cov3D_precomp[6 * idx + k];
  } else {
    scl = make_float3(scales[3 * idx], scales[3 * idx + 1], scales[3 * idx + 2]);
    rot = reinterpret_cast<const float4*>(rotations)[idx];
    cov3d_recompute(scl, scale_modifier, rot, cov6);
  }

  const float3 dL_dconic = make_float3(a1.x, a1.y, a1.z);
  float3 t = xform_point_4x3(mean, view);
  const float limx = 1.3f * tan_fovx;
  const float limy = 1.3f * tan_fovy;
  const float txtz = t.x / t.z;
  const float tytz = t.y / t.z;
  t.x = min(limx, max(-limx, txtz)) * t.z;
  t.y = min(limy, max(-limy, tytz)) * t.z;
  const float x_grad_mul = txtz < -limx || txtz > limx ? 0 : 1;
  const float y_grad_mul = tytz < -limy || tytz > limy ? 0 : 1;

  const Mat3 J = mat3_cols(h_x / t.z, 0.0f, -(h_x * t.x) / (t.z * t.z),
                           0.0f, h_y / t.z, -(h_y * t.y) / (t.z * t.z),
                           0.f, 0.f, 0.f);
  const Mat3 Wm = mat3_cols(view[0], view[4], view[8], view[1], view[5], view[9], view[2], view[6],
                            view[10]);
  const Mat3 Vrk = mat3_cols(cov6[0], cov6[1], cov6[2], cov6[1], cov6[3], cov6[4], cov6[2], cov6[4],
                             cov6[5]);
  const Mat3 T = mul(Wm, J);
  Mat3 cov2D = mul(mul(transpose(T), transpose(Vrk)), T);
  const float a = cov2D.c[0][0] += 0.3f;
  const float b = cov2D.c[0][1];
  const float c = cov2D.c[1][1] += 0.3f;

  const float denom = a * c - b * b;
  float dL_da = 0, dL_db = 0, dL_dc = 0;
  const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
  float dcov[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (denom2inv != 0) {
    dL_da = denom2inv * (-c * c * dL_dconic.x + 2 * b * c * dL_dconic.y + (denom - a * c) * dL_dconic.z);
    dL_dc = denom2inv * (-a * a * dL_dconic.z + 2 * a * b * dL_dconic.y + (denom - a * c) * dL_dconic.x);
    dL_db = denom2inv * 2 * (b * c * dL_dconic.x - (denom + 2 * b * b) * dL_dconic.y + a * b * dL_dconic.z);

    dcov[0] = (T.c[0][0] * T.c[0][0] * dL_da + T.c[0][0] * T.c[1][0] * dL_db + T.c[1][0] * T.c[1][0] * dL_dc);
    dcov[3] = (T.c[0][1] * T.c[0][1] * dL_da + T.c[0][1] * T.c[1][1] * dL_db + T.c[1][1] * T.c[1][1] * dL_dc);
    dcov[5] = (T.c[0][2] * T.c[0][2] * dL_da + T.c[0][2] * T.c[1][2] * dL_db + T.c[1][2] * T.c[1][2] * dL_dc);
    dcov[1] = 2 * T.c[0][0] * T.c[0][1] * dL_da + (T.c[0][0] * T.c[1][1] + T.c[0][1] * T.c[1][0]) * dL_db + 2 * T.c[1][0] * T.c[1][1] * dL_dc;
    dcov[2] = 2 * T.c[0][0] * T.c[0][2] * dL_da + (T.c[0][0] * T.c[1][2] + T.c[0][2] * T.c[1][0]) * dL_db + 2 * T.c[1][0] * T.c[1][2] * dL_dc;
    dcov[4] = 2 * T.c[0][2] * T.c[0][1] * dL_da + (T.c[0][1] * T.c[1][2] + T.c[0][2] * T.c[1][1]) * dL_db + 2 * T.c[1][1] * T.c[1][2] * dL_dc;
  }

  // dL/dT (upper 2x3), dL/dJ (4 non-zero entries), dL/dt
  const float dL_dT00 = 2 * (T.c[0][0] * Vrk.c[0][0] + T.c[0][1] * Vrk.c[0][1] + T.c[0][2] * Vrk.c[0][2]) * dL_da +
                        (T.c[1][0] * Vrk.c[0][0] + T.c[1][1] * Vrk.c[0][1] + T.c[1][2] * Vrk.c[0][2]) * dL_db;
  const float dL_dT01 = 2 * (T.c[0][0] * Vrk.c[1][0] + T.c[0][1] * Vrk.c[1][1] + T.c[0][2] * Vrk.c[1][2]) * dL_da +
                        (T.c[1][0] * Vrk.c[1][0] + T.c[1][1] * Vrk.c[1][1] + T.c[1][2] * Vrk.c[1][2]) * dL_db;
  const float dL_dT02 = 2 * (T.c[0][0] * Vrk.c[2][0] + T.c[0][1] * Vrk.c[2][1] + T.c[0][2] * Vrk.c[2][2]) * dL_da +
                        (T.c[1][0] * Vrk.c[2][0] + T.c[1][1] * Vrk.c[2][1] + T.c[1][2] * Vrk.c[2][2]) * dL_db;
  const float dL_dT10 = 2 * (T.c[1][0] * Vrk.c[0][0] + T.c[1][1] * Vrk.c[0][1] + T.c[1][2] * Vrk.c[0][2]) * dL_dc +
                        (T.c[0][0] * Vrk.c[0][0] + T.c[0][1] * Vrk.c[0][1] + T.c[0][2] * Vrk.c[0][2]) * dL_db;
  const float dL_dT11 = 2 * (T.c[1][0] * Vrk.c[1][0] + T.c[1][1] * Vrk.c[1][1] + T.c[1][2] * Vrk.c[1][2]) * dL_dc +
                        (T.c[0][0] * Vrk.c[1][0] + T.c[0][1] * Vrk.c[1][1] + T.c[0][2] * Vrk.c[1][2]) * dL_db;
  const float dL_dT12 = 2 * (T.c[1][0] * Vrk.c[2][0] + T.c[1][1] * Vrk.c[2][1] + T.c[1][2] * Vrk.c[2][2]) * dL_dc +
                        (T.c[0][0] * Vrk.c[2][0] + T.c[0][1] * Vrk.c[2][1] + T.c[0][2] * Vrk.c[2][2]) * dL_db;

  const float dL_dJ00 = Wm.c[0][0] * dL_dT00 + Wm.c[0][1] * dL_dT01 + Wm.c[0][2] * dL_dT02;
  const float dL_dJ02 = Wm.c[2][0] * dL_dT00 + Wm.c[2][1] * dL_dT01 + Wm.c[2][2] * dL_dT02;
  const float dL_dJ11 = Wm.c[1][0] * dL_dT10 + Wm.c[1][1] * dL_dT11 + Wm.c[1][2] * dL_dT12;
  const float dL_dJ12 = Wm.c[2][0] * dL_dT10 + Wm.c[2][1] * dL_dT11 + Wm.c[2][2] * dL_dT12;

  const float tz = 1.f / t.z;
  const float tz2 = tz * tz;
  const float tz3 = tz2 * tz;
  const float dL_dtx = x_grad_mul * -h_x * tz2 * dL_dJ02;
  const float dL_dty = y_grad_mul * -h_y * tz2 * dL_dJ12;
  const float dL_dtz = -h_x * tz2 * dL_dJ00 - h_y * tz2 * dL_dJ11 + (2 * h_x * t.x) * tz3 * dL_dJ02 +
                       (2 * h_y * t.y) * tz3 * dL_dJ12;
  float3 dmean = xform_vec_4x3_T(make_float3(dL_dtx, dL_dty, dL_dtz), view);

  // ---- projection part (App. A.8c (1)) ------------------------------------------------------
  {
    const float4 m_hom = xform_point_4x4(mean, proj);
    const float m_w = 1.0f / (m_hom.w + 0.0000001f);
    const float mul1 = (proj[0] * mean.x + proj[4] * mean.y + proj[8] * mean.z + proj[12]) * m_w * m_w;
    const float mul2 = (proj[1] * mean.x + proj[5] * mean.y + proj[9] * mean.z + proj[13]) * m_w * m_w;
    float3 g;
    g.x = (proj[0] * m_w - proj[3] * mul1) * a0.x + (proj[1] * m_w - proj[3] * mul2) * a0.y;
    g.y = (proj[4] * m_w - proj[7] * mul1) * a0.x + (proj[5] * m_w - proj[7] * mul2) * a0.y;
    g.z = (proj[8] * m_w - proj[11] * mul1) * a0.x + (proj[9] * m_w - proj[11] * mul2) * a0.y;
    dmean.x += g.x; dmean.y += g.y; dmean.z += g.z;
  }
  // ---- depth part (App. A.8c (2)) -----------------------------------------------------------
  {
    const float mul3 = view[2] * mean.x + view[6] * mean.y + view[10] * mean.z + view[14];
    float3 g;
    g.x = (view[2] - view[3] * mul3) * dL_ddepth;
    g.y = (view[6] - view[7] * mul3) * dL_ddepth;
    g.z = (view[10] - view[11] * mul3) * dL_ddepth;
    dmean.x += g.x; dmean.y += g.y; dmean.z += g.z;
  }
  // ---- colour part --------------------------------------------------------------------------
  if (has_sh) {
    F3 drgb = {a2.x, a2.y, a2.z};
    const float3 g = sh_backward<VEC>(D, M, mean, make_float3(s_cam[0], s_cam[1], s_cam[2]), sh_in, clamped[idx], drgb,
                                      sh_out);
    dmean.x += g.x; dmean.y += g.y; dmean.z += g.z;
    drgb_out = make_float3(drgb.x, drgb.y, drgb.z);
  } else {
    dL_dcolors[3 * idx + 0] = a2.x;
    dL_dcolors[3 * idx + 1] = a2.y;
    dL_dcolors[3 * idx + 2] = a2.z;
  }
  dL_dmeans3D[3 * idx + 0] = dmean.x;
  dL_dmeans3D[3 * idx + 1] = dmean.y;
  dL_dmeans3D[3 * idx + 2] = dmean.z;

  // ---- covariance parametrisation -------------------------------------------------------------
  if (raw_geom) {
    float ds[3], dq[4];
    raw_covariance_backward(raw, idx, s_rot, dcov, ds, dq, dT);
    rg.d_scaling_raw[3 * idx + 0] = ds[0];
    rg.d_scaling_raw[3 * idx + 1] = ds[1];
    rg.d_scaling_raw[3 * idx + 2] = ds[2];
    reinterpret_cast<float4*>(rg.d_rotation_raw)[idx] = make_float4(dq[0], dq[1], dq[2], dq[3]);
  } else if (cov3D_precomp != nullptr) {
#pragma unroll
    for (int k = 0; k < 6; ++k) dL_dcov3D[6 * idx + k] = dcov[k];
  } else {
    float3 ds;
    float4 dq;
    cov3d_backward(scl, scale_modifier, rot, dcov, ds, dq);
    dL_dscales[3 * idx + 0] = ds.x;
    dL_dscales[3 * idx + 1] = ds.y;
    dL_dscales[3 * idx + 2] = ds.z;
    reinterpret_cast<float4*>(dL_drot)[idx] = dq;
  }
}

// One thread per Gaussian, 128-thread blocks. With 16 coefficients (SH degree 3 storage) the SH rows
// of a warp's 32 Gaussians -- 6 KB, contiguous in memory -- are staged through shared memory with
// float4 accesses both ways: a thread reading or writing its own 192-byte row straight from global
// memory makes every warp instruction touch 32 sectors, which left this kernel bound by L2
// transactions (ncu: lts 71 %, dram 20 %) instead of by HBM.
constexpr int PB_THREADS = 128;
constexpr int PB_WARPS = PB_THREADS / 32;
constexpr int SH_ROW = 48;        // 16 coefficients x RGB
// pitch of a staged row: 52 floats for the contiguous [P,16,3] tensor (16-byte aligned rows, read and
// written with 16-byte shared accesses), 49 (odd: conflict-free scalar accesses) for the split dc /
// rest rows, whose pieces start at 12- and 180-byte offsets
constexpr int SH_PITCH = 52;
constexpr int SH_PITCH_SPLIT = 49;

// SH_MODE as in the forward kernel: 0 = rows straight from global memory, 1 = staged float4 rows of
// one contiguous [P,16,3] tensor, 2 = rows staged from / written back to the split features_dc /
// features_rest tensors of the raw-parameter mode.
// Destinations of the factored SH gradient (EgsBackwardArgs.dL_drgb, drgb_remote, drgb_multicast): the
// local slot and this view's slot in the other ranks' gather buffers.
struct DrgbSlots {
  float* local;
  float* remote[EGS_MAX_PEERS];
  float* multicast;
  int n_remote;
};

__device__ __forceinline__ void st_sys_f4(float* p, float4 v) {
  asm volatile("st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}
__device__ __forceinline__ void st_multicast_f4(float* p, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y),
               "f"(v.z), "f"(v.w)
               : "memory");
}
// 16 bytes of a slot to every destination (float offset `at`, a multiple of 4)
__device__ __forceinline__ void store_slot_f4(const DrgbSlots& d, size_t at, float4 v) {
  *reinterpret_cast<float4*>(d.local + at) = v;
  if (d.multicast != nullptr) {
    st_multicast_f4(d.multicast + at, v);
  } else {
    for (int q = 0; q < d.n_remote; ++q) st_sys_f4(d.remote[q] + at, v);
  }
}

// FACTORED: dL/dsh is not written (neither the rows in shared memory nor the 12 M bytes per Gaussian in
// HBM); the clamp-masked dL/dRGB rows and the camera centre go to the slots instead.
template <int SH_MODE, bool RAW, bool FACTORED>
__global__ void __launch_bounds__(PB_THREADS, RAW ? 5 : 6)   // raw mode carries the activation chain: 102 registers
preprocess_backward_kernel(RawModel raw, EgsModelGrads rg, int P, int D, int M, const float* __restrict__ means3D,
                           const int* __restrict__ radii, const float* __restrict__ shs,
                           const unsigned char* __restrict__ clamped,
                           const float* __restrict__ scales, const float* __restrict__ rotations,
                           const float* __restrict__ cov3D_precomp,
                           const float* __restrict__ colors_precomp, float scale_modifier,
                           const float* __restrict__ viewmatrix, const float* __restrict__ projmatrix,
                           const float* __restrict__ campos, float h_x, float h_y, float tan_fovx,
                           float tan_fovy, const float4* __restrict__ acc,
                           float* __restrict__ dL_dmeans2D, float* __restrict__ dL_dcolors,
                           float* __restrict__ dL_dopacity, float* __restrict__ dL_dmeans3D,
                           float* __restrict__ dL_dcov3D, float* __restrict__ dL_dsh,
                           float* __restrict__ dL_dscales, float* __restrict__ dL_drot, bool colors_only,
                           const DrgbSlots slots) {
  __shared__ float view[16], proj[16], s_cam[3];
  __shared__ __align__(16) float s_sh[SH_MODE != 0 ? PB_WARPS * 32 * SH_PITCH : 4];
  // 16 coefficients: 16-byte aligned rows and vector accesses (also for the split dc | rest rows, which
  // sit side by side in one row); fewer: the odd pitch and scalar accesses
  const bool vec = SH_MODE == 1 || M == 16;
  const int PITCH = vec ? SH_PITCH : SH_PITCH_SPLIT;
  __shared__ float s_rot[RAW ? 18 : 1];
  __shared__ float s_dT[RAW ? PB_WARPS : 1][9];
  if (threadIdx.x < 16) {
    view[threadIdx.x] = viewmatrix[threadIdx.x];
    proj[threadIdx.x] = projmatrix[threadIdx.x];
  }
  if (threadIdx.x < 3) s_cam[threadIdx.x] = campos[threadIdx.x];
  if (RAW) stage_rotations(raw, s_rot);
  __syncthreads();
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float dT[RAW ? 9 : 1];
#pragma unroll
  for (int k = 0; k < (RAW ? 9 : 1); ++k) dT[k] = 0.f;
  float3 drgb = make_float3(0.f, 0.f, 0.f);
  if (FACTORED && blockIdx.x == 0 && threadIdx.x == 0)   // the view's camera centre closes the slot
    store_slot_f4(slots, ((size_t)P * 3 + 3) / 4 * 4, make_float4(s_cam[0], s_cam[1], s_cam[2], 0.f));

#define EGS_GB_TAIL                                                                                          \
  clamped, scales, rotations, cov3D_precomp, colors_precomp, scale_modifier, view, proj, s_cam, h_x, h_y,     \
      tan_fovx, tan_fovy, acc, dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dscales,      \
      dL_drot, raw, s_rot, rg, dT, colors_only, drgb

  if (SH_MODE != 0) {
    // rows [first, first + rows) of this warp
    const int first = blockIdx.x * blockDim.x + warp * 32;
    const int rows = min(32, P - first);
    float* tile = s_sh + warp * 32 * SH_PITCH;
    const int wrest = 3 * M - 3;
    if (rows > 0) {
      if (SH_MODE == 1) {   // one contiguous run of rows * 48 floats, 16 bytes at a time
        stage_rows_async16(tile, PITCH, SH_ROW, shs + (size_t)first * SH_ROW, rows * SH_ROW, lane);
      } else {              // two contiguous runs: rows * 3 and rows * (3M - 3) floats
        stage_rows_async(tile, PITCH, 0, 3, raw.features_dc + (size_t)first * 3, rows * 3, lane);
        if (wrest > 0)
          stage_rows_async(tile, PITCH, 3, wrest, raw.features_rest + (size_t)first * wrest, rows * wrest, lane);
      }
    }
    stage_rows_wait();
    if (idx < P) {
      float* row = tile + lane * PITCH;
      if (vec) gaussian_backward<RAW, true>(idx, D, M, means3D, radii, row, FACTORED ? nullptr : row, true, EGS_GB_TAIL);
      else gaussian_backward<RAW, false>(idx, D, M, means3D, radii, row, FACTORED ? nullptr : row, true, EGS_GB_TAIL);
    }
    __syncwarp();
    if (FACTORED && rows > 0) {
      // the warp's 32 triples are 384 contiguous bytes of the slot: through the (now free) tile, then
      // 24 lanes store 16 bytes each to every destination
      tile[3 * lane] = drgb.x;
      tile[3 * lane + 1] = drgb.y;
      tile[3 * lane + 2] = drgb.z;
      __syncwarp();
      const size_t at = (size_t)first * 3 + lane * 4;
      if (lane < 24 && at < (size_t)P * 3)   // the last 16 bytes may run into the slot's padding
        store_slot_f4(slots, at, make_float4(tile[4 * lane], tile[4 * lane + 1], tile[4 * lane + 2], tile[4 * lane + 3]));
    }
    if (rows > 0 && !FACTORED) {
      if (SH_MODE == 1) {
        float4* dst = reinterpret_cast<float4*>(dL_dsh + (size_t)first * SH_ROW);
        for (int f = lane; f < rows * (SH_ROW / 4); f += 32) {
          const int row = f / (SH_ROW / 4), k = (f % (SH_ROW / 4)) * 4;
          dst[f] = *reinterpret_cast<const float4*>(tile + row * PITCH + k);
        }
      } else {
        warp_rows_copy<false>(tile, PITCH, 0, 3, rg.d_features_dc + (size_t)first * 3, rows * 3, lane);
        if (wrest > 0)
          warp_rows_copy<false>(tile, PITCH, 3, wrest, rg.d_features_rest + (size_t)first * wrest, rows * wrest,
                                lane);
      }
    }
  } else if (idx < P) {
    const bool has_sh = shs != nullptr;
    const size_t off = (size_t)idx * M * 3;
    gaussian_backward<RAW, false>(idx, D, M, means3D, radii, has_sh ? shs + off : nullptr,
                                  has_sh && !FACTORED ? dL_dsh + off : nullptr, has_sh, EGS_GB_TAIL);
    if (FACTORED) {   // unstaged rows (sh_coeffs != 16): scalar stores to every destination
      const float v[3] = {drgb.x, drgb.y, drgb.z};
      for (int k = 0; k < 3; ++k) {
        slots.local[3 * (size_t)idx + k] = v[k];
        for (int q = 0; q < slots.n_remote; ++q)
          asm volatile("st.relaxed.sys.global.f32 [%0], %1;" ::"l"(slots.remote[q] + 3 * (size_t)idx + k), "f"(v[k]) : "memory");
      }
    }
  }
#undef EGS_GB_TAIL
  // dL/d train_R: every thread of the CTA takes part (no early return above)
  if (RAW) {
    if (raw.train_R != nullptr && rg.d_train_R != nullptr) reduce_pose_gradient(dT, s_dT, rg.d_train_R);
  }
}

int launch_preprocess_backward(const EgsBackwardArgs& a, const EgsModelArgs* model, const EgsModelGrads* grads,
                               const GeomPtrs& g, const BackwardScratch& s, cudaStream_t stream) {
  const float focal_y = a.height / (2.0f * a.tan_fovy);
  const float focal_x = a.width / (2.0f * a.tan_fovx);
  StageTimer timer(stream, ST_PREPROCESS_BWD);
  const int blocks = (a.P + PB_THREADS - 1) / PB_THREADS;
  const bool staged = a.shs != nullptr && a.sh_coeffs == 16 &&
                      (reinterpret_cast<uintptr_t>(a.shs) & 15u) == 0 &&
                      (a.dL_drgb != nullptr || (reinterpret_cast<uintptr_t>(a.dL_dsh) & 15u) == 0);
  RawModel raw{};
  EgsModelGrads rg{};
  if (model != nullptr) {
    raw = raw_model_from(model);
    rg = *grads;
    if (rg.d_train_R != nullptr && raw.train_R != nullptr)
      EGS_CUDA_TRY(cudaMemsetAsync(rg.d_train_R, 0, 9 * sizeof(float), stream));
  }
  const bool split = model != nullptr && a.colors_precomp == nullptr;
  const int sh_coeffs = split ? model->sh_coeffs : a.sh_coeffs;
#define EGS_PB_ARGS                                                                                   \
  raw, rg, a.P, a.sh_degree, sh_coeffs, a.means3D, a.radii, a.shs, g.clamped, a.scales, a.rotations,         \
      a.cov3D_precomp, a.colors_precomp, a.scale_modifier, a.viewmatrix, a.projmatrix, a.campos,      \
      focal_x, focal_y, a.tan_fovx, a.tan_fovy, s.acc, a.dL_dmeans2D, a.dL_dcolors, a.dL_dopacity,    \
      a.dL_dmeans3D, a.dL_dcov3D, a.dL_dsh, a.dL_dscales, a.dL_drotations, a.colors_only != 0, slots
  const bool factored = a.dL_drgb != nullptr;
  DrgbSlots slots = {};
  slots.local = a.dL_drgb;
  slots.n_remote = a.drgb_n_remote;
  for (int q = 0; q < a.drgb_n_remote && q < EGS_MAX_PEERS; ++q) slots.remote[q] = a.drgb_remote[q];
  // one multicast store replaces the peer stores only where rows are stored 16 bytes at a time
  slots.multicast = (staged || split) ? a.drgb_multicast : nullptr;
  if (slots.multicast != nullptr) slots.n_remote = 0;
#define EGS_PB_LAUNCH(MODE, RAWMODE)                                                                          \
  do {                                                                                                        \
    if (factored) preprocess_backward_kernel<MODE, RAWMODE, true><<<blocks, PB_THREADS, 0, stream>>>(EGS_PB_ARGS); \
    else preprocess_backward_kernel<MODE, RAWMODE, false><<<blocks, PB_THREADS, 0, stream>>>(EGS_PB_ARGS);     \
  } while (0)
  if (split)
    EGS_PB_LAUNCH(2, true);
  else if (model != nullptr)
    EGS_PB_LAUNCH(0, true);
  else if (staged)
    EGS_PB_LAUNCH(1, false);
  else
    EGS_PB_LAUNCH(0, false);
#undef EGS_PB_LAUNCH
#undef EGS_PB_ARGS
  EGS_AFTER_LAUNCH(stream, a.debug);
  return EGS_OK;
}

}  // namespace egs
