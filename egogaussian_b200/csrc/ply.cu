// ply.cu -- the checkpoint format either side of the path: GaussianModel.save_ply /
// load_ply (scene/gaussian_model.py:340-481) keep one record of 4-byte floats per Gaussian,
//   x y z  nx ny nz  f_dc_0..2  f_rest_0..3(M-1)-1  opacity  scale_0..2  rot_0..3  [label generation is_object]
// with the SH coefficients channel-major (f_rest_j, j = c * (M - 1) + k, is features_rest[p][k][c]).
// The reference builds the records on the host through a Python tuple per Gaussian; here the rows are
// packed / unpacked on the device (one thread per float, coalesced on the row side) and cross PCIe
// once as one block. Pure data movement: bit-exact by construction.
#include "common.cuh"

namespace egs {

constexpr int PLY_FIXED = 6 + 3;   // xyz, normals, f_dc

struct PlyLayout {
  int P, M, K;          // Gaussians, SH coefficients per channel, floats per record
  int rest;             // 3 * (M - 1)
  int with_extras;
};

// canonical column -> value
__device__ __forceinline__ float ply_value(const PlyLayout& L, const EgsPlyColumns& c, int p, int k) {
  if (k < 3) return c.xyz[3 * p + k];
  if (k < 6) return 0.f;                                        // normals: always zero
  if (k < 9) return c.features_dc[3 * p + (k - 6)];
  k -= PLY_FIXED;
  if (k < L.rest) {
    const int ch = k / (L.M - 1), kk = k - ch * (L.M - 1);
    return c.features_rest[((size_t)p * (L.M - 1) + kk) * 3 + ch];
  }
  k -= L.rest;
  if (k == 0) return c.opacity[p];
  if (k < 4) return c.scaling[3 * p + (k - 1)];
  if (k < 8) return c.rotation[4 * p + (k - 4)];
  if (k == 8) return c.label[p];
  if (k == 9) return (float)c.generation[p];
  return (float)c.is_object[p];
}

__global__ void __launch_bounds__(256)
ply_pack_kernel(const PlyLayout L, const EgsPlyColumns c, float* __restrict__ rows) {
  const size_t total = (size_t)L.P * L.K;
  for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < total; i += (size_t)gridDim.x * 256) {
    const int p = (int)(i / L.K), k = (int)(i - (size_t)p * L.K);
    rows[i] = ply_value(L, c, p, k);
  }
}

struct PlyColumnMap {
  int src[EGS_PLY_MAX_COLUMNS];   // canonical column -> column of the file's record, or -1: absent
};

__global__ void __launch_bounds__(256)
ply_unpack_kernel(const PlyLayout L, const EgsPlyColumnsOut c, const float* __restrict__ rows, int stride,
                  const PlyColumnMap map, float label_default, int is_object_default) {
  const int K = 14 + 3 * L.M + 3;   // always produce the extras
  const size_t total = (size_t)L.P * K;
  for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < total; i += (size_t)gridDim.x * 256) {
    const int p = (int)(i / K);
    int k = (int)(i - (size_t)p * K);
    const int s = map.src[k];
    const float v = s >= 0 ? rows[(size_t)p * stride + s] : 0.f;
    if (k < 3) { c.xyz[3 * p + k] = v; continue; }
    if (k < 6) continue;
    if (k < 9) { c.features_dc[3 * p + (k - 6)] = v; continue; }
    k -= PLY_FIXED;
    if (k < L.rest) {
      const int ch = k / (L.M - 1), kk = k - ch * (L.M - 1);
      c.features_rest[((size_t)p * (L.M - 1) + kk) * 3 + ch] = v;
      continue;
    }
    k -= L.rest;
    if (k == 0) c.opacity[p] = v;
    else if (k < 4) c.scaling[3 * p + (k - 1)] = v;
    else if (k < 8) c.rotation[4 * p + (k - 4)] = v;
    else if (k == 8) c.label[p] = s >= 0 ? v : label_default;
    else if (k == 9) c.generation[p] = s >= 0 ? (int)v : 0;
    else c.is_object[p] = s >= 0 ? (int)v : is_object_default;
  }
}

static unsigned ply_grid(size_t total) {
  size_t b = (total + 255) / 256;
  return (unsigned)(b < 1 ? 1 : b > 148 * 16 ? 148 * 16 : b);
}

}  // namespace egs

extern "C" int egs_ply_record_floats(int sh_coeffs, int with_extras) {
  return sh_coeffs < 1 ? 0 : 14 + 3 * sh_coeffs + (with_extras ? 3 : 0);
}

extern "C" int egs_ply_pack(int P, int sh_coeffs, const EgsPlyColumns* cols, int with_extras, float* rows,
                            void* stream) {
  using namespace egs;
  if (P < 0 || sh_coeffs < 1 || !cols || (P > 0 && !rows)) {
    set_error("egs_ply_pack: negative P, sh_coeffs < 1 or NULL argument");
    return EGS_ERR_INVALID_ARGUMENT;
  }
  if (P == 0) return EGS_OK;
  if (!cols->xyz || !cols->features_dc || (sh_coeffs > 1 && !cols->features_rest) || !cols->opacity || !cols->scaling ||
      !cols->rotation || (with_extras && (!cols->label || !cols->generation || !cols->is_object))) {
    set_error("egs_ply_pack: a column is missing");
    return EGS_ERR_INVALID_ARGUMENT;
  }
  PlyLayout L = {P, sh_coeffs, egs_ply_record_floats(sh_coeffs, with_extras), 3 * (sh_coeffs - 1), with_extras};
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  ply_pack_kernel<<<ply_grid((size_t)P * L.K), 256, 0, st>>>(L, *cols, rows);
  EGS_AFTER_LAUNCH(st, false);
  return EGS_OK;
}

extern "C" int egs_ply_unpack(int P, int sh_coeffs, const float* rows, int record_floats, const int* column_of,
                              const EgsPlyColumnsOut* cols, float label_default, int is_object_default, void* stream) {
  using namespace egs;
  const int K = egs_ply_record_floats(sh_coeffs, 1);
  if (P < 0 || sh_coeffs < 1 || K > EGS_PLY_MAX_COLUMNS || !cols || !column_of || (P > 0 && !rows) || record_floats < 1) {
    set_error("egs_ply_unpack: bad sizes (at most %d canonical columns) or NULL argument", EGS_PLY_MAX_COLUMNS);
    return EGS_ERR_INVALID_ARGUMENT;
  }
  if (P == 0) return EGS_OK;
  if (!cols->xyz || !cols->features_dc || (sh_coeffs > 1 && !cols->features_rest) || !cols->opacity || !cols->scaling ||
      !cols->rotation || !cols->label || !cols->generation || !cols->is_object) {
    set_error("egs_ply_unpack: an output column is missing");
    return EGS_ERR_INVALID_ARGUMENT;
  }
  PlyColumnMap map;
  for (int k = 0; k < K; ++k) {
    const bool optional = (k >= 3 && k < 6) || k >= K - 3;   // normals and the three extras may be absent
    if (column_of[k] >= record_floats || (column_of[k] < 0 && !optional)) {
      set_error("egs_ply_unpack: canonical column %d is missing from the file or out of range", k);
      return EGS_ERR_INVALID_ARGUMENT;
    }
    map.src[k] = column_of[k];
  }
  PlyLayout L = {P, sh_coeffs, K, 3 * (sh_coeffs - 1), 1};
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  ply_unpack_kernel<<<ply_grid((size_t)P * K), 256, 0, st>>>(L, *cols, rows, record_floats, map, label_default,
                                                             is_object_default);
  EGS_AFTER_LAUNCH(st, false);
  return EGS_OK;
}
