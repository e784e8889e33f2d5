// model_math.cuh -- per-Gaussian arithmetic of the GaussianModel activations that sit between the
// optimiser's raw parameters and the rasterizer inputs:
//
//   get_scaling  = exp(_scaling)                                scene/gaussian_model.py:126-128
//   get_opacity  = sigmoid(_opacity)                            scene/gaussian_model.py:151-153
//   build_rotation(_rotation): q = r / |r|, R(q)                utils/general_utils.py:124-146
//   build_scaling_rotation: L = R diag(mod * s)                 utils/general_utils.py:148-158
//   build_covariance_from_scaling_rotation[_w_rot]:
//        L' = [train_R] accum_R L on the selected rows, Sigma = L' L'^T, 6 upper entries
//                                                               scene/gaussian_model.py:29-63
//
// Shared by the standalone activation kernels (model.cu) and the raw-parameter mode of the
// preprocess kernels, so both produce bit-identical covariances.
//
// The reference evaluates these with separate torch kernels, i.e. every product and sum is rounded
// on its own. The quaternion -> R part is written with explicit round-to-nearest intrinsics so
// that nvcc does not contract it into FMAs and R, L match torch bit for bit; the 3-term dot
// products behind `@` (cuBLAS bmm, summation order unspecified) are plain FMA chains and are
// tolerance-checked (tests/test_model_gpu.py).
#pragma once
#include <cuda_runtime.h>
#include "../../include/egs_rasterizer.h"

namespace egs {

struct M33 {
  float m[3][3];   // row-major m[row][col], like the torch tensors it stands for
};

__device__ __forceinline__ M33 m33_load(const float* __restrict__ p) {
  M33 r;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) r.m[i][j] = p[3 * i + j];
  return r;
}

// C = A B
__device__ __forceinline__ M33 m33_mul(const M33& A, const M33& B) {
  M33 C;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j)
      C.m[i][j] = fmaf(A.m[i][2], B.m[2][j], fmaf(A.m[i][1], B.m[1][j], A.m[i][0] * B.m[0][j]));
  return C;
}

// C = A^T B
__device__ __forceinline__ M33 m33_tmul(const M33& A, const M33& B) {
  M33 C;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j)
      C.m[i][j] = fmaf(A.m[2][i], B.m[2][j], fmaf(A.m[1][i], B.m[1][j], A.m[0][i] * B.m[0][j]));
  return C;
}

// C = A B^T
__device__ __forceinline__ M33 m33_mult(const M33& A, const M33& B) {
  M33 C;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j)
      C.m[i][j] = fmaf(A.m[i][2], B.m[j][2], fmaf(A.m[i][1], B.m[j][1], A.m[i][0] * B.m[j][0]));
  return C;
}

// What the forward of one Gaussian leaves behind for its backward.
struct RawFrame {
  float s[3];      // exp(_scaling)
  float sm[3];     // scale_modifier * s
  float q[4];      // normalised quaternion (r, x, y, z)
  float norm;      // |raw quaternion|
  M33 R;           // rotation of the normalised quaternion
};

// general_utils.py:124-146, one rounding per torch op.
__device__ __forceinline__ void raw_frame(const float s_raw[3], const float4 q_raw, float mod, RawFrame& f) {
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    f.s[j] = expf(s_raw[j]);
    f.sm[j] = __fmul_rn(mod, f.s[j]);
  }
  float n2 = __fadd_rn(__fmul_rn(q_raw.x, q_raw.x), __fmul_rn(q_raw.y, q_raw.y));
  n2 = __fadd_rn(n2, __fmul_rn(q_raw.z, q_raw.z));
  n2 = __fadd_rn(n2, __fmul_rn(q_raw.w, q_raw.w));
  f.norm = __fsqrt_rn(n2);
  const float r = __fdiv_rn(q_raw.x, f.norm), x = __fdiv_rn(q_raw.y, f.norm);
  const float y = __fdiv_rn(q_raw.z, f.norm), z = __fdiv_rn(q_raw.w, f.norm);
  f.q[0] = r; f.q[1] = x; f.q[2] = y; f.q[3] = z;
  const float xx = __fmul_rn(x, x), yy = __fmul_rn(y, y), zz = __fmul_rn(z, z);
  const float xy = __fmul_rn(x, y), xz = __fmul_rn(x, z), yz = __fmul_rn(y, z);
  const float rx = __fmul_rn(r, x), ry = __fmul_rn(r, y), rz = __fmul_rn(r, z);
  f.R.m[0][0] = __fsub_rn(1.f, __fmul_rn(2.f, __fadd_rn(yy, zz)));
  f.R.m[0][1] = __fmul_rn(2.f, __fsub_rn(xy, rz));
  f.R.m[0][2] = __fmul_rn(2.f, __fadd_rn(xz, ry));
  f.R.m[1][0] = __fmul_rn(2.f, __fadd_rn(xy, rz));
  f.R.m[1][1] = __fsub_rn(1.f, __fmul_rn(2.f, __fadd_rn(xx, zz)));
  f.R.m[1][2] = __fmul_rn(2.f, __fsub_rn(yz, rx));
  f.R.m[2][0] = __fmul_rn(2.f, __fsub_rn(xz, ry));
  f.R.m[2][1] = __fmul_rn(2.f, __fadd_rn(yz, rx));
  f.R.m[2][2] = __fsub_rn(1.f, __fmul_rn(2.f, __fadd_rn(xx, yy)));
}

// L = R diag(sm): the bmm against a diagonal matrix adds exact zeros, so one product per entry.
__device__ __forceinline__ M33 frame_factor(const RawFrame& f) {
  M33 L;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) L.m[i][j] = __fmul_rn(f.R.m[i][j], f.sm[j]);
  return L;
}

// strip_symmetric(L L^T): [xx, xy, xz, yy, yz, zz]   (general_utils.py:110-122)
__device__ __forceinline__ void factor_to_cov6(const M33& L, float c[6]) {
  auto dot = [&](int a, int b) {
    return fmaf(L.m[a][2], L.m[b][2], fmaf(L.m[a][1], L.m[b][1], L.m[a][0] * L.m[b][0]));
  };
  c[0] = dot(0, 0); c[1] = dot(0, 1); c[2] = dot(0, 2);
  c[3] = dot(1, 1); c[4] = dot(1, 2); c[5] = dot(2, 2);
}

// The selected rows' factor goes through accum_R and then, during pose training, through
// ObjectMove.rot_L (geometry_utils.py:26-28): L' = train_R (accum_R L).
__device__ __forceinline__ M33 rotate_factor(const M33& L0, const M33* accum_R, const M33* train_R, M33* mid) {
  M33 L = L0;
  if (accum_R) L = m33_mul(*accum_R, L);
  if (mid) *mid = L;
  if (train_R) L = m33_mul(*train_R, L);
  return L;
}

// Gradient of the six stored covariance entries w.r.t. the final factor: only the upper triangle of
// Sigma is read, so dSigma is upper-triangular and dL = (dSigma + dSigma^T) L.
__device__ __forceinline__ M33 cov6_backward_to_factor(const float d[6], const M33& L) {
  M33 G;
  G.m[0][0] = 2.f * d[0]; G.m[0][1] = d[1];       G.m[0][2] = d[2];
  G.m[1][0] = d[1];       G.m[1][1] = 2.f * d[3]; G.m[1][2] = d[4];
  G.m[2][0] = d[2];       G.m[2][1] = d[4];       G.m[2][2] = 2.f * d[5];
  return m33_mul(G, L);
}

// dL0 (gradient of L = R diag(sm)) -> raw scaling and raw quaternion gradients.
__device__ __forceinline__ void frame_backward(const RawFrame& f, const M33& dL0, float mod,
                                               float ds_raw[3], float dq_raw[4]) {
  M33 dR;
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    float dsm = 0.f;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      dsm = fmaf(dL0.m[i][j], f.R.m[i][j], dsm);
      dR.m[i][j] = dL0.m[i][j] * f.sm[j];
    }
    ds_raw[j] = dsm * mod * f.s[j];   // through mod * exp(raw)
  }
  const float r = f.q[0], x = f.q[1], y = f.q[2], z = f.q[3];
  float dq[4];
  dq[0] = 2.f * (-z * dR.m[0][1] + y * dR.m[0][2] + z * dR.m[1][0] - x * dR.m[1][2] - y * dR.m[2][0] + x * dR.m[2][1]);
  dq[1] = 2.f * (y * dR.m[0][1] + z * dR.m[0][2] + y * dR.m[1][0] - 2.f * x * dR.m[1][1] - r * dR.m[1][2] +
                 z * dR.m[2][0] + r * dR.m[2][1] - 2.f * x * dR.m[2][2]);
  dq[2] = 2.f * (-2.f * y * dR.m[0][0] + x * dR.m[0][1] + r * dR.m[0][2] + x * dR.m[1][0] + z * dR.m[1][2] -
                 r * dR.m[2][0] + z * dR.m[2][1] - 2.f * y * dR.m[2][2]);
  dq[3] = 2.f * (-2.f * z * dR.m[0][0] - r * dR.m[0][1] + x * dR.m[0][2] + r * dR.m[1][0] - 2.f * z * dR.m[1][1] +
                 y * dR.m[1][2] + x * dR.m[2][0] + y * dR.m[2][1]);
  // q = raw / |raw|
  const float qdq = f.q[0] * dq[0] + f.q[1] * dq[1] + f.q[2] * dq[2] + f.q[3] * dq[3];
  const float inv = 1.f / f.norm;
#pragma unroll
  for (int k = 0; k < 4; ++k) dq_raw[k] = (dq[k] - f.q[k] * qdq) * inv;
}

// Device view of EgsModelArgs, passed to kernels by value.
struct RawModel {
  int P;
  const float* scaling_raw;
  const float* rotation_raw;
  const float* opacity_raw;
  const float* features_dc;
  const float* features_rest;
  float mod;
  const int* is_object;
  int which;
  const float* accum_R;
  const float* train_R;
  int row0_copies;
};

inline RawModel raw_model_from(const EgsModelArgs* m) {
  RawModel k;
  k.P = m->P;
  k.scaling_raw = m->scaling_raw;
  k.rotation_raw = m->rotation_raw;
  k.opacity_raw = m->opacity_raw;
  k.features_dc = m->features_dc;
  k.features_rest = m->features_rest;
  k.mod = m->scale_modifier;
  k.is_object = m->is_object;
  k.which = m->which_object;
  k.accum_R = m->accum_R;
  k.train_R = m->train_R;
  k.row0_copies = (m->accum_R || m->train_R) ? m->row0_copies : 0;
  return k;
}

// Selection of build_covariance_from_scaling_rotation_w_rot (gaussian_model.py:46-63). The mask
// tensor is [P,1], so `nonzero(mask).squeeze()` holds (row, 0) PAIRS and indexing L with it touches
// row 0 once per selected row as well: row 0 is rotated whenever any row is. `row0_copies` (> 0
// when that happens) is how many index entries name row 0; the host counts it (cabi / binding).
__device__ __forceinline__ bool row_selected(const RawModel& a, int i) {
  if (a.accum_R == nullptr && a.train_R == nullptr) return false;
  if (i == 0 && a.row0_copies > 0) return true;
  return a.is_object == nullptr || a.is_object[i] == a.which;
}

// The two optional 3x3 rotations, staged once per CTA: s_rot[0..9) = accum_R, [9..18) = train_R.
__device__ __forceinline__ void stage_rotations(const RawModel& a, float* s_rot) {
  if (threadIdx.x < 9) s_rot[threadIdx.x] = a.accum_R ? a.accum_R[threadIdx.x] : 0.f;
  else if (threadIdx.x < 18) s_rot[threadIdx.x] = a.train_R ? a.train_R[threadIdx.x - 9] : 0.f;
}

// Forward of one row: frame, final factor, six covariance entries. The second form takes the raw
// scale and quaternion from registers (callers that issue their loads early).
__device__ __forceinline__ void raw_covariance_from(const RawModel& a, int i, const float s_raw[3], const float4 q_raw,
                                                    const float* s_rot, float cov6[6]);
__device__ __forceinline__ void raw_covariance(const RawModel& a, int i, const float* s_rot, float cov6[6]) {
  const float s_raw[3] = {a.scaling_raw[3 * i], a.scaling_raw[3 * i + 1], a.scaling_raw[3 * i + 2]};
  const float4 q_raw = reinterpret_cast<const float4*>(a.rotation_raw)[i];
  raw_covariance_from(a, i, s_raw, q_raw, s_rot, cov6);
}
__device__ __forceinline__ void raw_covariance_from(const RawModel& a, int i, const float s_raw[3], const float4 q_raw,
                                                    const float* s_rot, float cov6[6]) {
  RawFrame f;
  raw_frame(s_raw, q_raw, a.mod, f);
  M33 L = frame_factor(f);
  if (row_selected(a, i)) {
    const M33 A = m33_load(s_rot), T = m33_load(s_rot + 9);
    L = rotate_factor(L, a.accum_R ? &A : nullptr, a.train_R ? &T : nullptr, nullptr);
  }
  factor_to_cov6(L, cov6);
}

// Backward of one row: d6 = dL/d(six covariance entries) -> raw scaling / quaternion gradients and
// this row's contribution dT[9] to dL/d train_R (zero when the row is not rotated by train_R).
__device__ __forceinline__ void raw_covariance_backward(const RawModel& a, int i, const float* s_rot,
                                                        const float d6[6], float ds[3], float dq[4], float dT[9]) {
  const float s_raw[3] = {a.scaling_raw[3 * i], a.scaling_raw[3 * i + 1], a.scaling_raw[3 * i + 2]};
  const float4 q_raw = reinterpret_cast<const float4*>(a.rotation_raw)[i];
  RawFrame f;
  raw_frame(s_raw, q_raw, a.mod, f);
  const M33 L0 = frame_factor(f);
  M33 dL0;
  if (row_selected(a, i)) {
    const M33 A = m33_load(s_rot), T = m33_load(s_rot + 9);
    M33 mid;
    const M33 Lf = rotate_factor(L0, a.accum_R ? &A : nullptr, a.train_R ? &T : nullptr, &mid);
    M33 g = cov6_backward_to_factor(d6, Lf);
    // every index entry that names this row receives the row's full gradient (index_put with
    // duplicate indices): row 0 counts `row0_copies` times, any other selected row once
    const float copies = (i == 0 && a.row0_copies > 0) ? (float)a.row0_copies : 1.f;
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
      for (int c = 0; c < 3; ++c) g.m[r][c] *= copies;
    if (a.train_R) {
      const M33 gT = m33_mult(g, mid);   // dT = g mid^T
#pragma unroll
      for (int k = 0; k < 9; ++k) dT[k] = gT.m[k / 3][k % 3];
      g = m33_tmul(T, g);
    }
    if (a.accum_R) g = m33_tmul(A, g);
    dL0 = g;
  } else {
    dL0 = cov6_backward_to_factor(d6, L0);
  }
  frame_backward(f, dL0, a.mod, ds, dq);
}

// Sum dT over the CTA and add it to d_train_R (device [9]) with nine atomics. Every thread of the
// CTA must call this. `s_part` is [blockDim.x / 32][9] floats of shared memory.
__device__ __forceinline__ void reduce_pose_gradient(const float dT[9], float (*s_part)[9], float* d_train_R) {
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5, warps = blockDim.x >> 5;
#pragma unroll
  for (int k = 0; k < 9; ++k) {
    float v = dT[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) s_part[warp][k] = v;
  }
  __syncthreads();
  if (threadIdx.x < 9) {
    float v = 0.f;
    for (unsigned w = 0; w < warps; ++w) v += s_part[w][threadIdx.x];
    if (v != 0.f) atomicAdd(d_train_R + threadIdx.x, v);
  }
}

// Warp-wide copy between one contiguous global run of `count` floats, holding rows of `width`
// floats back to back, and a pitched shared-memory tile: element e of the run is
// tile[(e / width) * pitch + col0 + e % width]. Used to stage the split features_dc / features_rest
// rows (width 3 and 3M - 3) next to each other in one SH row per Gaussian, and to write the
// gradient rows back. float4 global accesses when the run is 16-byte aligned, all loads of an
// unrolled group issued before their stores (a one-load-per-iteration loop exposes the full DRAM
// latency 45 times per warp); e / width by reciprocal multiplication, exact for e < 2^20 / width.
template <bool TO_TILE>
__device__ __forceinline__ void warp_rows_copy(float* tile, int pitch, int col0, int width, float* run, int count,
                                               unsigned lane) {
  const uint32_t inv = ((1u << 20) + (uint32_t)width - 1u) / (uint32_t)width;
  auto offset_of = [&](int e, int& k) -> int {   // tile offset of element e; k = its column in the row
    const int r = (int)(((uint32_t)e * inv) >> 20);
    k = e - r * width;
    return r * pitch + col0 + k;
  };
  int done = 0;
  if ((reinterpret_cast<uintptr_t>(run) & 15u) == 0) {
    const int nvec = count >> 2;
    float4* run4 = reinterpret_cast<float4*>(run);
    constexpr int GROUP = 4;   // float4 loads issued back to back before any of them is used
    for (int f0 = (int)lane; f0 < nvec; f0 += 32 * GROUP) {
      float4 v[GROUP];
      if (TO_TILE) {
#pragma unroll
        for (int j = 0; j < GROUP; ++j)
          if (f0 + 32 * j < nvec) v[j] = __ldg(run4 + f0 + 32 * j);
      }
#pragma unroll
      for (int j = 0; j < GROUP; ++j) {
        const int f = f0 + 32 * j;
        if (f >= nvec) break;
        // one division for the first component, the other three step along (and across) the rows
        int k;
        int o0 = offset_of(4 * f, k), o1, o2, o3;
        const int wrap = pitch - width;   // extra tile offset when a component starts the next row
        ++k; o1 = o0 + 1; if (k >= width) { k -= width; o1 += wrap; }
        ++k; o2 = o1 + 1; if (k >= width) { k -= width; o2 += wrap; }
        ++k; o3 = o2 + 1; if (k >= width) { k -= width; o3 += wrap; }
        if (TO_TILE) {
          tile[o0] = v[j].x; tile[o1] = v[j].y; tile[o2] = v[j].z; tile[o3] = v[j].w;
        } else {
          run4[f] = make_float4(tile[o0], tile[o1], tile[o2], tile[o3]);
        }
      }
    }
    done = nvec << 2;
  }
#pragma unroll 4
  for (int e = done + (int)lane; e < count; e += 32) {
    int k;
    const int o = offset_of(e, k);
    if (TO_TILE) tile[o] = __ldg(run + e);
    else run[e] = tile[o];
  }
}

// Staging of such rows with asynchronous copies (cp.async, 4 bytes each, global -> shared without a
// register in between): a lane queues all of its elements at once, so the whole run of a warp is in
// flight together and the queue depth does not cost registers; the odd pitch keeps the later
// row-wise reads conflict-free (16-byte copies would need 16-byte aligned rows). Call
// stage_rows_wait() before the tile is read.
__device__ __forceinline__ void stage_rows_async(float* tile, int pitch, int col0, int width, const float* run,
                                                 int count, unsigned lane) {
  const uint32_t inv = ((1u << 20) + (uint32_t)width - 1u) / (uint32_t)width;
  const uint32_t base = (uint32_t)__cvta_generic_to_shared(tile);
#pragma unroll 4
  for (int e = (int)lane; e < count; e += 32) {
    const int r = (int)(((uint32_t)e * inv) >> 20);
    const uint32_t dst = base + 4u * (uint32_t)(r * pitch + col0 + (e - r * width));
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(run + e) : "memory");
  }
}
// The same for rows whose length is a multiple of 4 floats and whose run is 16-byte aligned (the
// [P,16,3] SH tensor: 192-byte rows): 16-byte copies into a tile whose pitch is a multiple of 4 floats,
// a quarter of the copy instructions, and the rows can then be read with 16-byte shared loads (pitch 52:
// the eight lanes of a quarter warp hit eight different 16-byte bank groups).
__device__ __forceinline__ void stage_rows_async16(float* tile, int pitch, int width, const float* run, int count,
                                                   unsigned lane) {
  const int chunks_per_row = width >> 2;
  const uint32_t inv = ((1u << 20) + (uint32_t)chunks_per_row - 1u) / (uint32_t)chunks_per_row;
  const uint32_t base = (uint32_t)__cvta_generic_to_shared(tile);
  const int chunks = count >> 2;
#pragma unroll 4
  for (int e = (int)lane; e < chunks; e += 32) {
    const int r = (int)(((uint32_t)e * inv) >> 20);
    const uint32_t dst = base + 4u * (uint32_t)(r * pitch + 4 * (e - r * chunks_per_row));
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(run + 4 * e) : "memory");
  }
}
// a contiguous run of `count` floats (a multiple of 4, 16-byte aligned) to a contiguous tile
__device__ __forceinline__ void stage_run_async16(float* tile, const float* run, int count, unsigned lane) {
  const uint32_t base = (uint32_t)__cvta_generic_to_shared(tile);
  for (int e = (int)lane; e < (count >> 2); e += 32)
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(base + 16u * (uint32_t)e), "l"(run + 4 * e) : "memory");
}
__device__ __forceinline__ void stage_rows_wait() {
  asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
  __syncwarp();
}

// torch.sigmoid and its autograd formula grad * (1 - y) * y
__device__ __forceinline__ float sigmoid_fwd(float x) { return 1.f / (1.f + expf(-x)); }
__device__ __forceinline__ float sigmoid_bwd(float y, float g) { return g * (1.f - y) * y; }

}  // namespace egs
