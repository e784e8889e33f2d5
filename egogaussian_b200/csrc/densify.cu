// densify.cu -- the growing half of densification: what
// GaussianModel.densify_and_clone / densify_and_split do with boolean-mask indexing, repeat() and
// torch.cat per tensor, and cat_tensors_to_optimizer with the Adam moments
// (scene/gaussian_model.py:243-272, 559-677). The pruning half is compact.cu.
//
//   select   grads = xyz_gradient_accum / denom (NaN -> 0), selected = grads >= max_grad and the
//            largest scale on the clone / split side of percent_dense * extent (and, optionally, the
//            object id and the generation): one pass, one byte per row
//   plan     egs_compact_plan on that mask: the selected rows in order and their count (the one host
//            read: the outputs have to be sized)
//   append   every tensor at once: dst = cat(src, src[selected].repeat(copies)) -- or zeros / a constant
//            in the new rows (Adam moments; the generation stamp) -- one launch, 4-byte words
//   split    the new rows' positions and scales: xyz + R(q / |q|) (noise * s), log(s / (0.8 copies)),
//            with the standard-normal noise supplied by the caller (torch.normal(0, s) = s * noise)
#include "common.cuh"

#include <math.h>

namespace egs {

__global__ void __launch_bounds__(256)
densify_select_kernel(EgsDensifySelect a, unsigned char* __restrict__ selected) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= a.P) return;
  float g = 0.f;
  if (i < a.n_graded) {   // rows appended since the statistic was taken count as zero (gaussian_model.py:604-605)
    g = a.grad_accum[i] / a.denom[i];
    if (isnan(g)) g = 0.f;   // grads[grads.isnan()] = 0.0 (:690)
  }
  bool sel = g >= a.grad_threshold;
  const float s = fmaxf(fmaxf(expf(a.scaling_raw[3 * i]), expf(a.scaling_raw[3 * i + 1])), expf(a.scaling_raw[3 * i + 2]));
  sel = sel && (a.split ? s > a.size_threshold : s <= a.size_threshold);
  if (a.is_object != nullptr) sel = sel && a.is_object[i] == a.which_object;
  if (a.generation != nullptr) sel = sel && a.generation[i] == a.only_generation;
  selected[i] = sel ? 1 : 0;
}

struct AppendSpans {
  const uint32_t* src[EGS_COMPACT_MAX_TENSORS];
  uint32_t* dst[EGS_COMPACT_MAX_TENSORS];
  int row_words[EGS_COMPACT_MAX_TENSORS];
  int fill[EGS_COMPACT_MAX_TENSORS];
  uint32_t value[EGS_COMPACT_MAX_TENSORS];
};

// blockIdx.y = tensor; a CTA copies 256 output rows of it, one 4-byte word per thread per trip. Output
// row r < P is source row r; row P + n * count + k is the k-th selected row (x[mask].repeat(copies, 1)).
__global__ void __launch_bounds__(256)
append_rows_kernel(AppendSpans sp, int P, uint32_t count, int copies, const uint32_t* __restrict__ src_of) {
  const int t = blockIdx.y;
  const int rw = sp.row_words[t];
  const uint32_t* __restrict__ src = sp.src[t];
  uint32_t* __restrict__ dst = sp.dst[t];
  const int fill = sp.fill[t];
  const uint32_t value = sp.value[t];
  const uint32_t total = (uint32_t)P + (uint32_t)copies * count;
  for (uint32_t row0 = blockIdx.x * 256u; row0 < total; row0 += gridDim.x * 256u) {
    const uint32_t rows = min(256u, total - row0);
    const uint32_t words = rows * (uint32_t)rw;
    for (uint32_t w = threadIdx.x; w < words; w += 256) {
      const uint32_t r = w / (uint32_t)rw, c = w - r * (uint32_t)rw;
      const uint32_t row = row0 + r;
      uint32_t v;
      if (row < (uint32_t)P) v = src[(size_t)row * rw + c];
      else if (fill == 1) v = 0u;
      else if (fill == 2) v = value;
      else v = src[(size_t)src_of[(row - (uint32_t)P) % count] * rw + c];
      dst[(size_t)row * rw + c] = v;
    }
  }
}

// build_rotation normalises the quaternion (utils/general_utils.py:124-149); torch.bmm(rots, samples)
__global__ void __launch_bounds__(256)
split_samples_kernel(uint32_t count, int copies, const uint32_t* __restrict__ src_of, const float* __restrict__ xyz,
                     const float* __restrict__ scaling_raw, const float* __restrict__ rotation_raw,
                     const float* __restrict__ noise, float* __restrict__ xyz_new, float* __restrict__ scaling_new) {
  const uint32_t j = blockIdx.x * 256u + threadIdx.x;   // new row
  if (j >= count * (uint32_t)copies) return;
  const uint32_t i = src_of[j % count];
  const float sx = expf(scaling_raw[3 * i]), sy = expf(scaling_raw[3 * i + 1]), sz = expf(scaling_raw[3 * i + 2]);
  const float ax = noise[3 * j] * sx, ay = noise[3 * j + 1] * sy, az = noise[3 * j + 2] * sz;
  float r = rotation_raw[4 * i], x = rotation_raw[4 * i + 1], y = rotation_raw[4 * i + 2], z = rotation_raw[4 * i + 3];
  const float n = sqrtf(r * r + x * x + y * y + z * z);
  r /= n; x /= n; y /= n; z /= n;
  const float R00 = 1.f - 2.f * (y * y + z * z), R01 = 2.f * (x * y - r * z), R02 = 2.f * (x * z + r * y);
  const float R10 = 2.f * (x * y + r * z), R11 = 1.f - 2.f * (x * x + z * z), R12 = 2.f * (y * z - r * x);
  const float R20 = 2.f * (x * z - r * y), R21 = 2.f * (y * z + r * x), R22 = 1.f - 2.f * (x * x + y * y);
  xyz_new[3 * j] = R00 * ax + R01 * ay + R02 * az + xyz[3 * i];
  xyz_new[3 * j + 1] = R10 * ax + R11 * ay + R12 * az + xyz[3 * i + 1];
  xyz_new[3 * j + 2] = R20 * ax + R21 * ay + R22 * az + xyz[3 * i + 2];
  const float shrink = 0.8f * (float)copies;
  scaling_new[3 * j] = logf(sx / shrink);
  scaling_new[3 * j + 1] = logf(sy / shrink);
  scaling_new[3 * j + 2] = logf(sz / shrink);
}

}  // namespace egs

extern "C" int egs_densify_select(const EgsDensifySelect* args, unsigned char* selected, void* stream) {
  using namespace egs;
  if (!args || args->P < 0 || (args->P > 0 && (!selected || !args->scaling_raw ||
                                                (args->n_graded > 0 && (!args->grad_accum || !args->denom))))) {
    set_error("egs_densify_select: NULL argument or negative P");
    return EGS_ERR_INVALID_ARGUMENT;
  }
  if (args->P == 0) return EGS_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  densify_select_kernel<<<(args->P + 255) / 256, 256, 0, st>>>(*args, selected);
  EGS_AFTER_LAUNCH(st, false);
  return EGS_OK;
}

extern "C" int egs_append_rows(int P, const void* plan_scratch, int count, int copies, const EgsAppendTensor* tensors,
                               int n_tensors, void* stream) {
  using namespace egs;
  if (P < 0 || count < 0 || copies < 1 || n_tensors < 0 || n_tensors > EGS_COMPACT_MAX_TENSORS ||
      (n_tensors > 0 && !tensors) || (count > 0 && !plan_scratch)) {
    set_error("egs_append_rows: bad argument (at most %d tensors per call, copies >= 1)", EGS_COMPACT_MAX_TENSORS);
    return EGS_ERR_INVALID_ARGUMENT;
  }
  if (n_tensors == 0 || P + count == 0) return EGS_OK;
  AppendSpans sp;
  for (int t = 0; t < n_tensors; ++t) {
    const EgsAppendTensor& x = tensors[t];
    if ((P > 0 && !x.src) || !x.dst || x.row_bytes <= 0 || (x.row_bytes & 3) || x.fill < 0 || x.fill > 2 ||
        (reinterpret_cast<uintptr_t>(x.src) & 3u) || (reinterpret_cast<uintptr_t>(x.dst) & 3u)) {
      set_error("egs_append_rows: tensor %d needs 4-byte aligned src/dst, row_bytes %% 4 == 0 and fill in 0..2", t);
      return EGS_ERR_INVALID_ARGUMENT;
    }
    sp.src[t] = static_cast<const uint32_t*>(x.src);
    sp.dst[t] = static_cast<uint32_t*>(x.dst);
    sp.row_words[t] = x.row_bytes / 4;
    sp.fill[t] = x.fill;
    sp.value[t] = (uint32_t)x.value;
  }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int nb = (P + 1023) / 1024;   // layout of egs_compact_plan's scratch: total, block sums, src_of
  const uint32_t* src_of = count > 0 ? static_cast<const uint32_t*>(plan_scratch) + 1 + nb : nullptr;
  const dim3 grid(148 * 2, n_tensors, 1);
  append_rows_kernel<<<grid, 256, 0, st>>>(sp, P, (uint32_t)count, copies, src_of);
  EGS_AFTER_LAUNCH(st, false);
  return EGS_OK;
}

extern "C" int egs_split_samples(int P, const void* plan_scratch, int count, int copies, const float* xyz,
                                 const float* scaling_raw, const float* rotation_raw, const float* noise, float* xyz_new,
                                 float* scaling_new, void* stream) {
  using namespace egs;
  if (P < 0 || count < 0 || copies < 1 ||
      (count > 0 && (!plan_scratch || !xyz || !scaling_raw || !rotation_raw || !noise || !xyz_new || !scaling_new))) {
    set_error("egs_split_samples: NULL argument or bad size");
    return EGS_ERR_INVALID_ARGUMENT;
  }
  if (count == 0) return EGS_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int nb = (P + 1023) / 1024;
  const uint32_t* src_of = static_cast<const uint32_t*>(plan_scratch) + 1 + nb;
  const uint32_t n_new = (uint32_t)count * (uint32_t)copies;
  split_samples_kernel<<<(n_new + 255) / 256, 256, 0, st>>>((uint32_t)count, copies, src_of, xyz, scaling_raw, rotation_raw,
                                                           noise, xyz_new, scaling_new);
  EGS_AFTER_LAUNCH(st, false);
  return EGS_OK;
}
