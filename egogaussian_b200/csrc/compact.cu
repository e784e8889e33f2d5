// compact.cu -- stable row compaction of many tensors at once (the periodic
// half): what GaussianModel.prune_points / _prune_optimizer do with boolean-mask indexing
// (scene/gaussian_model.py:506-557): every parameter, both of its Adam moments and the per-Gaussian
// bookkeeping tensors are cut down to the rows with keep[i] != 0, in order. In torch that is one
// nonzero() + gather (and one host sync) per tensor, ~26 times per prune; here the mask is scanned
// once (one host read: the new row count, needed to size the outputs) and all tensors are gathered
// by one launch with coalesced 4-byte accesses.
#include "common.cuh"

namespace egs {

constexpr int CP_BLOCK = 1024;            // rows per CTA in the counting / scatter kernels
constexpr int CP_MAX_SPANS = EGS_COMPACT_MAX_TENSORS;

// scratch layout (uint32): [0] total, [1 .. 1 + nblocks) block sums -> exclusive block offsets,
// then src_of[P] (kept row k came from row src_of[k])
__device__ __forceinline__ uint32_t* cp_block_sums(uint32_t* s) { return s + 1; }
__device__ __forceinline__ uint32_t* cp_src_of(uint32_t* s, int nblocks) { return s + 1 + nblocks; }

__global__ void __launch_bounds__(CP_BLOCK)
compact_count_kernel(int P, const unsigned char* __restrict__ keep, uint32_t* __restrict__ scratch) {
  __shared__ uint32_t s_w[CP_BLOCK / 32];
  const int i = blockIdx.x * CP_BLOCK + threadIdx.x;
  const bool k = i < P && keep[i] != 0;
  const uint32_t b = __ballot_sync(0xffffffffu, k);
  if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5] = __popc(b);
  __syncthreads();
  if (threadIdx.x < 32) {
    uint32_t v = s_w[threadIdx.x];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (threadIdx.x == 0) cp_block_sums(scratch)[blockIdx.x] = v;
  }
}

// one CTA: exclusive scan of the block sums in place, total to scratch[0]
__global__ void __launch_bounds__(1024)
compact_scan_kernel(int nblocks, uint32_t* __restrict__ scratch) {
  __shared__ uint32_t s_w[32];
  __shared__ uint32_t s_carry;
  uint32_t* sums = cp_block_sums(scratch);
  if (threadIdx.x == 0) s_carry = 0;
  __syncthreads();
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int base = 0; base < nblocks; base += 1024) {
    const int i = base + (int)threadIdx.x;
    const uint32_t v = i < nblocks ? sums[i] : 0u;
    uint32_t incl = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += t;
    }
    if (lane == 31) s_w[warp] = incl;
    __syncthreads();
    uint32_t woff = 0;
    for (unsigned w = 0; w < warp; ++w) woff += s_w[w];
    const uint32_t carry = s_carry;
    if (i < nblocks) sums[i] = carry + woff + incl - v;
    __syncthreads();
    if (threadIdx.x == 1023) s_carry = carry + woff + incl;
    __syncthreads();
  }
  if (threadIdx.x == 0) scratch[0] = s_carry;
}

__global__ void __launch_bounds__(CP_BLOCK)
compact_scatter_kernel(int P, const unsigned char* __restrict__ keep, uint32_t* __restrict__ scratch, int nblocks) {
  __shared__ uint32_t s_w[CP_BLOCK / 32];
  const int i = blockIdx.x * CP_BLOCK + threadIdx.x;
  const bool k = i < P && keep[i] != 0;
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t b = __ballot_sync(0xffffffffu, k);
  if (lane == 0) s_w[warp] = __popc(b);
  __syncthreads();
  uint32_t before = 0;
  for (unsigned w = 0; w < warp; ++w) before += s_w[w];
  if (k) {
    const uint32_t pos = cp_block_sums(scratch)[blockIdx.x] + before + __popc(b & ((1u << lane) - 1u));
    cp_src_of(scratch, nblocks)[pos] = (uint32_t)i;
  }
}

struct CompactSpans {
  const uint32_t* src[CP_MAX_SPANS];
  uint32_t* dst[CP_MAX_SPANS];
  int row_words[CP_MAX_SPANS];
  int n;
};

// blockIdx.y = tensor; a CTA copies 256 output rows of it, one 4-byte word per thread per trip
__global__ void __launch_bounds__(256)
compact_gather_kernel(CompactSpans sp, const uint32_t* __restrict__ scratch, int nblocks) {
  const uint32_t count = scratch[0];
  const uint32_t* src_of = scratch + 1 + nblocks;
  const int t = blockIdx.y;
  const int rw = sp.row_words[t];
  const uint32_t* __restrict__ src = sp.src[t];
  uint32_t* __restrict__ dst = sp.dst[t];
  for (uint32_t row0 = blockIdx.x * 256u; row0 < count; row0 += gridDim.x * 256u) {
    const uint32_t rows = min(256u, count - row0);
    const uint32_t words = rows * (uint32_t)rw;
    for (uint32_t w = threadIdx.x; w < words; w += 256) {
      const uint32_t r = w / (uint32_t)rw, c = w - r * (uint32_t)rw;
      dst[(size_t)(row0 + r) * rw + c] = src[(size_t)src_of[row0 + r] * rw + c];
    }
  }
}

static int cp_blocks(int P) { return (P + CP_BLOCK - 1) / CP_BLOCK; }

}  // namespace egs

extern "C" size_t egs_compact_scratch_bytes(int P) {
  if (P < 0) P = 0;
  return ((size_t)1 + egs::cp_blocks(P) + (size_t)P) * sizeof(uint32_t) + 256;
}

extern "C" int egs_compact_plan(int P, const unsigned char* keep, void* scratch, size_t scratch_bytes, int* count,
                                void* stream) {
  using namespace egs;
  if (P < 0 || !count || (P > 0 && (!keep || !scratch)) || scratch_bytes < egs_compact_scratch_bytes(P)) {
    set_error("egs_compact_plan: NULL argument, negative P or scratch too small");
    return EGS_ERR_INVALID_ARGUMENT;
  }
  *count = 0;
  if (P == 0) return EGS_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  uint32_t* s = static_cast<uint32_t*>(scratch);
  const int nb = cp_blocks(P);
  compact_count_kernel<<<nb, CP_BLOCK, 0, st>>>(P, keep, s);
  EGS_AFTER_LAUNCH(st, false);
  compact_scan_kernel<<<1, 1024, 0, st>>>(nb, s);
  EGS_AFTER_LAUNCH(st, false);
  compact_scatter_kernel<<<nb, CP_BLOCK, 0, st>>>(P, keep, s, nb);
  EGS_AFTER_LAUNCH(st, false);
  uint32_t total = 0;
  EGS_CUDA_TRY(cudaMemcpyAsync(&total, s, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
  EGS_CUDA_TRY(cudaStreamSynchronize(st));
  *count = (int)total;
  return EGS_OK;
}

extern "C" int egs_compact_rows(int P, const void* scratch, const EgsRowTensor* tensors, int n_tensors, void* stream) {
  using namespace egs;
  if (P < 0 || n_tensors < 0 || n_tensors > CP_MAX_SPANS || (n_tensors > 0 && !tensors) || (P > 0 && !scratch)) {
    set_error("egs_compact_rows: bad argument (at most %d tensors per call)", CP_MAX_SPANS);
    return EGS_ERR_INVALID_ARGUMENT;
  }
  if (P == 0 || n_tensors == 0) return EGS_OK;
  CompactSpans sp;
  sp.n = n_tensors;
  for (int t = 0; t < n_tensors; ++t) {
    if (!tensors[t].src || !tensors[t].dst || tensors[t].row_bytes <= 0 || (tensors[t].row_bytes & 3) != 0 ||
        (reinterpret_cast<uintptr_t>(tensors[t].src) & 3u) || (reinterpret_cast<uintptr_t>(tensors[t].dst) & 3u)) {
      set_error("egs_compact_rows: tensor %d needs non-NULL 4-byte aligned src/dst and row_bytes %% 4 == 0", t);
      return EGS_ERR_INVALID_ARGUMENT;
    }
    sp.src[t] = static_cast<const uint32_t*>(tensors[t].src);
    sp.dst[t] = static_cast<uint32_t*>(tensors[t].dst);
    sp.row_words[t] = tensors[t].row_bytes / 4;
  }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const dim3 grid(148 * 2, n_tensors, 1);
  compact_gather_kernel<<<grid, 256, 0, st>>>(sp, static_cast<const uint32_t*>(scratch), cp_blocks(P));
  EGS_AFTER_LAUNCH(st, false);
  return EGS_OK;
}
