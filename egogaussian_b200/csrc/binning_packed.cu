// binning_packed.cu -- the fast binning path: (tile | depth | index) packed in ONE 64-bit word,
// sorted by a stable LSD "onesweep" radix sort with 9-bit digits and coalesced scatters.
//
// Replaces duplicateWithKeys + cub::DeviceRadixSort::SortPairs of the reference
// (DGR/cuda_rasterizer/rasterizer_impl.cu:70-111, 301-309) with the same result:
//   * the reference sorts (u64 key, u32 value) pairs on key bits [0, 32+bit): 12 B per pair read
//     and 12 B written per pass, 6 passes of 8 bits at 1280x720;
//   * here the Gaussian index rides in the low `ib` bits of the same word as the key,
//       word = ((tile << 32 | depth bits) << ib) | index,        ib = bit_length(P - 1),
//     whenever ib + 32 + bit <= 64 (e.g. 20 + 32 + 12 at 1M Gaussians, 1280x720). Passes run over
//     word bits [ib, ib + 32 + bit) only, so ties keep generation (= index) order exactly like the
//     reference's stable sort -- and each pass moves 8 B per pair each way instead of 12 B;
//   * 9-bit digits: 5 passes for 44 key bits instead of 6;
//   * inside a block the ranked items are first exchanged through shared memory, so that every
//     warp then stores runs of consecutive addresses (direct scatter from registers made each
//     store instruction touch up to 32 sectors and halved the throughput);
//   * the last pass unpacks: it writes the u32 index list the blend kernels read and the 64-bit
//     keys (tile | depth) that identifyTileRanges and the parity tests read.
// Scenes whose indices do not fit (P > 2^(32 - bit)) take the pair-sorting path in binning.cu.
#include "common.cuh"

#include <stddef.h>

namespace egs {

constexpr int PK_BITS = 9;
constexpr int PK_RADIX = 1 << PK_BITS;   // 512
constexpr int PK_THREADS = 512;
constexpr int PK_WARPS = PK_THREADS / 32;
constexpr int PK_ITEMS = 8;
constexpr int PK_TILE = PK_THREADS * PK_ITEMS;   // 4096 words per block
constexpr int PK_MAX_PASSES = 6;
static_assert(PK_TILE == SORT_TILE, "status array sizing assumes the same block tile as the pair sort");
static_assert(PK_MAX_PASSES * PK_RADIX <= SORT_STATUS_WORDS, "status array too small");

__global__ void __launch_bounds__(256)
duplicate_packed_kernel(int P, const float2* __restrict__ means2D, const float* __restrict__ depths,
                        const uint32_t* __restrict__ offsets, const int* __restrict__ radii,
                        const uint32_t* __restrict__ order /* NULL: slot j of the scan is Gaussian j */,
                        uint64_t* __restrict__ words, int grid_x, int grid_y, int idx_bits) {
  const int slot = blockIdx.x * blockDim.x + threadIdx.x;
  if (slot >= P) return;
  const int idx = order ? (int)order[slot] : slot;
  const int r = radii[idx];
  if (r <= 0) return;
  uint32_t off = (slot == 0) ? 0u : offsets[slot - 1];
  uint2 lo, hi;
  tile_rect(means2D[idx], r, grid_x, grid_y, lo, hi);
  const uint64_t low = ((uint64_t)__float_as_uint(depths[idx]) << idx_bits) | (uint32_t)idx;
  const int tile_shift = 32 + idx_bits;
  for (uint32_t y = lo.y; y < hi.y; ++y)
    for (uint32_t x = lo.x; x < hi.x; ++x)
      words[off++] = ((uint64_t)(y * (uint32_t)grid_x + x) << tile_shift) | low;
}

template <int PASSES>
__global__ void __launch_bounds__(256)
packed_histogram_kernel(const uint64_t* __restrict__ words, int n, int begin_bit, int end_bit,
                        uint32_t* __restrict__ hist) {
  __shared__ uint32_t s_hist[PASSES * PK_RADIX];
  for (int i = threadIdx.x; i < PASSES * PK_RADIX; i += blockDim.x) s_hist[i] = 0;
  __syncthreads();
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < (size_t)n;
       i += (size_t)gridDim.x * blockDim.x) {
    const uint64_t w = words[i];
#pragma unroll
    for (int p = 0; p < PASSES; ++p) {
      const int shift = begin_bit + p * PK_BITS;
      const int width = min(PK_BITS, end_bit - shift);
      const uint32_t d = (uint32_t)(w >> shift) & ((1u << width) - 1u);
      atomicAdd(&s_hist[p * PK_RADIX + d], 1u);
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < PASSES * PK_RADIX; i += blockDim.x) {
    const uint32_t c = s_hist[i];
    if (c) atomicAdd(&hist[i], c);
  }
}

// exclusive scan of each pass's 512-bin histogram, in place (one block per pass)
__global__ void __launch_bounds__(PK_RADIX)
packed_scan_hist_kernel(uint32_t* __restrict__ hist) {
  __shared__ uint32_t s_w[PK_RADIX / 32];
  uint32_t* h = hist + blockIdx.x * PK_RADIX;
  const uint32_t v = h[threadIdx.x];
  const unsigned lane = lane_id(), warp = threadIdx.x >> 5;
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
  h[threadIdx.x] = woff + incl - v;
}

constexpr uint32_t PK_FLAG_AGG = 1u << 30;
constexpr uint32_t PK_FLAG_PREFIX = 1u << 31;
constexpr uint32_t PK_VALUE_MASK = (1u << 30) - 1u;

struct PackedSmem {
  uint32_t warp_hist[PK_WARPS][PK_RADIX];   // 32 KB
  uint64_t exch[PK_TILE];                   // 32 KB
  uint32_t count[PK_RADIX];                 // this block's digit histogram (published early)
  uint32_t digit_local[PK_RADIX];
  uint32_t digit_global[PK_RADIX];
  uint32_t warp_sums[PK_WARPS];
  uint32_t tile;
  uint32_t exch_v[PK_TILE];                 // 16 KB, pair mode only (kept last: not allocated otherwise)
};

// status words carry flag and value together, so relaxed device-scope accesses are enough
__device__ __forceinline__ uint32_t ld_status(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_status(uint32_t* p, uint32_t v) {
  asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// PAIRS = false: `in` holds packed words (key and index in one u64); the LAST pass unpacks.
// PAIRS = true : `in` holds bare keys and a u32 value array travels with them (scenes whose
//                indices do not fit beside the key); every pass moves both arrays.
template <bool LAST, bool PAIRS>
__global__ void __launch_bounds__(PK_THREADS)
packed_onesweep_kernel(const uint64_t* __restrict__ in, const uint32_t* __restrict__ vals_in,
                       uint64_t* __restrict__ out,
                       uint32_t* __restrict__ vals_out, int n, int shift, uint32_t digit_mask,
                       int idx_bits, const uint32_t* __restrict__ digit_base,
                       uint32_t* __restrict__ ticket, uint32_t* __restrict__ status_raw) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  PackedSmem& S = *reinterpret_cast<PackedSmem*>(smem_raw);
  const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) S.tile = atomicAdd(ticket, 1u);
#pragma unroll
  for (int w = 0; w < PK_WARPS; ++w) S.warp_hist[w][tid] = 0;
  S.count[tid] = 0;
  __syncthreads();
  const uint32_t tile = S.tile;
  const unsigned lt_mask = (1u << lane) - 1u;
  const size_t block_base = (size_t)tile * PK_TILE;
  const size_t warp_base = block_base + (size_t)warp * (32 * PK_ITEMS);

  uint64_t word[PK_ITEMS];
  uint32_t val[PAIRS ? PK_ITEMS : 1];
  uint32_t rank[PK_ITEMS];
#pragma unroll
  for (int i = 0; i < PK_ITEMS; ++i) {
    const size_t g = warp_base + (size_t)i * 32 + lane;
    word[i] = g < (size_t)n ? in[g] : ~0ull;
    if (PAIRS) val[i] = g < (size_t)n ? vals_in[g] : 0u;
  }
  // 1. block histogram (unordered shared-memory atomics) and EARLY publication of the aggregate,
  //    so that successors' look-backs find it while this block is still ranking
#pragma unroll
  for (int i = 0; i < PK_ITEMS; ++i) {
    const size_t g = warp_base + (size_t)i * 32 + lane;
    if (g < (size_t)n) atomicAdd(&S.count[(uint32_t)(word[i] >> shift) & digit_mask], 1u);
  }
  __syncthreads();
  uint32_t* st = status_raw + (size_t)tile * PK_RADIX + tid;
  const uint32_t block_count = S.count[tid];
  if (tile != 0) st_status(st, PK_FLAG_AGG | block_count);

  // 2. stable ranks inside each warp's contiguous segment
#pragma unroll
  for (int i = 0; i < PK_ITEMS; ++i) {
    const size_t g = warp_base + (size_t)i * 32 + lane;
    const bool ok = g < (size_t)n;
    const uint32_t d = (uint32_t)(word[i] >> shift) & digit_mask;
    const uint32_t peers = match_digit<PK_BITS>(d, ok);
    const uint32_t before = ok ? S.warp_hist[warp][d] : 0u;
    __syncwarp();
    if (ok && (peers & lt_mask) == 0u) S.warp_hist[warp][d] = before + __popc(peers);
    __syncwarp();
    rank[i] = before + __popc(peers & lt_mask);
  }
  __syncthreads();

  // 3. thread d owns digit d: warp offsets, block-local start, cross-block look-back
  {
    const uint32_t d = tid;
    uint32_t run = 0;
#pragma unroll
    for (int w = 0; w < PK_WARPS; ++w) {
      const uint32_t c = S.warp_hist[w][d];
      S.warp_hist[w][d] = run;
      run += c;
    }
    uint32_t incl = block_count;
#pragma unroll
    for (int s = 1; s < 32; s <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, incl, s);
      if (lane >= s) incl += t;
    }
    if (lane == 31) S.warp_sums[warp] = incl;
    __syncthreads();
    uint32_t woff = 0;
    for (unsigned w = 0; w < warp; ++w) woff += S.warp_sums[w];
    const uint32_t local_start = woff + incl - block_count;
    S.digit_local[d] = local_start;

    uint32_t excl = 0;
    if (tile != 0) {
      const uint32_t* prev = st - PK_RADIX;
      for (;;) {
        uint32_t s;
        do { s = ld_status(prev); } while ((s & (PK_FLAG_AGG | PK_FLAG_PREFIX)) == 0u);
        excl += s & PK_VALUE_MASK;
        if (s & PK_FLAG_PREFIX) break;
        prev -= PK_RADIX;
      }
    }
    st_status(st, PK_FLAG_PREFIX | (excl + block_count));
    S.digit_global[d] = digit_base[d] + excl - local_start;   // dst = digit_global[d] + local position
  }
  __syncthreads();

  // exchange: each item goes to its block-local sorted position
#pragma unroll
  for (int i = 0; i < PK_ITEMS; ++i) {
    const size_t g = warp_base + (size_t)i * 32 + lane;
    if (g < (size_t)n) {
      const uint32_t d = (uint32_t)(word[i] >> shift) & digit_mask;
      const uint32_t pos = S.digit_local[d] + S.warp_hist[warp][d] + rank[i];
      S.exch[pos] = word[i];
      if (PAIRS) S.exch_v[pos] = val[i];
    }
  }
  __syncthreads();

  const uint32_t valid = (uint32_t)min((size_t)PK_TILE, (size_t)n - block_base);
  const uint64_t idx_mask = (1ull << idx_bits) - 1ull;
#pragma unroll
  for (int m = 0; m < PK_ITEMS; ++m) {
    const uint32_t k = m * PK_THREADS + tid;
    if (k < valid) {
      const uint64_t w = S.exch[k];
      const uint32_t d = (uint32_t)(w >> shift) & digit_mask;
      const uint32_t dst = S.digit_global[d] + k;
      if (PAIRS) {
        out[dst] = w;
        vals_out[dst] = S.exch_v[k];
      } else if (LAST) {
        out[dst] = w >> idx_bits;                 // (tile << 32) | depth bits
        vals_out[dst] = (uint32_t)(w & idx_mask); // Gaussian index
      } else {
        out[dst] = w;
      }
    }
  }
}

template <int PASSES>
static void launch_packed_hist(const uint64_t* words, int n, int begin_bit, int end_bit, uint32_t* hist,
                               cudaStream_t stream) {
  int blocks = (n + 256 * 16 - 1) / (256 * 16);
  blocks = blocks < 1 ? 1 : (blocks > 148 * 8 ? 148 * 8 : blocks);
  packed_histogram_kernel<PASSES><<<blocks, 256, 0, stream>>>(words, n, begin_bit, end_bit, hist);
}

bool packed_path_fits(int P, int W, int H) {
  const int ib = tile_id_bits((uint32_t)(P > 1 ? P - 1 : 1));
  return ib + 32 + tile_id_bits((uint32_t)(tiles_x(W) * tiles_y(H))) <= 64;
}

// (u64 key, u32 value) generation for the pair mode, same order as duplicateWithKeys
__global__ void __launch_bounds__(256)
duplicate_pairs_kernel(int P, const float2* __restrict__ means2D, const float* __restrict__ depths,
                       const uint32_t* __restrict__ offsets, const int* __restrict__ radii,
                       const uint32_t* __restrict__ order, uint64_t* __restrict__ keys, uint32_t* __restrict__ vals,
                       int grid_x, int grid_y) {
  const int slot = blockIdx.x * blockDim.x + threadIdx.x;
  if (slot >= P) return;
  const int idx = order ? (int)order[slot] : slot;
  const int r = radii[idx];
  if (r <= 0) return;
  uint32_t off = (slot == 0) ? 0u : offsets[slot - 1];
  uint2 lo, hi;
  tile_rect(means2D[idx], r, grid_x, grid_y, lo, hi);
  const uint64_t depth_bits = (uint64_t)__float_as_uint(depths[idx]);
  for (uint32_t y = lo.y; y < hi.y; ++y)
    for (uint32_t x = lo.x; x < hi.x; ++x) {
      keys[off] = ((uint64_t)(y * (uint32_t)grid_x + x) << 32) | depth_bits;
      vals[off] = (uint32_t)idx;
      ++off;
    }
}

template <bool PAIRS>
static int run_onesweep_passes(const BinningPtrs& b, int R, int passes, int begin_bit, int end_bit, int ib,
                               int src, cudaStream_t stream, bool debug) {
  const int blocks = (int)b.sort_blocks;
  const size_t smem = PAIRS ? sizeof(PackedSmem) : offsetof(PackedSmem, exch_v);
  EGS_CUDA_TRY(cudaFuncSetAttribute(packed_onesweep_kernel<false, PAIRS>,
                                    cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  EGS_CUDA_TRY(cudaFuncSetAttribute(packed_onesweep_kernel<true, PAIRS>,
                                    cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  for (int p = 0; p < passes; ++p) {
    const int shift = begin_bit + p * PK_BITS;
    const int width = (end_bit - shift) < PK_BITS ? (end_bit - shift) : PK_BITS;
    const int dst = src ^ 1;
    uint32_t* status = b.status + (size_t)p * b.sort_blocks * PK_RADIX;
    if (p + 1 < passes) {
      packed_onesweep_kernel<false, PAIRS><<<blocks, PK_THREADS, smem, stream>>>(
          b.keys[src], b.vals[src], b.keys[dst], PAIRS ? b.vals[dst] : nullptr, R, shift, (1u << width) - 1u, ib,
          b.hist + p * PK_RADIX, b.ticket + p, status);
    } else {
      packed_onesweep_kernel<true, PAIRS><<<blocks, PK_THREADS, smem, stream>>>(
          b.keys[src], b.vals[src], b.keys[dst], b.vals[dst], R, shift, (1u << width) - 1u, ib,
          b.hist + p * PK_RADIX, b.ticket + p, status);
    }
    EGS_AFTER_LAUNCH(stream, debug);
    src = dst;
  }
  return EGS_OK;
}

// pairs = false: packed words (index inside the word); pairs = true: keys + a value array
int launch_binning_packed(const GeomPtrs& g, const BinningPtrs& b, const ImagePtrs& im, const int* radii,
                          const uint32_t* order, int P, int R, int W, int H, bool pairs, cudaStream_t stream,
                          bool debug) {
  const int gx = tiles_x(W), gy = tiles_y(H);
  EGS_CUDA_TRY(cudaMemsetAsync(im.ranges, 0, (size_t)gx * gy * sizeof(uint2), stream));
  if (R == 0) return EGS_OK;
  const int ib = pairs ? 0 : tile_id_bits((uint32_t)(P > 1 ? P - 1 : 1));
  const int key_bits = 32 + tile_id_bits((uint32_t)(gx * gy));
  const int begin_bit = ib, end_bit = ib + key_bits;
  const int passes = (key_bits + PK_BITS - 1) / PK_BITS;
  // ping-pong so that the last pass reads buffer 1 and leaves the sorted lists in buffer 0
  const int src = (passes & 1) ? 1 : 0;
  {
    StageTimer timer(stream, ST_DUPLICATE);
    if (pairs)
      duplicate_pairs_kernel<<<(P + 255) / 256, 256, 0, stream>>>(P, g.means2D, g.depths, g.point_offsets, radii,
                                                                  order, b.keys[src], b.vals[src], gx, gy);
    else
      duplicate_packed_kernel<<<(P + 255) / 256, 256, 0, stream>>>(P, g.means2D, g.depths, g.point_offsets,
                                                                   radii, order, b.keys[src], gx, gy, ib);
    EGS_AFTER_LAUNCH(stream, debug);
  }
  {
    StageTimer timer(stream, ST_SORT_HIST);
    EGS_CUDA_TRY(cudaMemsetAsync(b.hist, 0, b.zero_bytes, stream));
    switch (passes) {
      case 4: launch_packed_hist<4>(b.keys[src], R, begin_bit, end_bit, b.hist, stream); break;
      case 5: launch_packed_hist<5>(b.keys[src], R, begin_bit, end_bit, b.hist, stream); break;
      default: launch_packed_hist<6>(b.keys[src], R, begin_bit, end_bit, b.hist, stream); break;
    }
    EGS_AFTER_LAUNCH(stream, debug);
    packed_scan_hist_kernel<<<passes, PK_RADIX, 0, stream>>>(b.hist);
    EGS_AFTER_LAUNCH(stream, debug);
  }
  {
    StageTimer timer(stream, ST_SORT_PASSES);
    const int rc = pairs ? run_onesweep_passes<true>(b, R, passes, begin_bit, end_bit, ib, src, stream, debug)
                         : run_onesweep_passes<false>(b, R, passes, begin_bit, end_bit, ib, src, stream, debug);
    if (rc) return rc;
  }
  return launch_tile_ranges(R, b.keys[0], im, stream, debug);
}

}  // namespace egs
