// binning.cu -- tiles_touched prefix sum, (tile | depth) key duplication, in-house LSD radix
// sort of (u64 key, u32 value) pairs, and per-tile range identification. No CUB.
//
// Replaces, with identical results:
//   cub::DeviceScan::InclusiveSum            rasterizer_impl.cu:278
//   duplicateWithKeys                        rasterizer_impl.cu:70-111
//   cub::DeviceRadixSort::SortPairs          rasterizer_impl.cu:304-309  (bits [0, 32+bit))
//   cudaMemset + identifyTileRanges          rasterizer_impl.cu:311-319,116-138
//
// The sort is a stable least-significant-digit radix sort, 8-bit digits, one "onesweep" pass per
// digit: every block ranks its 4096 items per digit, publishes its per-digit count and resolves
// its global offsets with a decoupled look-back over the preceding blocks' status words. Blocks
// take their logical tile number from an atomic ticket so that a block never waits on one that
// has not started. Stability (ties keep Gaussian-index order) is what makes the result equal to
// the reference's stable sort.
#include "common.cuh"

namespace egs {

// ---------------------------------------------------------------------------------------------
// inclusive scan of tiles_touched -> point_offsets (single pass, decoupled look-back)
// ---------------------------------------------------------------------------------------------
constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = SCAN_TILE / SCAN_THREADS;  // 16
constexpr unsigned long long SCAN_FLAG_AGG = 1ull << 62;
constexpr unsigned long long SCAN_FLAG_PREFIX = 1ull << 63;
constexpr unsigned long long SCAN_VALUE_MASK = (1ull << 62) - 1;

// GATHER: element i of the scanned sequence is in[perm[i]]
template <bool GATHER>
__global__ void __launch_bounds__(SCAN_THREADS)
scan_tiles_kernel(const uint32_t* __restrict__ in, const uint32_t* __restrict__ perm, uint32_t* __restrict__ out,
                  int n, unsigned long long* __restrict__ status, uint32_t* __restrict__ total /* may be null */) {
  __shared__ uint32_t s_tile;
  __shared__ uint32_t s_warp_sums[SCAN_THREADS / 32];
  __shared__ uint32_t s_block_prefix;
  if (threadIdx.x == 0) s_tile = (uint32_t)atomicAdd(&status[0], 1ull);
  __syncthreads();
  const uint32_t tile = s_tile;
  volatile unsigned long long* st = status + 1;

  const size_t base = (size_t)tile * SCAN_TILE + (size_t)threadIdx.x * SCAN_ITEMS;
  uint32_t v[SCAN_ITEMS];
  if (GATHER) {
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) v[k] = (base + k < (size_t)n) ? perm[base + k] : 0u;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) v[k] = (base + k < (size_t)n) ? in[v[k]] : 0u;
  } else if (base + SCAN_ITEMS <= (size_t)n) {
    const uint4* p = reinterpret_cast<const uint4*>(in + base);
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS / 4; ++k) {
      uint4 q = p[k];
      v[4 * k] = q.x; v[4 * k + 1] = q.y; v[4 * k + 2] = q.z; v[4 * k + 3] = q.w;
    }
  } else {
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) v[k] = (base + k < (size_t)n) ? in[base + k] : 0u;
  }
  uint32_t run = 0;
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; ++k) { run += v[k]; v[k] = run; }

  // block-wide exclusive scan of the per-thread totals
  const unsigned lane = lane_id(), warp = threadIdx.x >> 5;
  uint32_t incl = run;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    uint32_t t = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= d) incl += t;
  }
  if (lane == 31) s_warp_sums[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    uint32_t w = lane < SCAN_THREADS / 32 ? s_warp_sums[lane] : 0;
    uint32_t wi = w;
#pragma unroll
    for (int d = 1; d < SCAN_THREADS / 32; d <<= 1) {
      uint32_t t = __shfl_up_sync(0xffffffffu, wi, d);
      if (lane >= d) wi += t;
    }
    if (lane < SCAN_THREADS / 32) s_warp_sums[lane] = wi - w;  // exclusive warp offsets
    const uint32_t block_total = __shfl_sync(0xffffffffu, wi, SCAN_THREADS / 32 - 1);
    // decoupled look-back, 32 predecessors per step: lane l reads tile (t0 - l); the sum runs up to
    // the nearest tile that already holds an inclusive prefix (one lane walking back tile by tile
    // paid an L2 round trip per predecessor: 14 us for 245 tiles)
    uint32_t excl = 0;
    if (tile == 0) {
      if (lane == 0) st[0] = SCAN_FLAG_PREFIX | block_total;
    } else {
      if (lane == 0) st[tile] = SCAN_FLAG_AGG | block_total;
      for (int t0 = (int)tile - 1;; t0 -= 32) {
        const int t = t0 - (int)lane;
        unsigned long long s = SCAN_FLAG_PREFIX;   // before tile 0: an empty prefix
        if (t >= 0) {
          do { s = st[t]; } while ((s & (SCAN_FLAG_AGG | SCAN_FLAG_PREFIX)) == 0);
        }
        const unsigned prefixes = __ballot_sync(0xffffffffu, (s & SCAN_FLAG_PREFIX) != 0);
        const int first = prefixes ? __ffs((int)prefixes) - 1 : 32;
        uint32_t val = (int)lane <= first ? (uint32_t)(s & SCAN_VALUE_MASK) : 0u;
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) val += __shfl_xor_sync(0xffffffffu, val, d);
        excl += val;
        if (prefixes) break;
      }
      if (lane == 0) st[tile] = SCAN_FLAG_PREFIX | (unsigned long long)(excl + block_total);
    }
    if (lane == 0) s_block_prefix = excl;
  }
  __syncthreads();
  const uint32_t off = s_block_prefix + s_warp_sums[warp] + (incl - run);
  if (base + SCAN_ITEMS <= (size_t)n) {
    uint4* p = reinterpret_cast<uint4*>(out + base);
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS / 4; ++k)
      p[k] = make_uint4(v[4 * k] + off, v[4 * k + 1] + off, v[4 * k + 2] + off, v[4 * k + 3] + off);
  } else {
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k)
      if (base + k < (size_t)n) out[base + k] = v[k] + off;
  }
  // the sum of everything, next to the other words the host reads back
  if (total != nullptr) {
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k)
      if (base + k + 1 == (size_t)n) *total = v[k] + off;
  }
}

int launch_scan_tiles(const GeomPtrs& g, int P, const uint32_t* order, cudaStream_t stream, bool debug) {
  const int blocks = (P + SCAN_TILE - 1) / SCAN_TILE;
  StageTimer timer(stream, ST_SCAN);
  // g.scan_status was cleared with the rest of the zeroed run when the forward began
  if (order)
    scan_tiles_kernel<true><<<blocks, SCAN_THREADS, 0, stream>>>(g.tiles_touched, order, g.point_offsets, P,
                                                                 g.scan_status, g.readback);
  else
    scan_tiles_kernel<false><<<blocks, SCAN_THREADS, 0, stream>>>(g.tiles_touched, nullptr, g.point_offsets, P,
                                                                  g.scan_status, g.readback);
  EGS_AFTER_LAUNCH(stream, debug);
  return EGS_OK;
}

// the same scan for any u32 array (the cell counts of knn.cu); `status` holds n / SCAN_TILE + 2 words
int launch_inclusive_scan(const uint32_t* in, uint32_t* out, int n, unsigned long long* status, cudaStream_t stream) {
  const int blocks = (n + SCAN_TILE - 1) / SCAN_TILE;
  EGS_CUDA_TRY(cudaMemsetAsync(status, 0, (size_t)(blocks + 1) * 8, stream));
  scan_tiles_kernel<false><<<blocks, SCAN_THREADS, 0, stream>>>(in, nullptr, out, n, status, nullptr);
  EGS_AFTER_LAUNCH(stream, false);
  return EGS_OK;
}

// ---------------------------------------------------------------------------------------------
// key duplication: one (tile << 32 | depth bits, gaussian) pair per covered tile, rows first
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
duplicate_keys_kernel(int P, const float2* __restrict__ means2D, const float* __restrict__ depths,
                      const uint32_t* __restrict__ offsets, const int* __restrict__ radii,
                      uint64_t* __restrict__ keys, uint32_t* __restrict__ vals, int grid_x,
                      int grid_y) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= P) return;
  const int r = radii[idx];
  if (r <= 0) return;
  uint32_t off = (idx == 0) ? 0u : offsets[idx - 1];
  uint2 lo, hi;
  tile_rect(means2D[idx], r, grid_x, grid_y, lo, hi);
  const uint64_t depth_bits = (uint64_t)__float_as_uint(depths[idx]);
  for (uint32_t y = lo.y; y < hi.y; ++y) {
    for (uint32_t x = lo.x; x < hi.x; ++x) {
      const uint64_t tile = (uint64_t)(y * (uint32_t)grid_x + x);
      keys[off] = (tile << 32) | depth_bits;
      vals[off] = (uint32_t)idx;
      ++off;
    }
  }
}

// ---------------------------------------------------------------------------------------------
// radix sort
// ---------------------------------------------------------------------------------------------
// Digit histograms of every pass in one sweep over the keys.
template <int PASSES>
__global__ void __launch_bounds__(256)
sort_histogram_kernel(const uint64_t* __restrict__ keys, int n, int end_bit,
                      uint32_t* __restrict__ hist) {
  __shared__ uint32_t s_hist[PASSES * SORT_RADIX];
  for (int i = threadIdx.x; i < PASSES * SORT_RADIX; i += blockDim.x) s_hist[i] = 0;
  __syncthreads();
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < (size_t)n;
       i += (size_t)gridDim.x * blockDim.x) {
    const uint64_t k = keys[i];
#pragma unroll
    for (int p = 0; p < PASSES; ++p) {
      const int shift = p * SORT_RADIX_BITS;
      const int width = min(SORT_RADIX_BITS, end_bit - shift);
      const uint32_t d = (uint32_t)(k >> shift) & ((1u << width) - 1u);
      atomicAdd(&s_hist[p * SORT_RADIX + d], 1u);
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < PASSES * SORT_RADIX; i += blockDim.x) {
    const uint32_t c = s_hist[i];
    if (c) atomicAdd(&hist[i], c);
  }
}

// In-place exclusive scan of each pass's 256-bin histogram (one block per pass).
__global__ void __launch_bounds__(SORT_RADIX)
sort_scan_hist_kernel(uint32_t* __restrict__ hist) {
  __shared__ uint32_t s_w[SORT_RADIX / 32];
  uint32_t* h = hist + blockIdx.x * SORT_RADIX;
  const uint32_t v = h[threadIdx.x];
  const unsigned lane = lane_id(), warp = threadIdx.x >> 5;
  uint32_t incl = v;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    uint32_t t = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= d) incl += t;
  }
  if (lane == 31) s_w[warp] = incl;
  __syncthreads();
  uint32_t woff = 0;
  for (unsigned w = 0; w < warp; ++w) woff += s_w[w];
  h[threadIdx.x] = woff + incl - v;
}

constexpr uint32_t SORT_FLAG_AGG = 1u << 30;
constexpr uint32_t SORT_FLAG_PREFIX = 1u << 31;
constexpr uint32_t SORT_VALUE_MASK = (1u << 30) - 1u;

__global__ void __launch_bounds__(SORT_THREADS)
sort_onesweep_kernel(const uint64_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                     uint64_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out, int n,
                     int shift, uint32_t digit_mask, const uint32_t* __restrict__ digit_base,
                     uint32_t* __restrict__ ticket, uint32_t* __restrict__ status_raw) {
  constexpr int WARPS = SORT_THREADS / 32;
  __shared__ uint32_t s_warp_hist[WARPS][SORT_RADIX];
  __shared__ uint32_t s_digit_off[SORT_RADIX];
  __shared__ uint32_t s_tile;
  if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1u);
#pragma unroll
  for (int w = 0; w < WARPS; ++w) s_warp_hist[w][threadIdx.x] = 0;
  __syncthreads();
  const uint32_t tile = s_tile;
  volatile uint32_t* status = status_raw;
  const unsigned lane = lane_id(), warp = threadIdx.x >> 5;
  const unsigned lt_mask = (1u << lane) - 1u;

  // warp-striped arrangement inside a warp-contiguous segment keeps index order = (warp, i, lane)
  const size_t warp_base = (size_t)tile * SORT_TILE + (size_t)warp * (32 * SORT_ITEMS);
  uint64_t key[SORT_ITEMS];
  uint32_t val[SORT_ITEMS];
  uint32_t rank[SORT_ITEMS];
#pragma unroll
  for (int i = 0; i < SORT_ITEMS; ++i) {
    const size_t g = warp_base + (size_t)i * 32 + lane;
    const bool ok = g < (size_t)n;
    key[i] = ok ? keys_in[g] : ~0ull;
    val[i] = ok ? vals_in[g] : 0u;
  }
#pragma unroll
  for (int i = 0; i < SORT_ITEMS; ++i) {
    const size_t g = warp_base + (size_t)i * 32 + lane;
    const bool ok = g < (size_t)n;
    const uint32_t d = (uint32_t)(key[i] >> shift) & digit_mask;
    // out-of-range lanes get a private pseudo digit so they never join a peer group
    const uint32_t peers = match_digit<SORT_RADIX_BITS>(d, ok);
    const uint32_t before = ok ? s_warp_hist[warp][d] : 0u;
    __syncwarp();
    if (ok && (peers & lt_mask) == 0u) s_warp_hist[warp][d] = before + __popc(peers);
    __syncwarp();
    rank[i] = before + __popc(peers & lt_mask);
  }
  __syncthreads();

  // thread d owns digit d: exclusive scan over the warps, then the cross-block look-back
  {
    const uint32_t d = threadIdx.x;
    uint32_t run = 0;
#pragma unroll
    for (int w = 0; w < WARPS; ++w) {
      const uint32_t c = s_warp_hist[w][d];
      s_warp_hist[w][d] = run;
      run += c;
    }
    const uint32_t block_count = run;
    uint32_t excl = 0;
    volatile uint32_t* st = status + d;
    if (tile == 0) {
      st[0] = SORT_FLAG_PREFIX | block_count;
    } else {
      st[(size_t)tile * SORT_RADIX] = SORT_FLAG_AGG | block_count;
      for (int t = (int)tile - 1;; --t) {
        uint32_t s;
        do { s = st[(size_t)t * SORT_RADIX]; } while ((s & (SORT_FLAG_AGG | SORT_FLAG_PREFIX)) == 0u);
        excl += s & SORT_VALUE_MASK;
        if (s & SORT_FLAG_PREFIX) break;
      }
      st[(size_t)tile * SORT_RADIX] = SORT_FLAG_PREFIX | (excl + block_count);
    }
    s_digit_off[d] = digit_base[d] + excl;
  }
  __syncthreads();

#pragma unroll
  for (int i = 0; i < SORT_ITEMS; ++i) {
    const size_t g = warp_base + (size_t)i * 32 + lane;
    if (g < (size_t)n) {
      const uint32_t d = (uint32_t)(key[i] >> shift) & digit_mask;
      const uint32_t dst = s_digit_off[d] + s_warp_hist[warp][d] + rank[i];
      keys_out[dst] = key[i];
      vals_out[dst] = val[i];
    }
  }
}

// ---------------------------------------------------------------------------------------------
// per-tile [start, end) ranges in the sorted list
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
tile_ranges_kernel(int n, const uint64_t* __restrict__ keys, uint2* __restrict__ ranges) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n) return;
  const uint32_t cur = (uint32_t)(keys[idx] >> 32);
  if (idx == 0) {
    ranges[cur].x = 0;
  } else {
    const uint32_t prev = (uint32_t)(keys[idx - 1] >> 32);
    if (cur != prev) {
      ranges[prev].y = idx;
      ranges[cur].x = idx;
    }
  }
  if (idx == n - 1) ranges[cur].y = n;
}

int launch_tile_ranges(int R, const uint64_t* sorted_keys, const ImagePtrs& im, cudaStream_t stream,
                       bool debug) {
  StageTimer timer(stream, ST_RANGES);
  tile_ranges_kernel<<<(R + 255) / 256, 256, 0, stream>>>(R, sorted_keys, im.ranges);
  EGS_AFTER_LAUNCH(stream, debug);
  return EGS_OK;
}

template <int PASSES>
static void launch_hist(const uint64_t* keys, int n, int end_bit, uint32_t* hist,
                        cudaStream_t stream) {
  int blocks = (n + 256 * 16 - 1) / (256 * 16);
  blocks = blocks < 1 ? 1 : (blocks > 148 * 8 ? 148 * 8 : blocks);
  sort_histogram_kernel<PASSES><<<blocks, 256, 0, stream>>>(keys, n, end_bit, hist);
}

int launch_binning(const GeomPtrs& g, const BinningPtrs& b, const ImagePtrs& im, const int* radii,
                   int P, int R, int W, int H, cudaStream_t stream, bool debug) {
  const int gx = tiles_x(W), gy = tiles_y(H);
  EGS_CUDA_TRY(cudaMemsetAsync(im.ranges, 0, (size_t)gx * gy * sizeof(uint2), stream));
  if (R == 0) return EGS_OK;

  const int end_bit = 32 + tile_id_bits((uint32_t)(gx * gy));
  const int passes = (end_bit + SORT_RADIX_BITS - 1) / SORT_RADIX_BITS;
  // ping-pong so that the sorted lists always end in keys[0] / vals[0]
  const int first = passes & 1;
  {
    StageTimer timer(stream, ST_DUPLICATE);
    duplicate_keys_kernel<<<(P + 255) / 256, 256, 0, stream>>>(P, g.means2D, g.depths,
                                                               g.point_offsets, radii, b.keys[first],
                                                               b.vals[first], gx, gy);
    EGS_AFTER_LAUNCH(stream, debug);
  }

  {
    StageTimer timer(stream, ST_SORT_HIST);
    EGS_CUDA_TRY(cudaMemsetAsync(b.hist, 0, b.zero_bytes, stream));
    switch (passes) {
      case 4: launch_hist<4>(b.keys[first], R, end_bit, b.hist, stream); break;
      case 5: launch_hist<5>(b.keys[first], R, end_bit, b.hist, stream); break;
      case 6: launch_hist<6>(b.keys[first], R, end_bit, b.hist, stream); break;
      case 7: launch_hist<7>(b.keys[first], R, end_bit, b.hist, stream); break;
      default: launch_hist<8>(b.keys[first], R, end_bit, b.hist, stream); break;
    }
    EGS_AFTER_LAUNCH(stream, debug);
    sort_scan_hist_kernel<<<passes, SORT_RADIX, 0, stream>>>(b.hist);
    EGS_AFTER_LAUNCH(stream, debug);
  }
  {
    StageTimer timer(stream, ST_SORT_PASSES);
    const int blocks = (int)b.sort_blocks;
    for (int p = 0; p < passes; ++p) {
      const int shift = p * SORT_RADIX_BITS;
      const int width = (end_bit - shift) < SORT_RADIX_BITS ? (end_bit - shift) : SORT_RADIX_BITS;
      const int src = (p + first) & 1, dst = src ^ 1;
      sort_onesweep_kernel<<<blocks, SORT_THREADS, 0, stream>>>(
          b.keys[src], b.vals[src], b.keys[dst], b.vals[dst], R, shift, (1u << width) - 1u,
          b.hist + p * SORT_RADIX, b.ticket + p, b.status + (size_t)p * b.sort_blocks * SORT_RADIX);
      EGS_AFTER_LAUNCH(stream, debug);
    }
  }
  return launch_tile_ranges(R, b.keys[0], im, stream, debug);
}

}  // namespace egs
