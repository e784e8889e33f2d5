// binning_tiles.cu -- tile-bucket binning: the B200-first replacement of "duplicate keys, sort all
// R pairs on 32+bit key bits, then find the tile boundaries"
// (DGR/cuda_rasterizer/rasterizer_impl.cu:290-319).
//
// The reference's global sort orders pairs by (tile, depth bits) and, being stable over a list
// generated in Gaussian-index order, breaks ties by Gaussian index. The result is therefore the
// unique ordering by the triple (tile, depth bits, index) -- any algorithm that produces that
// ordering is bit-identical, stable or not. We exploit the structure of the key instead of
// radix-sorting 44 bits six times through HBM (152 B per pair):
//   1. preprocess already counted the instances of every tile (reductions into tile_counts);
//   2. one small block turns the counts into the tile ranges (an exclusive scan) -- the ranges
//      are known BEFORE anything is sorted, and identifyTileRanges disappears;
//   3. a scatter pass drops each instance's (depth bits << 32 | index) word into its tile's
//      segment at an atomically claimed slot (order inside the segment is arbitrary);
//   4. one CTA per tile sorts its segment in shared memory (bitonic network on 64-bit words) and
//      writes the sorted index list (and the 64-bit keys, for parity checks).
// HBM traffic: 8 B write + 8 B read + 4 B (+8 B keys) write per pair instead of 152 B.
// Tiles longer than TILE_SORT_MAX fall back to the global radix sort in binning.cu.
#include "common.cuh"

namespace egs {

// ---- 2. counts -> ranges ------------------------------------------------------------------
constexpr int TSCAN_THREADS = 1024;

__global__ void __launch_bounds__(TSCAN_THREADS)
tile_scan_kernel(const uint32_t* __restrict__ counts, int n_tiles, uint2* __restrict__ ranges,
                 uint32_t* __restrict__ meta) {
  __shared__ uint32_t s_warp[TSCAN_THREADS / 32];
  __shared__ uint32_t s_max[TSCAN_THREADS / 32];
  const int per = (n_tiles + TSCAN_THREADS - 1) / TSCAN_THREADS;
  const int begin = threadIdx.x * per;
  const int end = min(begin + per, n_tiles);
  uint32_t sum = 0, mx = 0;
  for (int i = begin; i < end; ++i) {
    const uint32_t c = counts[i];
    sum += c;
    mx = max(mx, c);
  }
  const unsigned lane = lane_id(), warp = threadIdx.x >> 5;
  uint32_t incl = sum;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const uint32_t t = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= d) incl += t;
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, d));
  if (lane == 31) s_warp[warp] = incl;
  if (lane == 0) s_max[warp] = mx;
  __syncthreads();
  uint32_t woff = 0, total = 0, gmax = 0;
  for (unsigned w = 0; w < TSCAN_THREADS / 32; ++w) {
    const uint32_t v = s_warp[w];
    if (w < warp) woff += v;
    total += v;
    gmax = max(gmax, s_max[w]);
  }
  uint32_t run = woff + incl - sum;
  for (int i = begin; i < end; ++i) {
    const uint32_t c = counts[i];
    // empty tiles keep (0, 0) like the reference's zero-filled ranges (rasterizer_impl.cu:311)
    ranges[i] = c ? make_uint2(run, run + c) : make_uint2(0u, 0u);
    run += c;
  }
  if (threadIdx.x == 0) {
    meta[0] = total;
    meta[1] = gmax;
  }
}

int launch_tile_scan(const ImagePtrs& im, int W, int H, cudaStream_t stream, bool debug) {
  StageTimer timer(stream, ST_TILE_SCAN);
  tile_scan_kernel<<<1, TSCAN_THREADS, 0, stream>>>(im.tile_counts, tiles_x(W) * tiles_y(H), im.ranges,
                                                    im.bin_meta);
  EGS_AFTER_LAUNCH(stream, debug);
  return EGS_OK;
}

// ---- 3. scatter ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
tile_scatter_kernel(int P, const float2* __restrict__ means2D, const float* __restrict__ depths,
                    const int* __restrict__ radii, const uint2* __restrict__ ranges,
                    uint32_t* __restrict__ cursor, uint64_t* __restrict__ pairs, int grid_x, int grid_y) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= P) return;
  const int r = radii[idx];
  if (r <= 0) return;
  uint2 lo, hi;
  tile_rect(means2D[idx], r, grid_x, grid_y, lo, hi);
  const uint64_t word = ((uint64_t)__float_as_uint(depths[idx]) << 32) | (uint32_t)idx;
  for (uint32_t y = lo.y; y < hi.y; ++y) {
    for (uint32_t x = lo.x; x < hi.x; ++x) {
      const uint32_t t = y * (uint32_t)grid_x + x;
      const uint32_t slot = atomicAdd(&cursor[t], 1u);
      pairs[ranges[t].x + slot] = word;
    }
  }
}

// ---- 4. per-tile sort ---------------------------------------------------------------------
constexpr int TSORT_THREADS = 512;

__global__ void __launch_bounds__(TSORT_THREADS)
tile_sort_kernel(const uint2* __restrict__ ranges, const uint64_t* __restrict__ pairs,
                 uint32_t* __restrict__ point_list, uint64_t* __restrict__ keys_sorted) {
  extern __shared__ uint64_t s_w[];
  const uint32_t tile = blockIdx.x;
  const uint2 range = ranges[tile];
  const uint32_t n = range.y - range.x;
  if (n == 0) return;
  uint32_t np2 = 2;
  while (np2 < n) np2 <<= 1;
  const uint32_t tid = threadIdx.x;
  for (uint32_t i = tid; i < np2; i += TSORT_THREADS) s_w[i] = i < n ? pairs[range.x + i] : ~0ull;
  __syncthreads();
  const uint32_t half = np2 >> 1;
  for (uint32_t k = 2; k <= np2; k <<= 1) {
    for (uint32_t j = k >> 1; j > 0; j >>= 1) {
      for (uint32_t t = tid; t < half; t += TSORT_THREADS) {
        const uint32_t i = ((t & ~(j - 1)) << 1) | (t & (j - 1));   // index with bit j clear
        uint64_t a = s_w[i], b = s_w[i | j];
        const bool asc = (i & k) == 0;
        if ((a > b) == asc) {
          s_w[i] = b;
          s_w[i | j] = a;
        }
      }
      __syncthreads();
    }
  }
  const uint64_t tile_hi = (uint64_t)tile << 32;
  for (uint32_t i = tid; i < n; i += TSORT_THREADS) {
    const uint64_t w = s_w[i];
    point_list[range.x + i] = (uint32_t)w;
    keys_sorted[range.x + i] = tile_hi | (w >> 32);
  }
}

int launch_tile_binning(const GeomPtrs& g, const BinningPtrs& b, const ImagePtrs& im, const int* radii,
                        int P, int R, int max_tile, int W, int H, cudaStream_t stream, bool debug) {
  if (R == 0) return EGS_OK;
  const int gx = tiles_x(W), gy = tiles_y(H);
  // sorted data must land where the radix path leaves it (see abi.cu): buffer index passes & 1
  const int sel = sort_passes(W, H) & 1;
  uint64_t* pairs = b.keys[sel ^ 1];
  {
    StageTimer timer(stream, ST_TILE_SCATTER);
    tile_scatter_kernel<<<(P + 255) / 256, 256, 0, stream>>>(P, g.means2D, g.depths, radii, im.ranges,
                                                             im.tile_cursor, pairs, gx, gy);
    EGS_AFTER_LAUNCH(stream, debug);
  }
  {
    StageTimer timer(stream, ST_TILE_SORT);
    uint32_t cap = 2;
    while (cap < (uint32_t)max_tile) cap <<= 1;
    const size_t smem = (size_t)cap * sizeof(uint64_t);
    if (smem > 48 * 1024)   // opt in to large dynamic shared memory (per device, cheap to repeat)
      EGS_CUDA_TRY(cudaFuncSetAttribute(tile_sort_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        (int)(TILE_SORT_MAX * sizeof(uint64_t))));
    tile_sort_kernel<<<gx * gy, TSORT_THREADS, smem, stream>>>(im.ranges, pairs, b.vals[sel], b.keys[sel]);
    EGS_AFTER_LAUNCH(stream, debug);
  }
  return EGS_OK;
}

}  // namespace egs
