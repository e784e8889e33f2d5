// knn.cu -- mean squared distance to the three nearest neighbours of every point (init-time scale
// of the Gaussians: scene/gaussian_model.py:301,312 `distCUDA2`).
//
// Same result as the reference's second native submodule (submodules/simple-knn/simple_knn.cu:
// Morton sort + 1024-point boxes pruned by their distance to the query, then the mean of the three
// smallest squared distances, self excluded, duplicates included). The pruning there is exact, so
// the result is simply the exact 3-NN statistic; two ways to get it:
//   * below KNN_GRID_MIN points: a tiled all-pairs sweep (shared-memory tiles of 1024 points,
//     13 instructions per pair), no scratch at all;
//   * above: a grid of cubic cells over the bulk of the cloud (mean +- 3 sigma per axis, ~4 points
//     per cell, points beyond it clamped into the border cells, which therefore reach to infinity), built
//     by a counting sort (count, scan, scatter) entirely on the device, and searched ring by ring
//     until the third-best distance cannot be beaten by anything outside the searched cube.
// Neither needs a host round trip. The squared distance uses the reference's expression, and the
// three smallest values do not depend on the visiting order, so both paths reproduce the reference
// bit for bit (tests/test_knn_gpu.py).
#include "common.cuh"

#include <float.h>
#include <math.h>

namespace egs {

constexpr int KNN_THREADS = 256;
constexpr int KNN_TILE = 1024;
constexpr int KNN_GRID_MIN = 32768;   // below this the all-pairs sweep (0.6 ms here) beats building a grid
constexpr int KNN_POINTS_PER_CELL = 4;  // grid resolution: cells = P / 4 at most

__device__ __forceinline__ void knn_insert(float dist, float& b0, float& b1, float& b2) {
  const float m0 = fminf(b0, dist), r0 = fmaxf(b0, dist);
  const float m1 = fminf(b1, r0), r1 = fmaxf(b1, r0);
  b0 = m0;
  b1 = m1;
  b2 = fminf(b2, r1);
}

// the reference's expression (simple_knn.cu:133-134), so the squared distances round alike
__device__ __forceinline__ float knn_dist2(const float3 ref, float px, float py, float pz) {
  const float3 d = {px - ref.x, py - ref.y, pz - ref.z};
  return d.x * d.x + d.y * d.y + d.z * d.z;
}

__global__ void __launch_bounds__(KNN_THREADS)
knn_all_pairs_kernel(int P, const float* __restrict__ points, float* __restrict__ mean_dist2) {
  __shared__ float4 s_pts[KNN_TILE];
  const int i = blockIdx.x * KNN_THREADS + threadIdx.x;
  float3 ref = make_float3(0.f, 0.f, 0.f);
  if (i < P) ref = make_float3(points[3 * i], points[3 * i + 1], points[3 * i + 2]);
  float b0 = FLT_MAX, b1 = FLT_MAX, b2 = FLT_MAX;   // ascending
  for (int base = 0; base < P; base += KNN_TILE) {
    __syncthreads();
    for (int t = threadIdx.x; t < KNN_TILE; t += KNN_THREADS) {
      const int j = base + t;
      s_pts[t] = j < P ? make_float4(points[3 * j], points[3 * j + 1], points[3 * j + 2], 0.f)
                       : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    __syncthreads();
    const int n = min(KNN_TILE, P - base);
    const int self = i - base;   // position of this thread's own point in the tile, if any
#pragma unroll 8
    for (int t = 0; t < n; ++t) {
      const float4 p = s_pts[t];
      float dist = knn_dist2(ref, p.x, p.y, p.z);
      if (t == self) dist = FLT_MAX;
      knn_insert(dist, b0, b1, b2);
    }
  }
  if (i < P) mean_dist2[i] = (b0 + b1 + b2) / 3.0f;
}

// ---- grid path -----------------------------------------------------------------------------------
// Workspace (all on the device):
//   double sums[6]              sum and sum of squares per axis
//   KnnGrid grid                origin, cell size, 1 / cell size (written by knn_grid_setup_kernel)
//   uint32 cell_count[G^3]      points per cell; counted down again by the scatter
//   uint32 cell_end[G^3]        inclusive scan of the counts
//   uint64 scan_status[...]     look-back words of the tiles scan
//   float4 sorted[P]            points in cell order, w = original index
struct KnnGrid {
  float3 origin;
  float cell, inv_cell;   // cubic cells: the reach of r rings is r * cell along every axis
  int3 dims;              // cells per axis, chosen on the device from the spread of the cloud
};

struct KnnWorkspace {
  double* sums;
  KnnGrid* grid;
  uint32_t* cell_count;
  uint32_t* cell_end;
  unsigned long long* scan_status;
  float4* sorted;
  size_t bytes;
  KnnWorkspace(void* base, int P, int cells) {
    size_t o = 0;
    auto take = [&](size_t n) { const size_t r = o; o = align_up(o + n); return r; };
    char* b = static_cast<char*>(base);
    sums = reinterpret_cast<double*>(b + take(6 * sizeof(double)));
    grid = reinterpret_cast<KnnGrid*>(b + take(sizeof(KnnGrid)));
    cell_count = reinterpret_cast<uint32_t*>(b + take((size_t)cells * 4));
    cell_end = reinterpret_cast<uint32_t*>(b + take((size_t)cells * 4));
    scan_status = reinterpret_cast<unsigned long long*>(b + take(((size_t)cells / SCAN_TILE + 2) * 8));
    sorted = reinterpret_cast<float4*>(b + take((size_t)P * 16));
    bytes = o + 256;
  }
};

// upper bound of the cell count (the device picks dims whose product stays below it)
static int knn_max_cells(int P) { return P / KNN_POINTS_PER_CELL + 64; }

__global__ void __launch_bounds__(256)
knn_moments_kernel(int P, const float* __restrict__ points, double* __restrict__ sums) {
  double s[6] = {0, 0, 0, 0, 0, 0};
  for (int i = blockIdx.x * 256 + threadIdx.x; i < P; i += gridDim.x * 256) {
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      const double v = points[3 * i + a];
      s[a] += v;
      s[3 + a] += v * v;
    }
  }
  __shared__ double s_part[8][6];
#pragma unroll
  for (int k = 0; k < 6; ++k) {
    double v = s[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5][k] = v;
  }
  __syncthreads();
  if (threadIdx.x < 6) {
    double v = 0;
    for (int w = 0; w < 8; ++w) v += s_part[w][threadIdx.x];
    atomicAdd(&sums[threadIdx.x], v);
  }
}

__global__ void knn_grid_setup_kernel(int P, int max_cells, const double* __restrict__ sums,
                                      KnnGrid* __restrict__ grid) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  double lo[3], ext[3], widest = 0.0;
  for (int a = 0; a < 3; ++a) {
    const double mean = sums[a] / P;
    const double var = fmax(sums[3 + a] / P - mean * mean, 0.0);
    lo[a] = mean - 3.0 * sqrt(var);
    ext[a] = 6.0 * sqrt(var);
    widest = fmax(widest, ext[a]);
  }
  widest = fmax(widest, 1e-30);
  // smallest cube edge whose grid fits in max_cells: start from the volume estimate (too fine when an
  // axis is flat, because such an axis still takes one cell) and coarsen until it fits
  double vol = 1.0;
  for (int a = 0; a < 3; ++a) vol *= fmax(ext[a], 1e-6 * widest);
  double h = cbrt(vol / max_cells);
  int dims[3];
  for (int it = 0; it < 4096; ++it) {
    double cells = 1.0;
    for (int a = 0; a < 3; ++a) {
      const double n = fmin(fmax(ceil(ext[a] / h), 1.0), 1e6);
      dims[a] = (int)n;
      cells *= n;
    }
    if (cells <= (double)max_cells) break;
    h *= 1.02;
  }
  KnnGrid g;
  g.origin = make_float3((float)lo[0], (float)lo[1], (float)lo[2]);
  g.cell = (float)h;
  g.inv_cell = 1.f / g.cell;
  g.dims = make_int3(dims[0], dims[1], dims[2]);
  *grid = g;
}

__device__ __forceinline__ int3 knn_cell_of(const KnnGrid& g, float x, float y, float z) {
  // points beyond the gridded bulk fall into the border cells
  const int cx = min(g.dims.x - 1, max(0, (int)floorf((x - g.origin.x) * g.inv_cell)));
  const int cy = min(g.dims.y - 1, max(0, (int)floorf((y - g.origin.y) * g.inv_cell)));
  const int cz = min(g.dims.z - 1, max(0, (int)floorf((z - g.origin.z) * g.inv_cell)));
  return make_int3(cx, cy, cz);
}

__global__ void __launch_bounds__(256)
knn_count_kernel(int P, const float* __restrict__ points, const KnnGrid* __restrict__ grid,
                 uint32_t* __restrict__ cell_count) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= P) return;
  const KnnGrid g = *grid;
  const int3 c = knn_cell_of(g, points[3 * i], points[3 * i + 1], points[3 * i + 2]);
  atomicAdd(&cell_count[(c.z * g.dims.y + c.y) * g.dims.x + c.x], 1u);
}

__global__ void __launch_bounds__(256)
knn_scatter_kernel(int P, const float* __restrict__ points, const KnnGrid* __restrict__ grid,
                   const uint32_t* __restrict__ cell_end, uint32_t* __restrict__ cell_count,
                   float4* __restrict__ sorted) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= P) return;
  const KnnGrid g = *grid;
  const float x = points[3 * i], y = points[3 * i + 1], z = points[3 * i + 2];
  const int3 c = knn_cell_of(g, x, y, z);
  const int cell = (c.z * g.dims.y + c.y) * g.dims.x + c.x;
  const uint32_t slot = cell_end[cell] - atomicSub(&cell_count[cell], 1u);   // any order within a cell
  sorted[slot] = make_float4(x, y, z, __int_as_float(i));
}

// One thread per point. Ring r = the cells at Chebyshev distance r from the point's own cell.
// Along every axis the searched slab reaches at least r cells beyond the point (a border cell
// reaches to infinity), so everything not yet visited is farther than r * cell.
__global__ void __launch_bounds__(128)
knn_grid_query_kernel(int P, const float4* __restrict__ sorted, const uint32_t* __restrict__ cell_end,
                      const KnnGrid* __restrict__ grid, float* __restrict__ mean_dist2) {
  const int s = blockIdx.x * 128 + threadIdx.x;   // position in cell order: neighbouring threads share cells
  if (s >= P) return;
  const KnnGrid g = *grid;
  const float4 me = sorted[s];
  const float3 ref = make_float3(me.x, me.y, me.z);
  const int3 c = knn_cell_of(g, me.x, me.y, me.z);
  const int3 G = g.dims;
  float b0 = FLT_MAX, b1 = FLT_MAX, b2 = FLT_MAX;
  for (int r = 0;; ++r) {
    const int z0 = max(0, c.z - r), z1 = min(G.z - 1, c.z + r);
    const int y0 = max(0, c.y - r), y1 = min(G.y - 1, c.y + r);
    const int x0 = max(0, c.x - r), x1 = min(G.x - 1, c.x + r);
    for (int z = z0; z <= z1; ++z) {
      const bool z_shell = (z == c.z - r) || (z == c.z + r);
      for (int y = y0; y <= y1; ++y) {
        const bool zy_shell = z_shell || (y == c.y - r) || (y == c.y + r);
        // on a z/y shell face the whole x run belongs to ring r; otherwise only its two ends do
        const int step = zy_shell ? 1 : max(1, x1 - x0);
        for (int x = x0; x <= x1; x += step) {
          if (!zy_shell && x != c.x - r && x != c.x + r) continue;
          const int cell = (z * G.y + y) * G.x + x;
          const uint32_t first = cell == 0 ? 0u : cell_end[cell - 1];
          const uint32_t last = cell_end[cell];
          for (uint32_t j = first; j < last; ++j) {
            const float4 p = sorted[j];
            float dist = knn_dist2(ref, p.x, p.y, p.z);
            if (j == (uint32_t)s) dist = FLT_MAX;
            knn_insert(dist, b0, b1, b2);
          }
        }
      }
    }
    const float reach = (float)r * g.cell;
    if (b2 <= reach * reach) break;
    if (x0 == 0 && y0 == 0 && z0 == 0 && x1 == G.x - 1 && y1 == G.y - 1 && z1 == G.z - 1) break;   // whole grid seen
  }
  mean_dist2[__float_as_int(me.w)] = (b0 + b1 + b2) / 3.0f;
}

}  // namespace egs

extern "C" size_t egs_knn_workspace_bytes(int P) {
  if (P < egs::KNN_GRID_MIN) return 0;
  return egs::KnnWorkspace(nullptr, P, egs::knn_max_cells(P)).bytes;
}

extern "C" int egs_knn_mean_dist2(int P, const float* points, float* mean_dist2, void* workspace,
                                  size_t workspace_bytes, void* stream) {
  using namespace egs;
  if (P < 0 || (P > 0 && (!points || !mean_dist2))) {
    set_error("egs_knn_mean_dist2: NULL argument or negative P");
    return EGS_ERR_INVALID_ARGUMENT;
  }
  if (P == 0) return EGS_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (P < KNN_GRID_MIN || workspace == nullptr) {   // no scratch given: the exact sweep works at any size
    knn_all_pairs_kernel<<<(P + KNN_THREADS - 1) / KNN_THREADS, KNN_THREADS, 0, st>>>(P, points, mean_dist2);
    EGS_AFTER_LAUNCH(st, false);
    return EGS_OK;
  }
  const int cells = knn_max_cells(P);   // the grid the device settles on uses a prefix of these
  const KnnWorkspace w(workspace, P, cells);
  if (workspace_bytes < w.bytes || (reinterpret_cast<uintptr_t>(workspace) & 255u) != 0) {
    set_error("egs_knn_mean_dist2: workspace too small (%zu < %zu bytes) or not 256-byte aligned", workspace_bytes,
              w.bytes);
    return EGS_ERR_INVALID_ARGUMENT;
  }
  // sums, grid and the counts are cleared in one go (they are contiguous)
  EGS_CUDA_TRY(cudaMemsetAsync(w.sums, 0, reinterpret_cast<char*>(w.cell_end) - reinterpret_cast<char*>(w.sums), st));
  knn_moments_kernel<<<148 * 4, 256, 0, st>>>(P, points, w.sums);
  EGS_AFTER_LAUNCH(st, false);
  knn_grid_setup_kernel<<<1, 32, 0, st>>>(P, cells, w.sums, w.grid);
  EGS_AFTER_LAUNCH(st, false);
  knn_count_kernel<<<(P + 255) / 256, 256, 0, st>>>(P, points, w.grid, w.cell_count);
  EGS_AFTER_LAUNCH(st, false);
  if (int rc = launch_inclusive_scan(w.cell_count, w.cell_end, cells, w.scan_status, st)) return rc;
  knn_scatter_kernel<<<(P + 255) / 256, 256, 0, st>>>(P, points, w.grid, w.cell_end, w.cell_count, w.sorted);
  EGS_AFTER_LAUNCH(st, false);
  knn_grid_query_kernel<<<(P + 127) / 128, 128, 0, st>>>(P, w.sorted, w.cell_end, w.grid, mean_dist2);
  EGS_AFTER_LAUNCH(st, false);
  return EGS_OK;
}
