// common.cuh -- shared device helpers, workspace layouts and host-side error plumbing.
//
// Workspace layouts replace GeometryState / ImageState / BinningState of the reference
// (DGR/cuda_rasterizer/rasterizer_impl.h:22-73, rasterizer_impl.cu:155-193). They are our own:
// the blend kernels want rgb+depth packed in one float4 so that a tile stages 40 B per list
// entry with three vector loads, and cov3D is never stored (backward recomputes it).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>
#include "../../include/egs_rasterizer.h"

namespace egs {

constexpr int TILE_X = EGS_TILE_X;
constexpr int TILE_Y = EGS_TILE_Y;
constexpr int TILE_PIXELS = TILE_X * TILE_Y;

// ---------------------------------------------------------------------------------------------
// host-side plumbing
// ---------------------------------------------------------------------------------------------
void set_error(const char* fmt, ...);
void count_launch(unsigned n = 1);

#define EGS_CUDA_TRY(expr)                                                              \
  do {                                                                                  \
    cudaError_t _e = (expr);                                                            \
    if (_e != cudaSuccess) {                                                            \
      egs::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__,  \
                     __LINE__);                                                         \
      return EGS_ERR_CUDA;                                                              \
    }                                                                                   \
  } while (0)

// After a kernel launch: always pick up launch-configuration errors; in debug mode also
// synchronise, like CHECK_CUDA in DGR/cuda_rasterizer/auxiliary.h:166-173.
#define EGS_AFTER_LAUNCH(stream, debug)                                                 \
  do {                                                                                  \
    egs::count_launch();                                                                \
    EGS_CUDA_TRY(cudaGetLastError());                                                   \
    if (debug) EGS_CUDA_TRY(cudaStreamSynchronize(stream));                             \
  } while (0)

inline size_t align_up(size_t x, size_t a = 256) { return (x + a - 1) / a * a; }

// Stage ids of the built-in event timing (egs_profile_* in the header).
enum Stage {
  ST_PREPROCESS = 0, ST_SCAN, ST_DUPLICATE, ST_SORT_HIST, ST_SORT_PASSES, ST_RANGES, ST_BLEND_FWD,
  ST_BWD_ZERO, ST_BLEND_BWD, ST_PREPROCESS_BWD, ST_ADAM, ST_SE3, ST_MODEL_FWD, ST_MODEL_BWD, ST_LOSS,
  ST_DEPTH_ORDER, ST_COUNT
};
static_assert(ST_COUNT <= EGS_PROFILE_STAGES, "grow EGS_PROFILE_STAGES");
bool profiling_enabled();
void profile_begin(cudaStream_t stream, int stage, void** token);
void profile_end(cudaStream_t stream, void* token);
// RAII bracket: records an event pair around the launches issued in its scope
struct StageTimer {
  cudaStream_t stream;
  void* token = nullptr;
  StageTimer(cudaStream_t s, int stage) : stream(s) {
    if (profiling_enabled()) profile_begin(s, stage, &token);
  }
  ~StageTimer() {
    if (token) profile_end(stream, token);
  }
};

// ---------------------------------------------------------------------------------------------
// workspace carving (pure functions of the sizes, so backward can re-derive the pointers)
// ---------------------------------------------------------------------------------------------
constexpr int SCAN_TILE = 4096;   // elements per block of the tiles_touched scan
// presort aux, in u32 words: hist[3][1024] start[3][1024] ticket[8] meta[8], then status[3][blocks][1024]
constexpr int DEPTH_ORDER_AUX_WORDS = 6 * 1024 + 8 + 8;
constexpr int SORT_RADIX_BITS = 8;
constexpr int SORT_RADIX = 1 << SORT_RADIX_BITS;
constexpr int SORT_THREADS = 256;
constexpr int SORT_ITEMS = 16;
constexpr int SORT_TILE = SORT_THREADS * SORT_ITEMS;

struct GeomLayout {
  size_t depths, means2D, conic_opacity, rgb_depth, clamped, tiles_touched, point_offsets,
      scan_status, sort_keys[2], sort_vals[2], sort_aux, sort_blocks, sort_aux_bytes, readback, zero_bytes, total;
  explicit GeomLayout(size_t P) {
    size_t o = 0;
    auto take = [&](size_t bytes) { size_t r = o; o = align_up(o + bytes); return r; };
    depths = take(P * 4);
    means2D = take(P * 8);
    conic_opacity = take(P * 16);
    rgb_depth = take(P * 16);
    clamped = take(P);
    tiles_touched = take(P * 4);
    point_offsets = take(P * 4);
    // Everything a forward needs zeroed lies in one run, cleared by ONE memset before the first kernel
    // (every memset between two kernels costs the stream ~4 us): the look-back words of the scan, the
    // presort's aux words (digit histograms, tickets, flags, the status words of three partition passes)
    // and the 16-byte record the host reads back: [0] instance count (written by the scan), [1] the
    // presort's veto, [2] set when a Gaussian is culled although `prefiltered` promised otherwise.
    scan_status = take(((P + SCAN_TILE - 1) / SCAN_TILE + 1) * 8 + 256);
    sort_blocks = (P + SORT_TILE - 1) / SORT_TILE;
    sort_aux_bytes = (DEPTH_ORDER_AUX_WORDS + 3 * sort_blocks * 1024) * 4;
    sort_aux = take(sort_aux_bytes);
    readback = take(16);
    zero_bytes = readback + 16 - scan_status;
    // front-to-back order of the Gaussians (binning presort, binning_buckets.cu): packed words
    // (30-bit depth key | index) in two buffers and the resulting order
    sort_keys[0] = take(P * 8);
    sort_keys[1] = take(P * 8);
    sort_vals[0] = take(P * 4);
    sort_vals[1] = 0;
    total = o + 256;
  }
};

struct GeomPtrs {
  float* depths;
  float2* means2D;
  float4* conic_opacity;
  float4* rgb_depth;
  unsigned char* clamped;
  uint32_t* tiles_touched;
  uint32_t* point_offsets;
  unsigned long long* scan_status;  // [0] ticket, [1..] per-block status words
  uint64_t* sort_keys[2];           // depth-order presort; the order ends in sort_vals[0]
  uint32_t* sort_vals[2];
  uint32_t* sort_aux;
  uint32_t* readback;               // [0] instance count, [1] presort veto, [2] prefilter fault
  uint32_t* prefilter_fault;
  void* zero_begin;                 // scan_status .. readback: cleared by one memset per forward
  size_t sort_blocks, sort_aux_bytes, zero_bytes;
  GeomPtrs(void* base, size_t P) {
    GeomLayout L(P);
    char* b = static_cast<char*>(base);
    depths = reinterpret_cast<float*>(b + L.depths);
    means2D = reinterpret_cast<float2*>(b + L.means2D);
    conic_opacity = reinterpret_cast<float4*>(b + L.conic_opacity);
    rgb_depth = reinterpret_cast<float4*>(b + L.rgb_depth);
    clamped = reinterpret_cast<unsigned char*>(b + L.clamped);
    tiles_touched = reinterpret_cast<uint32_t*>(b + L.tiles_touched);
    point_offsets = reinterpret_cast<uint32_t*>(b + L.point_offsets);
    scan_status = reinterpret_cast<unsigned long long*>(b + L.scan_status);
    for (int k = 0; k < 2; ++k) sort_keys[k] = reinterpret_cast<uint64_t*>(b + L.sort_keys[k]);
    sort_vals[0] = reinterpret_cast<uint32_t*>(b + L.sort_vals[0]);
    sort_vals[1] = nullptr;
    sort_aux = reinterpret_cast<uint32_t*>(b + L.sort_aux);
    readback = reinterpret_cast<uint32_t*>(b + L.readback);
    prefilter_fault = readback + 2;
    zero_begin = b + L.scan_status;
    zero_bytes = L.zero_bytes;
    sort_blocks = L.sort_blocks;
    sort_aux_bytes = L.sort_aux_bytes;
  }
};

inline int tiles_x(int W) { return (W + TILE_X - 1) / TILE_X; }
inline int tiles_y(int H) { return (H + TILE_Y - 1) / TILE_Y; }

// Number of key bits above the 32 depth bits that take part in the sort: the same count the
// reference derives with getHigherMsb (rasterizer_impl.cu:35-50): the smallest b with
// (n >> b) == 0, i.e. bit_length(n).
inline int tile_id_bits(uint32_t n_tiles) {
  int b = 0;
  while (b < 32 && (n_tiles >> b)) ++b;
  return b;
}

constexpr int SORT_MAX_PASSES = 8;
// status / histogram words per sort block: 8 passes x 256 digits (pair sort) or 6 x 512 (packed sort)
constexpr int SORT_STATUS_WORDS = 3072;
// histogram area shared by the paths: digit histograms of all passes (<= 3072 words), or bucket
// histogram + bucket starts + largest-first order (1024 + 1025 + 1024 words)
constexpr int SORT_HIST_WORDS = 4096;

inline int sort_passes(int W, int H) {
  int bits = 32 + tile_id_bits((uint32_t)(tiles_x(W) * tiles_y(H)));
  return (bits + SORT_RADIX_BITS - 1) / SORT_RADIX_BITS;
}

struct BinningLayout {
  size_t keys[2], vals[2], hist, status, ticket, total;
  size_t sort_blocks;
  // `instances` is num_rendered or a capacity >= num_rendered; the layout only depends on it
  // rounded up to a multiple of SORT_TILE, and `total` is strictly increasing in that capacity
  explicit BinningLayout(size_t instances) {
    size_t o = 0;
    auto take = [&](size_t bytes) { size_t r = o; o = align_up(o + bytes); return r; };
    sort_blocks = (instances + SORT_TILE - 1) / SORT_TILE;
    const size_t R = sort_blocks * SORT_TILE;
    keys[0] = take(R * 8);
    keys[1] = take(R * 8);
    vals[0] = take(R * 4);
    vals[1] = take(R * 4);
    // everything from `hist` to the end is zeroed with one memset per forward
    hist = take(SORT_HIST_WORDS * 4);
    ticket = take(SORT_MAX_PASSES * 4);
    status = take(sort_blocks * SORT_STATUS_WORDS * 4);
    total = o + 256;
  }
};

struct BinningPtrs {
  uint64_t* keys[2];
  uint32_t* vals[2];
  uint32_t* hist;
  uint32_t* ticket;
  uint32_t* status;
  size_t sort_blocks;
  size_t zero_bytes;  // bytes to clear starting at `hist`
  BinningPtrs(void* base, size_t R) {
    BinningLayout L(R);
    char* b = static_cast<char*>(base);
    keys[0] = reinterpret_cast<uint64_t*>(b + L.keys[0]);
    keys[1] = reinterpret_cast<uint64_t*>(b + L.keys[1]);
    vals[0] = reinterpret_cast<uint32_t*>(b + L.vals[0]);
    vals[1] = reinterpret_cast<uint32_t*>(b + L.vals[1]);
    hist = reinterpret_cast<uint32_t*>(b + L.hist);
    ticket = reinterpret_cast<uint32_t*>(b + L.ticket);
    status = reinterpret_cast<uint32_t*>(b + L.status);
    sort_blocks = L.sort_blocks;
    zero_bytes = L.total - 256 - L.hist;
  }
};

struct ImageLayout {
  size_t n_contrib, ranges, total;
  ImageLayout(int W, int H) {
    size_t o = 0;
    auto take = [&](size_t bytes) { size_t r = o; o = align_up(o + bytes); return r; };
    n_contrib = take((size_t)W * H * 4);
    ranges = take((size_t)tiles_x(W) * tiles_y(H) * 8);
    total = o + 256;
  }
};

struct ImagePtrs {
  uint32_t* n_contrib;
  uint2* ranges;
  ImagePtrs(void* base, int W, int H) {
    ImageLayout L(W, H);
    char* b = static_cast<char*>(base);
    n_contrib = reinterpret_cast<uint32_t*>(b + L.n_contrib);
    ranges = reinterpret_cast<uint2*>(b + L.ranges);
  }
};

// Backward scratch: one 48-byte accumulator record per Gaussian, three float4 so the tile
// kernel can use vector reductions:
//   acc[0] = (dL/dmean2D.x, dL/dmean2D.y, dL/dopacity, dL/ddepth)
//   acc[1] = (dL/dconic.x, dL/dconic.y, dL/dconic.w, unused)
//   acc[2] = (dL/dcolor.r, .g, .b, unused)
struct BackwardScratch {
  float4* acc;  // [P][3]
  static size_t bytes(size_t P) { return align_up(P * 48) + 256; }
  explicit BackwardScratch(void* base) : acc(static_cast<float4*>(base)) {}
};

// ---------------------------------------------------------------------------------------------
// stage entry points (each enqueues kernels on `stream`; return EgsStatus)
// ---------------------------------------------------------------------------------------------
// `model` non-NULL selects the raw-parameter mode (egs_forward_raw / egs_backward_raw): opacity,
// covariance and SH rows are derived from the optimiser's raw tensors inside the kernel.
int launch_preprocess_forward(const EgsForwardArgs& a, const EgsModelArgs* model, const GeomPtrs& g,
                              cudaStream_t stream);
// order != NULL: the scan runs over tiles_touched[order[i]] (front-to-back order of the Gaussians)
int launch_scan_tiles(const GeomPtrs& g, int P, const uint32_t* order, cudaStream_t stream, bool debug);
// g.sort_vals[0] <- Gaussian indices by ascending depth bits, ties by index; culled ones last
int launch_depth_order(const GeomPtrs& g, int P, const int* radii, cudaStream_t stream, bool debug);
int launch_inclusive_scan(const uint32_t* in, uint32_t* out, int n, unsigned long long* status, cudaStream_t stream);
// packed-word fast path (binning_packed.cu) and the shared range kernel
bool packed_path_fits(int P, int W, int H);
// `order` (all binning launchers): NULL, or the front-to-back order the scan ran over; instances are
// then emitted in that order, which is what the sorters' stability turns into less work
int launch_binning_packed(const GeomPtrs& g, const BinningPtrs& b, const ImagePtrs& im, const int* radii,
                          const uint32_t* order, int P, int R, int W, int H, bool pairs, cudaStream_t stream,
                          bool debug);
// MSD-bucket path (binning_buckets.cu): one global partition pass + one CTA per bucket
bool bucket_path_ok(int P, int R, int W, int H);
// The number of instances as the device sees it: `n_dev` points at the last element of the
// tiles_touched scan, `cap` is what the workspace was sized for. Kernels enqueued before the host
// knows the count read it here and skip themselves when it does not fit.
// `veto` (optional) is the flag the depth presort raises when its order is not the depth order (a
// depth outside its key range): the presorted path then skips itself too and forward re-runs the
// binning without relying on that order.
struct InstanceCount {
  const uint32_t* n_dev;
  uint32_t cap;
  const uint32_t* veto;
#ifdef __CUDACC__
  __device__ __forceinline__ bool get(uint32_t& n) const {
    n = *n_dev;
    return n <= cap && !(veto != nullptr && *veto != 0u);
  }
#endif
};
// presorted: `order` is the front-to-back order of launch_depth_order and the instances carry no
// depth bits (tile | index words of 32 or 64 bits: any number of Gaussians); otherwise full
// tile | depth | index words, which need packed_path_fits().
bool presorted_path_ok(int P, int W, int H);
const uint32_t* presort_veto_flag(const GeomPtrs& g);
int launch_binning_buckets(const GeomPtrs& g, const BinningPtrs& b, const ImagePtrs& im, const int* radii,
                           const uint32_t* order, bool presorted, int P, int capacity, InstanceCount count, int W,
                           int H, cudaStream_t stream, bool debug, bool zeroed);
int zero_binning_buckets(const BinningPtrs& b, cudaStream_t stream);
// keys[i] = tile << 32 | depth bits of point_list[i], rebuilt from the tile ranges (tests only: no
// kernel of the path reads the sorted keys, so the fast paths do not write them)
int launch_binning_keys(const GeomPtrs& g, const ImagePtrs& im, const uint32_t* point_list, uint64_t* keys, int W,
                        int H, cudaStream_t stream);

// Switches for tests and A/B timing, set through egs_debug_option() only: the library never reads
// the process environment, so nothing outside the caller's code can change which kernels run.
enum SortPath { SORT_AUTO = 0, SORT_PACKED, SORT_PAIRS, SORT_PAIRS8 };
struct DebugOptions {
  int sort;               // SortPath
  int presort;            // -1 automatic, 0 never, 1 whenever the scene has >= 64 Gaussians
  int presort_scramble;   // 1: presort keys unrelated to depth + veto raised (exercises the fallback)
  int fwd_v2;             // 1: the two-pixels-per-thread forward blend (independent implementation)
  int exchange_phases;    // timing only: 1 = only the slice sum, 2 = only the SH expansion, 0 = both
  int bwd_ppl;            // pixels per lane of the backward blend: 0 default (4), 2 for the A/B
};
DebugOptions debug_options();
int launch_tile_ranges(int R, const uint64_t* sorted_keys, const ImagePtrs& im, cudaStream_t stream,
                       bool debug);
int launch_binning(const GeomPtrs& g, const BinningPtrs& b, const ImagePtrs& im, const int* radii,
                   int P, int R, int W, int H, cudaStream_t stream, bool debug);
int launch_blend_forward(const EgsForwardArgs& a, const GeomPtrs& g, const uint32_t* point_list,
                         const ImagePtrs& im, cudaStream_t stream);
int launch_blend_backward(const EgsBackwardArgs& a, const GeomPtrs& g, const uint32_t* point_list,
                          const ImagePtrs& im, const BackwardScratch& s, cudaStream_t stream);
int launch_preprocess_backward(const EgsBackwardArgs& a, const EgsModelArgs* model, const EgsModelGrads* grads,
                               const GeomPtrs& g, const BackwardScratch& s, cudaStream_t stream);
int launch_mark_visible(int P, const float* means3D, const float* view, const float* proj,
                        unsigned char* present, cudaStream_t stream);

}  // namespace egs

// ---------------------------------------------------------------------------------------------
// device helpers
// ---------------------------------------------------------------------------------------------
#ifdef __CUDACC__
namespace egs {

// SH basis constants (real spherical harmonics, same values as utils/sh_utils.py:24-54 and
// DGR/cuda_rasterizer/auxiliary.h:22-39).
__device__ constexpr float kSH0 = 0.28209479177387814f;
__device__ constexpr float kSH1 = 0.4886025119029199f;
__device__ constexpr float kSH2[5] = {1.0925484305920792f, -1.0925484305920792f,
                                      0.31539156525252005f, -1.0925484305920792f,
                                      0.5462742152960396f};
__device__ constexpr float kSH3[7] = {-0.5900435899266435f, 2.890611442640554f,
                                      -0.4570457994644658f, 0.3731763325901154f,
                                      -0.4570457994644658f, 1.445305721320277f,
                                      -0.5900435899266435f};

// Values of the real SH basis functions in the unit direction (x, y, z), coefficients 0 .. (deg+1)^2 - 1
// (backward.cu:20-139 / forward.cu:20-71, grouped per coefficient). One definition for the two places
// that need dL/dsh[k] = basis_k * dL/dRGB: the per-Gaussian backward kernel and the factored gradient
// exchange, which must form bit-identical products.
__device__ __forceinline__ void sh_basis_values(int deg, float x, float y, float z, float* b) {
  b[0] = kSH0;
  if (deg > 0) {
    b[1] = -kSH1 * y;
    b[2] = kSH1 * z;
    b[3] = -kSH1 * x;
    if (deg > 1) {
      const float xx = x * x, yy = y * y, zz = z * z;
      const float xy = x * y, yz = y * z, xz = x * z;
      b[4] = kSH2[0] * xy;
      b[5] = kSH2[1] * yz;
      b[6] = kSH2[2] * (2.f * zz - xx - yy);
      b[7] = kSH2[3] * xz;
      b[8] = kSH2[4] * (xx - yy);
      if (deg > 2) {
        b[9] = kSH3[0] * y * (3.f * xx - yy);
        b[10] = kSH3[1] * xy * z;
        b[11] = kSH3[2] * y * (4.f * zz - xx - yy);
        b[12] = kSH3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy);
        b[13] = kSH3[4] * x * (4.f * zz - xx - yy);
        b[14] = kSH3[5] * z * (xx - yy);
        b[15] = kSH3[6] * x * (xx - 3.f * yy);
      }
    }
  }
}

// Column-major 3x3 with the same storage convention and the same product evaluation order as
// the reference's matrix library (glm/detail/type_mat3x3.inl:486-520): entry (col c, row r) is
// c[c][r], and (A*B)[c][r] = A[0][r]*B[c][0] + A[1][r]*B[c][1] + A[2][r]*B[c][2], summed left to
// right. Bit-exact radii and sort keys depend on this order.
struct Mat3 {
  float c[3][3];
};

__device__ __forceinline__ Mat3 mat3_cols(float a0, float a1, float a2, float b0, float b1,
                                          float b2, float c0, float c1, float c2) {
  Mat3 m;
  m.c[0][0] = a0; m.c[0][1] = a1; m.c[0][2] = a2;
  m.c[1][0] = b0; m.c[1][1] = b1; m.c[1][2] = b2;
  m.c[2][0] = c0; m.c[2][1] = c1; m.c[2][2] = c2;
  return m;
}

__device__ __forceinline__ Mat3 mul(const Mat3& A, const Mat3& B) {
  Mat3 R;
#pragma unroll
  for (int col = 0; col < 3; ++col)
#pragma unroll
    for (int row = 0; row < 3; ++row)
      R.c[col][row] = A.c[0][row] * B.c[col][0] + A.c[1][row] * B.c[col][1] + A.c[2][row] * B.c[col][2];
  return R;
}

__device__ __forceinline__ Mat3 transpose(const Mat3& A) {
  Mat3 R;
#pragma unroll
  for (int col = 0; col < 3; ++col)
#pragma unroll
    for (int row = 0; row < 3; ++row) R.c[col][row] = A.c[row][col];
  return R;
}

// Row-vector 4x4 transforms on the transposed matrices the reference passes around
// (auxiliary.h:58-76): memory index 4*c + r.
__device__ __forceinline__ float3 xform_point_4x3(const float3& p, const float* m) {
  float3 r;
  r.x = m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12];
  r.y = m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13];
  r.z = m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14];
  return r;
}

__device__ __forceinline__ float4 xform_point_4x4(const float3& p, const float* m) {
  float4 r;
  r.x = m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12];
  r.y = m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13];
  r.z = m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14];
  r.w = m[3] * p.x + m[7] * p.y + m[11] * p.z + m[15];
  return r;
}

// Pixel centre from NDC; evaluated in double like auxiliary.h:41-44 (double literals).
__device__ __forceinline__ float ndc_to_pix(float v, int S) {
  return ((v + 1.0) * S - 1.0) * 0.5;
}

// Tile rectangle covered by a disc of integer radius r around p (auxiliary.h:46-56).
__device__ __forceinline__ void tile_rect(const float2 p, int r, int gx, int gy, uint2& lo, uint2& hi) {
  lo.x = min(gx, max(0, (int)((p.x - r) / TILE_X)));
  lo.y = min(gy, max(0, (int)((p.y - r) / TILE_Y)));
  hi.x = min(gx, max(0, (int)((p.x + r + TILE_X - 1) / TILE_X)));
  hi.y = min(gy, max(0, (int)((p.y + r + TILE_Y - 1) / TILE_Y)));
}

__device__ __forceinline__ unsigned lane_id() { return threadIdx.x & 31; }

// Which BW x BH pixel blocks of a 16x16 tile can a Gaussian contribute to? Bit (by * NBX + bx) of
// the result is set when the block may contain a pixel with alpha = o * exp(power) >= 1/255, where
// power = -0.5 (A dx^2 + C dy^2) - B dx dy. That region is the ellipse A dx^2 + 2B dx dy + C dy^2
// <= 2 log(255 o), whose bounding box has half-extents sqrt(k C / det), sqrt(k A / det). The test
// is conservative, so skipping an unmarked block never changes a result: the per-pixel test
// evaluates `power` in fp32 with an absolute error of at most ~2e-7 * k * AC/det near the boundary
// (each of its three terms is bounded by k * AC/det there), so with AC/det <= 100 and k <= ~11 a
// margin of 1e-3 on the log covers it five times over; the extents get another 1e-3 px. Anything
// unusual (non-positive or non-finite opacity, degenerate or very elongated conic, non-finite
// centre) marks every block and leaves the decision to the exact per-pixel test. The tile kernels use one warp per block, so a warp skips a whole list entry with
// one shared-memory byte instead of evaluating 32 or 64 pixels.
template <int BW, int BH>
__device__ __forceinline__ unsigned footprint_mask(const float2 xy, const float4 conic_o, float tile_x0,
                                                   float tile_y0) {
  constexpr int NBX = TILE_X / BW, NBY = TILE_Y / BH;
  constexpr unsigned ALL = (1u << (NBX * NBY)) - 1u;
  const float A = conic_o.x, B = conic_o.y, C = conic_o.z, o = conic_o.w;
  const float det = A * C - B * B;
  const float k = 2.f * (__logf(255.f * o) + 1e-3f);
  const bool regular = o > 0.f && det > 0.f && A > 0.f && C > 0.f && det * 100.f >= A * C && isfinite(k) &&
                       isfinite(det) && isfinite(xy.x) && isfinite(xy.y);
  if (!regular) return ALL;
  if (k <= 0.f) return 0u;   // o < 1/255: alpha can never reach the threshold
  const float ex = sqrtf(k * C / det) * 1.0001f + 1e-3f;
  const float ey = sqrtf(k * A / det) * 1.0001f + 1e-3f;
  if (!(isfinite(ex) && isfinite(ey))) return ALL;
  const float lo_x = xy.x - ex - tile_x0, hi_x = xy.x + ex - tile_x0;   // tile-local pixel range
  const float lo_y = xy.y - ey - tile_y0, hi_y = xy.y + ey - tile_y0;
  // Blocks that pass the bounding-box test are then tested against the ellipse itself: the smallest
  // q(d) = A dx^2 + 2 B dx dy + C dy^2 over the block's rectangle of pixel centres (0 when the centre
  // lies inside, otherwise attained on one of the four edges, where q is a 1-D parabola) must not
  // exceed k. Rotated, elongated footprints fill their bounding box poorly, so this removes
  // another share of the (entry, block) pairs. The continuous minimum is <= the minimum over the
  // pixels and the comparison carries a 1 % + 0.05 margin (fp32 cancellation in q stays below
  // 1e-6 x the 3 terms <= 300 k), so the test remains conservative.
  const float cx = xy.x - tile_x0, cy = xy.y - tile_y0;   // centre in tile-local pixel coordinates
  const float inv_A = 1.f / A, inv_C = 1.f / C;
  const float k_lim = k * 1.01f + 0.05f;
  unsigned mask = 0u;
#pragma unroll
  for (int by = 0; by < NBY; ++by) {
    const bool ytest = lo_y <= (float)(by * BH + BH - 1) && hi_y >= (float)(by * BH);
#pragma unroll
    for (int bx = 0; bx < NBX; ++bx) {
      const bool xtest = lo_x <= (float)(bx * BW + BW - 1) && hi_x >= (float)(bx * BW);
      if (ytest && xtest) {
        const float x0 = (float)(bx * BW) - cx, x1 = (float)(bx * BW + BW - 1) - cx;
        const float y0 = (float)(by * BH) - cy, y1 = (float)(by * BH + BH - 1) - cy;
        bool hit = x0 <= 0.f && x1 >= 0.f && y0 <= 0.f && y1 >= 0.f;
        if (!hit) {
          auto on_vertical = [&](float xe) {   // min over y in [y0, y1] at x = xe
            const float ys = fminf(fmaxf(-B * xe * inv_C, y0), y1);
            return (A * xe + 2.f * B * ys) * xe + C * ys * ys;
          };
          auto on_horizontal = [&](float ye) {
            const float xs = fminf(fmaxf(-B * ye * inv_A, x0), x1);
            return (C * ye + 2.f * B * xs) * ye + A * xs * xs;
          };
          const float q = fminf(fminf(on_vertical(x0), on_vertical(x1)), fminf(on_horizontal(y0), on_horizontal(y1)));
          hit = !(q > k_lim);   // NaN counts as a hit
        }
        if (hit) mask |= 1u << (by * NBX + bx);
      }
    }
  }
  return mask;
}

// The exponent of a splat at offset (dx, dy) from its centre, conic (A, B, C):
//   -0.5 (A dx^2 + C dy^2) - B dx dy        (forward.cu:330-331, backward.cu:505-506).
// A long, rotated splat makes the three terms cancel (hundreds against a sum of a few units), so the
// last bits of the result depend on where the compiler fuses multiplies and adds -- and alpha = o
// exp(power) has to come out the same in the backward pass as in the forward pass, or T = T / (1 - alpha)
// drifts along the pixel's list (measured: 2e-4 instead of 2e-6 of the gradient maximum on the big
// splats of a 2M-Gaussian scene, flipping with the register allocation). The roundings are therefore
// written out, in the order nvcc gives the reference's expression (checked in the SASS of its forward
// kernel; the images stay bit-identical to the reference's), and every blend kernel uses this one
// definition. ax = A dx and bx = B dx may be formed once per entry for several pixels of a column.
__device__ __forceinline__ float splat_power_pre(float ax, float bx, float C, float dx, float dy) {
  const float cyy = __fmul_rn(dy, __fmul_rn(C, dy));
  const float bxy = __fmul_rn(dy, bx);
  return __fmaf_rn(__fmaf_rn(dx, ax, cyy), -0.5f, -bxy);
}
__device__ __forceinline__ float splat_power(float A, float B, float C, float dx, float dy) {
  return splat_power_pre(__fmul_rn(A, dx), __fmul_rn(B, dx), C, dx, dy);
}

// Shared-memory loads through a 32-bit shared-window address that is formed once outside the hot
// loop. Indexing a __shared__ array inside the tile loops makes nvcc re-derive the window base
// (S2R SR_CgaCtaId + LEA, ~4 instructions per access on sm_100) on every iteration.
__device__ __forceinline__ uint32_t smem_addr(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ float2 lds_f2(uint32_t addr) {
  float2 v;
  asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ float4 lds_f4(uint32_t addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ void sts_f32(uint32_t addr, float v) {
  asm volatile("st.shared.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory");
}
__device__ __forceinline__ uint32_t lds_u32(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
  return v;
}

// Lanes of the warp that are valid and hold the same BITS-bit digit as this lane. One ballot per
// digit bit (fixed cost, independent ballots pipeline) instead of the match.any instruction, whose
// cost grows with the number of distinct values in the warp.
template <int BITS>
__device__ __forceinline__ uint32_t match_digit(uint32_t d, bool ok) {
  uint32_t peers = __ballot_sync(0xffffffffu, ok);
#pragma unroll
  for (int b = 0; b < BITS; ++b) {
    const bool bit = (d >> b) & 1u;
    const uint32_t bal = __ballot_sync(0xffffffffu, bit);
    peers &= bit ? bal : ~bal;
  }
  return peers;
}

}  // namespace egs
#endif  // __CUDACC__
