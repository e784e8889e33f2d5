// blend_fwd.cu -- per-tile front-to-back alpha blend: colour, un-normalised depth, alpha, and
// the per-pixel contributor count that backward needs.
//
// Semantics follow the reference's forward renderCUDA (DGR/cuda_rasterizer/forward.cu:261-381)
// as restated in SURVEY.md App. A.7. One CTA per 16x16 tile, one thread per pixel. Differences
// in mechanism (not in results):
//   * a batch of 256 list entries is staged as xy (8 B) + conic/opacity (16 B) + rgb/depth
//     (16 B) in shared memory, so the inner loop never touches global memory (the reference
//     gathers colour and depth from global per blended fragment);
//   * each warp covers an 8x4 pixel block instead of 16x2, which makes whole-warp skips and
//     whole-warp termination more likely; a warp whose 32 pixels are done skips the inner loop,
//     and a warp only visits the list entries whose footprint can reach its block at all;
//   * the per-pixel arithmetic keeps the reference's expression shapes, because the alpha and
//     transmittance thresholds make the image sensitive to the last bit of `power`.
#include "common.cuh"

namespace egs {

constexpr int BLEND_BATCH = 256;

// The staging warps turn the footprint masks of their 32 entries (common.cuh: footprint_mask) into
// one 32-bit word per pixel warp (eight ballots), so a pixel warp walks only the entries that can
// reach its 8x4 block: it pops set bits in list order and never issues an instruction for the
// others (in the first version, which tested one mask byte per entry, a skipped entry still cost
// ~13 instructions, and more than half of all entries are skipped: 0.318 -> 0.245 ms). Shared
// memory is addressed through window addresses formed once (see lds_f4). The arithmetic of a
// blended fragment keeps the reference's expression shapes, so the images are bit-identical.
__global__ void __launch_bounds__(TILE_PIXELS, 6)
blend_forward3_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list,
                      int W, int H, int grid_x, const float2* __restrict__ means2D,
                      const float4* __restrict__ conic_opacity, const float4* __restrict__ rgb_depth,
                      const float* __restrict__ bg_color, float* __restrict__ out_color,
                      float* __restrict__ out_depth, float* __restrict__ out_alpha,
                      uint32_t* __restrict__ n_contrib) {
  __shared__ float2 s_xy[BLEND_BATCH];
  __shared__ float4 s_co[BLEND_BATCH];
  __shared__ float4 s_cd[BLEND_BATCH];
  __shared__ uint32_t s_bits[8][BLEND_BATCH / 32];   // [pixel warp][32-entry chunk]

  const unsigned tid = threadIdx.x;
  const unsigned warp = tid >> 5, lane = tid & 31;
  const unsigned lx = (warp & 1) * 8 + (lane & 7);
  const unsigned ly = (warp >> 1) * 4 + (lane >> 3);
  const unsigned px = blockIdx.x * TILE_X + lx;
  const unsigned py = blockIdx.y * TILE_Y + ly;
  const bool inside = px < (unsigned)W && py < (unsigned)H;
  const uint32_t pix_id = (uint32_t)W * py + px;
  const float2 pixf = make_float2((float)px, (float)py);

  const uint2 range = ranges[blockIdx.y * grid_x + blockIdx.x];
  const int total = (int)(range.y - range.x);
  const int rounds = (total + BLEND_BATCH - 1) / BLEND_BATCH;

  bool done = !inside;
  float T = 1.0f;
  uint32_t last_contributor = 0;
  const float tile_x0 = (float)(blockIdx.x * TILE_X), tile_y0 = (float)(blockIdx.y * TILE_Y);
  float C0 = 0.f, C1 = 0.f, C2 = 0.f, weight = 0.f, D = 0.f;
  uint32_t a_xy = smem_addr(s_xy), a_co = smem_addr(s_co), a_cd = smem_addr(s_cd);
  uint32_t a_bits = smem_addr(&s_bits[warp][0]);
  // opaque to the optimiser, or it re-derives the window addresses inside the loop instead of
  // keeping four registers
  asm volatile("" : "+r"(a_xy), "+r"(a_co), "+r"(a_cd), "+r"(a_bits));

  for (int r = 0; r < rounds; ++r) {
    if (__syncthreads_count(done) == TILE_PIXELS) break;
    const int progress = r * BLEND_BATCH + (int)tid;
    unsigned mask = 0u;
    if (progress < total) {
      const uint32_t id = point_list[range.x + progress];
      const float2 xy = means2D[id];
      const float4 co = conic_opacity[id];
      s_xy[tid] = xy;
      s_co[tid] = co;
      s_cd[tid] = rgb_depth[id];
      mask = footprint_mask<8, 4>(xy, co, tile_x0, tile_y0);
    }
    uint32_t mine = 0u;   // lane w ends up with the word of pixel warp w for this chunk
#pragma unroll
    for (unsigned w = 0; w < 8; ++w) {
      const uint32_t word = __ballot_sync(0xffffffffu, (mask >> w) & 1u);
      if (lane == w) mine = word;
    }
    if (lane < 8) s_bits[lane][warp] = mine;
    __syncthreads();
    if (__all_sync(0xffffffffu, done)) continue;  // warp-uniform: nothing left for these 32 pixels

    const uint32_t first = (uint32_t)(r * BLEND_BATCH) + 1u;
    for (uint32_t c = 0; c < BLEND_BATCH / 32 && !done; ++c) {
      uint32_t bits = lds_u32(a_bits + 4u * c);
      while (bits != 0u) {   // a pixel that saturates clears its own copy of the word
        const uint32_t j = c * 32u + (uint32_t)__ffs((int)bits) - 1u;
        bits &= bits - 1u;
        const float2 xy = lds_f2(a_xy + 8u * j);
        const float4 con_o = lds_f4(a_co + 16u * j);
        const float2 d = make_float2(xy.x - pixf.x, xy.y - pixf.y);
        const float power = splat_power(con_o.x, con_o.y, con_o.z, d.x, d.y);
        if (power > 0.0f) continue;
        const float alpha = min(0.99f, con_o.w * expf(power));
        if (alpha < 1.0f / 255.0f) continue;
        const float test_T = T * (1 - alpha);
        if (test_T < 0.0001f) {
          done = true;
          bits = 0u;
          continue;
        }
        const float4 cd = lds_f4(a_cd + 16u * j);
        C0 += cd.x * alpha * T;
        C1 += cd.y * alpha * T;
        C2 += cd.z * alpha * T;
        weight += alpha * T;
        D += cd.w * alpha * T;
        T = test_T;
        last_contributor = first + j;   // 1-based position in the tile's list
      }
    }
  }

  if (inside) {
    const size_t plane = (size_t)H * W;
    n_contrib[pix_id] = last_contributor;
    out_color[pix_id] = C0 + T * bg_color[0];
    out_color[plane + pix_id] = C1 + T * bg_color[1];
    out_color[2 * plane + pix_id] = C2 + T * bg_color[2];
    out_alpha[pix_id] = weight;
    out_depth[pix_id] = D;
  }
}

// ---------------------------------------------------------------------------------------------
// v2: two pixels per thread + a conservative log-domain pre-filter
// ---------------------------------------------------------------------------------------------
// Bit-identical results to the kernel above (and so to the reference): every fragment that is
// blended goes through the same expression shapes (exact expf, same products and sums). What
// changes is only how much work surrounds them:
//   * 128 threads per tile; a thread owns (x, y) and (x, y + 4) of its warp's 8x8 block, so the
//     staged entry is read from shared memory once per two pixels and the loop overhead halves;
//   * `alpha = o * exp(power) < 1/255` is decided without the exponential whenever
//     power < thr - 1e-4 with thr = -log(255 o): expf is accurate to 2 ulp and the threshold to a
//     few 1e-7, so everything below that margin is certainly skipped by the reference too;
//     fragments inside the margin take the exact test. No decision can differ.
constexpr int FWD2_THREADS = 128;
constexpr int FWD2_BATCH = 256;

struct FwdPixel {
  float px, py, T, C0, C1, C2, weight, D;
  uint32_t last;
  bool done;
};

__device__ __forceinline__ void fwd_fragment(FwdPixel& s, const float4 a4, const float4 con_o, const float4 cd,
                                             uint32_t contributor) {
  const float2 d = make_float2(a4.x - s.px, a4.y - s.py);
  const float power = splat_power(con_o.x, con_o.y, con_o.z, d.x, d.y);
  if (power > 0.0f) return;
  if (power < a4.z) return;   // certainly alpha < 1/255 (conservative margin, see above)
  const float alpha = min(0.99f, con_o.w * expf(power));
  if (alpha < 1.0f / 255.0f) return;
  const float test_T = s.T * (1 - alpha);
  if (test_T < 0.0001f) {
    s.done = true;
    return;
  }
  s.C0 += cd.x * alpha * s.T;
  s.C1 += cd.y * alpha * s.T;
  s.C2 += cd.z * alpha * s.T;
  s.weight += alpha * s.T;
  s.D += cd.w * alpha * s.T;
  s.T = test_T;
  s.last = contributor;
}

__global__ void __launch_bounds__(FWD2_THREADS)
blend_forward2_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list,
                      int W, int H, int grid_x, const float2* __restrict__ means2D,
                      const float4* __restrict__ conic_opacity, const float4* __restrict__ rgb_depth,
                      const float* __restrict__ bg_color, float* __restrict__ out_color,
                      float* __restrict__ out_depth, float* __restrict__ out_alpha,
                      uint32_t* __restrict__ n_contrib) {
  __shared__ float4 s_a[FWD2_BATCH];   // mean.x, mean.y, skip threshold, -
  __shared__ float4 s_co[FWD2_BATCH];
  __shared__ float4 s_cd[FWD2_BATCH];

  const unsigned tid = threadIdx.x;
  const unsigned warp = tid >> 5, lane = tid & 31;
  const unsigned bx = blockIdx.x * TILE_X + (warp & 1) * 8 + (lane & 7);
  const unsigned by = blockIdx.y * TILE_Y + (warp >> 1) * 8 + (lane >> 3);
  FwdPixel P0, P1;
  const bool in0 = bx < (unsigned)W && by < (unsigned)H;
  const bool in1 = bx < (unsigned)W && (by + 4) < (unsigned)H;
  P0 = FwdPixel{(float)bx, (float)by, 1.0f, 0.f, 0.f, 0.f, 0.f, 0.f, 0u, !in0};
  P1 = FwdPixel{(float)bx, (float)(by + 4), 1.0f, 0.f, 0.f, 0.f, 0.f, 0.f, 0u, !in1};

  const uint2 range = ranges[blockIdx.y * grid_x + blockIdx.x];
  const int total = (int)(range.y - range.x);
  const int rounds = (total + FWD2_BATCH - 1) / FWD2_BATCH;
  int to_do = total;
  uint32_t contributor = 0;

  for (int r = 0; r < rounds; ++r, to_do -= FWD2_BATCH) {
    if (__syncthreads_count(P0.done && P1.done) == FWD2_THREADS) break;
#pragma unroll
    for (int h = 0; h < FWD2_BATCH / FWD2_THREADS; ++h) {
      const int slot = h * FWD2_THREADS + (int)tid;
      const int progress = r * FWD2_BATCH + slot;
      if (progress < total) {
        const uint32_t id = point_list[range.x + progress];
        const float2 xy = means2D[id];
        const float4 co = conic_opacity[id];
        // below this power the fragment's alpha is certainly < 1/255; NaN / non-positive opacity
        // disables the pre-filter (the comparison power < NaN is false) and leaves the exact test
        const float thr = co.w > 0.f ? -logf(255.f * co.w) - 1e-4f : __int_as_float(0x7fc00000);
        s_a[slot] = make_float4(xy.x, xy.y, thr, 0.f);
        s_co[slot] = co;
        s_cd[slot] = rgb_depth[id];
      }
    }
    __syncthreads();
    if (__all_sync(0xffffffffu, P0.done && P1.done)) continue;

    const int n = min(FWD2_BATCH, to_do);
    for (int j = 0; j < n; ++j) {
      if (P0.done && P1.done) break;
      contributor++;
      const float4 a4 = s_a[j];
      const float4 con_o = s_co[j];
      const float4 cd = s_cd[j];
      if (!P0.done) fwd_fragment(P0, a4, con_o, cd, contributor);
      if (!P1.done) fwd_fragment(P1, a4, con_o, cd, contributor);
    }
  }

  const size_t plane = (size_t)H * W;
  if (in0) {
    const uint32_t pix = (uint32_t)W * by + bx;
    n_contrib[pix] = P0.last;
    out_color[pix] = P0.C0 + P0.T * bg_color[0];
    out_color[plane + pix] = P0.C1 + P0.T * bg_color[1];
    out_color[2 * plane + pix] = P0.C2 + P0.T * bg_color[2];
    out_alpha[pix] = P0.weight;
    out_depth[pix] = P0.D;
  }
  if (in1) {
    const uint32_t pix = (uint32_t)W * (by + 4) + bx;
    n_contrib[pix] = P1.last;
    out_color[pix] = P1.C0 + P1.T * bg_color[0];
    out_color[plane + pix] = P1.C1 + P1.T * bg_color[1];
    out_color[2 * plane + pix] = P1.C2 + P1.T * bg_color[2];
    out_alpha[pix] = P1.weight;
    out_depth[pix] = P1.D;
  }
}

int launch_blend_forward(const EgsForwardArgs& a, const GeomPtrs& g, const uint32_t* point_list,
                         const ImagePtrs& im, cudaStream_t stream) {
  const dim3 grid(tiles_x(a.width), tiles_y(a.height), 1);
  StageTimer timer(stream, ST_BLEND_FWD);
  // Default: one pixel per thread, per-warp bit lists. egs_debug_option("blend_forward", "v2") selects
  // the older two-pixels-per-thread kernel without bit lists, kept as an independent implementation
  // for tests/test_parity_gpu.py::test_kernel_variants_agree (bit-identical results, slower).
  if (!debug_options().fwd_v2) {
    blend_forward3_kernel<<<grid, TILE_PIXELS, 0, stream>>>(
        im.ranges, point_list, a.width, a.height, (int)grid.x, g.means2D, g.conic_opacity,
        g.rgb_depth, a.background, a.out_color, a.out_depth, a.out_alpha, im.n_contrib);
  } else {
    blend_forward2_kernel<<<grid, FWD2_THREADS, 0, stream>>>(
        im.ranges, point_list, a.width, a.height, (int)grid.x, g.means2D, g.conic_opacity,
        g.rgb_depth, a.background, a.out_color, a.out_depth, a.out_alpha, im.n_contrib);
  }
  EGS_AFTER_LAUNCH(stream, a.debug);
  return EGS_OK;
}

}  // namespace egs
