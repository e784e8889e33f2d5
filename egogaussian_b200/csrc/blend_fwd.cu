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
//     whole-warp termination more likely; a warp whose 32 pixels are done skips the inner loop;
//   * the per-pixel arithmetic keeps the reference's expression shapes, because the alpha and
//     transmittance thresholds make the image sensitive to the last bit of `power`.
#include "common.cuh"

namespace egs {

constexpr int BLEND_BATCH = 256;

__global__ void __launch_bounds__(TILE_PIXELS)
blend_forward_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list,
                     int W, int H, int grid_x, const float2* __restrict__ means2D,
                     const float4* __restrict__ conic_opacity, const float4* __restrict__ rgb_depth,
                     const float* __restrict__ bg_color, float* __restrict__ out_color,
                     float* __restrict__ out_depth, float* __restrict__ out_alpha,
                     uint32_t* __restrict__ n_contrib) {
  __shared__ float2 s_xy[BLEND_BATCH];
  __shared__ float4 s_co[BLEND_BATCH];
  __shared__ float4 s_cd[BLEND_BATCH];

  const unsigned tid = threadIdx.x;
  const unsigned warp = tid >> 5, lane = tid & 31;
  // 8 warps as a 2 (x) by 4 (y) arrangement of 8x4-pixel blocks
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
  int to_do = total;

  bool done = !inside;
  float T = 1.0f;
  uint32_t contributor = 0, last_contributor = 0;
  float C0 = 0.f, C1 = 0.f, C2 = 0.f, weight = 0.f, D = 0.f;

  for (int r = 0; r < rounds; ++r, to_do -= BLEND_BATCH) {
    if (__syncthreads_count(done) == TILE_PIXELS) break;
    const int progress = r * BLEND_BATCH + (int)tid;
    if (progress < total) {
      const uint32_t id = point_list[range.x + progress];
      s_xy[tid] = means2D[id];
      s_co[tid] = conic_opacity[id];
      s_cd[tid] = rgb_depth[id];
    }
    __syncthreads();
    if (__all_sync(0xffffffffu, done)) continue;  // warp-uniform: nothing left for these 32 pixels

    const int n = min(BLEND_BATCH, to_do);
    for (int j = 0; !done && j < n; ++j) {
      contributor++;
      const float2 xy = s_xy[j];
      const float2 d = make_float2(xy.x - pixf.x, xy.y - pixf.y);
      const float4 con_o = s_co[j];
      const float power = -0.5f * (con_o.x * d.x * d.x + con_o.z * d.y * d.y) - con_o.y * d.x * d.y;
      if (power > 0.0f) continue;
      const float alpha = min(0.99f, con_o.w * expf(power));
      if (alpha < 1.0f / 255.0f) continue;
      const float test_T = T * (1 - alpha);
      if (test_T < 0.0001f) {
        done = true;
        continue;
      }
      const float4 cd = s_cd[j];
      C0 += cd.x * alpha * T;
      C1 += cd.y * alpha * T;
      C2 += cd.z * alpha * T;
      weight += alpha * T;
      D += cd.w * alpha * T;
      T = test_T;
      last_contributor = contributor;
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

int launch_blend_forward(const EgsForwardArgs& a, const GeomPtrs& g, const uint32_t* point_list,
                         const ImagePtrs& im, cudaStream_t stream) {
  const dim3 grid(tiles_x(a.width), tiles_y(a.height), 1);
  StageTimer timer(stream, ST_BLEND_FWD);
  blend_forward_kernel<<<grid, TILE_PIXELS, 0, stream>>>(
      im.ranges, point_list, a.width, a.height, (int)grid.x, g.means2D, g.conic_opacity,
      g.rgb_depth, a.background, a.out_color, a.out_depth, a.out_alpha, im.n_contrib);
  EGS_AFTER_LAUNCH(stream, a.debug);
  return EGS_OK;
}

}  // namespace egs
