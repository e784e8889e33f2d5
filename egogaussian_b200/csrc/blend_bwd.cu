// blend_bwd.cu -- per-tile back-to-front re-walk of the blend: dL/d{mean2D, conic, opacity,
// colour, depth} accumulated per Gaussian.
//
// Per-pixel arithmetic follows the reference's backward renderCUDA
// (DGR/cuda_rasterizer/backward.cu:415-601) as restated in SURVEY.md App. A.8, including its
// non-autograd conventions: T_final = 1 - alpha image, the gradient flows through the 0.99
// clamp, and the background term uses T_final / (1 - alpha).
//
// Mechanism (ours): the reference issues 10 scalar atomicAdd per blended fragment. Here
//   * the walk starts at the tile's largest contributor count instead of the list end (entries
//     behind every pixel's last contributor are skipped without being fetched);
//   * all 32 lanes of a warp look at the same list entry, so the ten partial sums are first
//     reduced across the warp (a packed butterfly: 12 shuffles for 10 quantities) and only if
//     some lane contributed at all;
//   * the warp totals are added into a per-batch shared-memory accumulator (12 floats per list
//     entry), and after the batch each thread flushes one entry with three 16-byte vector
//     reductions (red.global.add.v4.f32) into the 48-byte per-Gaussian record.
#include "common.cuh"

namespace egs {

constexpr int BWD_BATCH = 256;
constexpr int ACC_STRIDE = 12;
constexpr int ACC_PITCH = 13;  // odd pitch: conflict-free both for the per-entry atomics and the flush

// Sum ten per-lane quantities over the warp with a packed butterfly. On return lane
// (2*slot) and (2*slot+1) hold the total of the quantity mapped to `slot` (see slot_of_quantity).
// Layout of the halving: 10 -> 5 (xor 16) -> 3 (xor 8) -> 2 (xor 4) -> 1 (xor 2) -> sum (xor 1).
__device__ __forceinline__ float warp_packed_reduce10(const float (&q)[10], unsigned lane) {
  const bool b4 = lane & 16, b3 = lane & 8, b2 = lane & 4, b1 = lane & 2;
  float a[5];
#pragma unroll
  for (int k = 0; k < 5; ++k) {
    const float send = b4 ? q[k] : q[k + 5];
    const float keep = b4 ? q[k + 5] : q[k];
    a[k] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
  }
  // a[0..4] (+ implicit zero a[5]); b3 = 0 keeps a[0..2], b3 = 1 keeps a[3], a[4], 0
  float b[3];
  {
    const float s0 = b3 ? a[0] : a[3], k0 = b3 ? a[3] : a[0];
    const float s1 = b3 ? a[1] : a[4], k1 = b3 ? a[4] : a[1];
    const float s2 = b3 ? a[2] : 0.f, k2 = b3 ? 0.f : a[2];
    b[0] = k0 + __shfl_xor_sync(0xffffffffu, s0, 8);
    b[1] = k1 + __shfl_xor_sync(0xffffffffu, s1, 8);
    b[2] = k2 + __shfl_xor_sync(0xffffffffu, s2, 8);
  }
  // b[0..2] (+ zero b[3]); b2 = 0 keeps b[0], b[1]; b2 = 1 keeps b[2], 0
  float c[2];
  {
    const float s0 = b2 ? b[0] : b[2], k0 = b2 ? b[2] : b[0];
    const float s1 = b2 ? b[1] : 0.f, k1 = b2 ? 0.f : b[1];
    c[0] = k0 + __shfl_xor_sync(0xffffffffu, s0, 4);
    c[1] = k1 + __shfl_xor_sync(0xffffffffu, s1, 4);
  }
  const float s = b1 ? c[0] : c[1], k = b1 ? c[1] : c[0];
  float r = k + __shfl_xor_sync(0xffffffffu, s, 2);
  r += __shfl_xor_sync(0xffffffffu, r, 1);
  return r;
}

// Which of the ten quantities a lane ends up holding after warp_packed_reduce10; -1 = padding.
__device__ __forceinline__ int quantity_of_lane(unsigned lane) {
  const int b4 = (lane >> 4) & 1, b3 = (lane >> 3) & 1, b2 = (lane >> 2) & 1, b1 = (lane >> 1) & 1;
  // within the group of five: index g5 in 0..5 (5 is padding)
  const int g3 = b2 * 2 + b1;           // 0..3 (3 is padding)
  if (g3 == 3) return -1;
  const int g5 = b3 * 3 + g3;           // 0..5
  if (g5 >= 5) return -1;
  return b4 * 5 + g5;
}

// Quantity order used by the tile kernel, and where each lands in the 12-float record:
//   q0 mean2D.x -> 0   q1 mean2D.y -> 1   q2 opacity -> 2   q3 depth -> 3
//   q4 conic.x  -> 4   q5 conic.y  -> 5   q6 conic.w -> 6
//   q7 colour.r -> 8   q8 colour.g -> 9   q9 colour.b -> 10
__device__ __forceinline__ int record_slot_of_quantity(int q) { return q < 7 ? q : q + 1; }

__global__ void __launch_bounds__(TILE_PIXELS)
blend_backward_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list,
                      int W, int H, int grid_x, const float* __restrict__ bg_color,
                      const float2* __restrict__ means2D, const float4* __restrict__ conic_opacity,
                      const float4* __restrict__ rgb_depth, const float* __restrict__ alphas,
                      const uint32_t* __restrict__ n_contrib, const float* __restrict__ dL_dpixels,
                      const float* __restrict__ dL_dpixel_depths, const float* __restrict__ dL_dalphas,
                      float4* __restrict__ acc) {
  __shared__ uint32_t s_id[BWD_BATCH];
  __shared__ float2 s_xy[BWD_BATCH];
  __shared__ float4 s_co[BWD_BATCH];
  __shared__ float4 s_cd[BWD_BATCH];
  __shared__ float s_acc[BWD_BATCH * ACC_PITCH];
  __shared__ uint32_t s_max_contrib;

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

  const float T_final = inside ? (1 - alphas[pix_id]) : 0;
  float T = T_final;
  const uint32_t last_contributor = inside ? n_contrib[pix_id] : 0;

  if (tid == 0) s_max_contrib = 0;
  __syncthreads();
  {
    uint32_t m = last_contributor;
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, d));
    if (lane == 0 && m) atomicMax(&s_max_contrib, m);
  }
  __syncthreads();
  const int walk = (int)s_max_contrib;  // positions [0, walk) of this tile's list matter
  if (walk == 0) return;
  uint32_t warp_last = last_contributor;
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) warp_last = max(warp_last, __shfl_xor_sync(0xffffffffu, warp_last, d));

  float accum_rec[3] = {0.f, 0.f, 0.f};
  float dL_dpixel[3] = {0.f, 0.f, 0.f};
  float accum_depth_rec = 0.f, dL_dpixel_depth = 0.f;
  float accum_alpha_rec = 0.f, dL_dalpha = 0.f;
  if (inside) {
    const size_t plane = (size_t)H * W;
#pragma unroll
    for (int ch = 0; ch < 3; ++ch) dL_dpixel[ch] = dL_dpixels[ch * plane + pix_id];
    dL_dpixel_depth = dL_dpixel_depths[pix_id];
    dL_dalpha = dL_dalphas[pix_id];
  }
  float last_alpha = 0.f;
  float last_color[3] = {0.f, 0.f, 0.f};
  float last_depth = 0.f;
  const float ddelx_dx = 0.5 * W;
  const float ddely_dy = 0.5 * H;
  float bg_dot_dpixel = 0;
#pragma unroll
  for (int ch = 0; ch < 3; ++ch) bg_dot_dpixel += bg_color[ch] * dL_dpixel[ch];

  const int my_quantity = quantity_of_lane(lane);
  const int my_slot = my_quantity >= 0 ? record_slot_of_quantity(my_quantity) : 0;
  const bool i_add = my_quantity >= 0 && (lane & 1) == 0;

  const int rounds = (walk + BWD_BATCH - 1) / BWD_BATCH;
  for (int r = 0; r < rounds; ++r) {
    __syncthreads();
    // stage positions walk-1-(r*B + tid), i.e. back to front
    const int pos = walk - 1 - (r * BWD_BATCH + (int)tid);
    if (pos >= 0) {
      const uint32_t id = point_list[range.x + pos];
      s_id[tid] = id;
      s_xy[tid] = means2D[id];
      s_co[tid] = conic_opacity[id];
      s_cd[tid] = rgb_depth[id];
    }
#pragma unroll
    for (int k = 0; k < ACC_PITCH; ++k) s_acc[k * BWD_BATCH + tid] = 0.f;
    __syncthreads();

    const int n = min(BWD_BATCH, walk - r * BWD_BATCH);
    for (int j = 0; j < n; ++j) {
      const int jpos = walk - 1 - (r * BWD_BATCH + j);  // 0-based list position of this entry
      if ((uint32_t)jpos >= warp_last) continue;        // warp-uniform
      bool active = (uint32_t)jpos < last_contributor;
      const float2 xy = s_xy[j];
      const float2 d = make_float2(xy.x - pixf.x, xy.y - pixf.y);
      const float4 con_o = s_co[j];
      const float power = -0.5f * (con_o.x * d.x * d.x + con_o.z * d.y * d.y) - con_o.y * d.x * d.y;
      active = active && !(power > 0.0f);
      const float G = expf(power);
      const float alpha = min(0.99f, con_o.w * G);
      active = active && !(alpha < 1.0f / 255.0f);
      if (!__any_sync(0xffffffffu, active)) continue;

      float q[10];
#pragma unroll
      for (int k = 0; k < 10; ++k) q[k] = 0.f;
      if (active) {
        T = T / (1.f - alpha);
        const float dchannel_dcolor = alpha * T;
        const float4 cd = s_cd[j];
        const float c[3] = {cd.x, cd.y, cd.z};
        float dL_dopa = 0.0f;
#pragma unroll
        for (int ch = 0; ch < 3; ++ch) {
          accum_rec[ch] = last_alpha * last_color[ch] + (1.f - last_alpha) * accum_rec[ch];
          last_color[ch] = c[ch];
          const float dL_dchannel = dL_dpixel[ch];
          dL_dopa += (c[ch] - accum_rec[ch]) * dL_dchannel;
          q[7 + ch] = dchannel_dcolor * dL_dchannel;
        }
        const float c_d = cd.w;
        accum_depth_rec = last_alpha * last_depth + (1.f - last_alpha) * accum_depth_rec;
        last_depth = c_d;
        dL_dopa += (c_d - accum_depth_rec) * dL_dpixel_depth;
        q[3] = dchannel_dcolor * dL_dpixel_depth;

        accum_alpha_rec = last_alpha + (1.f - last_alpha) * accum_alpha_rec;
        dL_dopa += (1 - accum_alpha_rec) * dL_dalpha;

        dL_dopa *= T;
        last_alpha = alpha;
        dL_dopa += (-T_final / (1.f - alpha)) * bg_dot_dpixel;

        const float dL_dG = con_o.w * dL_dopa;
        const float gdx = G * d.x;
        const float gdy = G * d.y;
        const float dG_ddelx = -gdx * con_o.x - gdy * con_o.y;
        const float dG_ddely = -gdy * con_o.z - gdx * con_o.y;
        q[0] = dL_dG * dG_ddelx * ddelx_dx;
        q[1] = dL_dG * dG_ddely * ddely_dy;
        q[4] = -0.5f * gdx * d.x * dL_dG;
        q[5] = -0.5f * gdx * d.y * dL_dG;
        q[6] = -0.5f * gdy * d.y * dL_dG;
        q[2] = G * dL_dopa;
      }
      const float total = warp_packed_reduce10(q, lane);
      if (i_add) atomicAdd(&s_acc[j * ACC_PITCH + my_slot], total);
    }
    __syncthreads();

    // flush: thread tid owns entry tid of the batch
    if ((int)tid < n) {
      float v[ACC_STRIDE];
      bool any = false;
#pragma unroll
      for (int k = 0; k < ACC_STRIDE; ++k) {
        v[k] = s_acc[tid * ACC_PITCH + k];
        any = any || (v[k] != 0.f);
      }
      if (any) {
        float4* dst = acc + (size_t)s_id[tid] * 3;
        atomicAdd(dst + 0, make_float4(v[0], v[1], v[2], v[3]));
        atomicAdd(dst + 1, make_float4(v[4], v[5], v[6], v[7]));
        atomicAdd(dst + 2, make_float4(v[8], v[9], v[10], v[11]));
      }
    }
  }
}

int launch_blend_backward(const EgsBackwardArgs& a, const GeomPtrs& g, const uint32_t* point_list,
                          const ImagePtrs& im, const BackwardScratch& s, cudaStream_t stream) {
  const dim3 grid(tiles_x(a.width), tiles_y(a.height), 1);
  StageTimer timer(stream, ST_BLEND_BWD);
  blend_backward_kernel<<<grid, TILE_PIXELS, 0, stream>>>(
      im.ranges, point_list, a.width, a.height, (int)grid.x, a.background, g.means2D,
      g.conic_opacity, g.rgb_depth, a.out_alpha, im.n_contrib, a.dL_dout_color, a.dL_dout_depth,
      a.dL_dout_alpha, s.acc);
  EGS_AFTER_LAUNCH(stream, a.debug);
  return EGS_OK;
}

}  // namespace egs
