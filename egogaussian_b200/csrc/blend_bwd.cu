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

// ---------------------------------------------------------------------------------------------
// two pixels per thread, exp only for fragments that pass a log-domain test
// ---------------------------------------------------------------------------------------------
// The per-fragment arithmetic of SURVEY.md App. A.8, organised for throughput:
//   * 128 threads per tile, each owning the pixels (x, y) and (x, y + 4) of its warp's 8x8 block:
//     a list entry costs 4 warp reductions per tile instead of 8, every staged entry is read from
//     shared memory half as often, and the two pixel chains give the scheduler independent work;
//   * whether a fragment contributes is decided without the exponential:
//       alpha = o * exp(power) >= 1/255   <=>   power >= -log(255 * o) =: thr (staged per entry),
//     so exp, the divide and everything after run only on warps where some lane contributes.
//     The forward pass decides with min(0.99, o * expf(power)) >= 1/255 in fp32; a fragment whose
//     power lies within CONTRIB_MARGIN of the threshold is therefore decided here with that very
//     expression, so both passes always agree on a pixel's contributor set (one fragment more or
//     less would shift T = T / (1 - alpha) for every earlier fragment of the pixel);
//   * backward is tolerance-checked (1e-4 of the tensor maximum), not bit-checked, so the
//     exponential and the reciprocal use the fast hardware approximations (ex2.approx, rcp.approx).
constexpr int BWD2_THREADS = 128;
constexpr int BWD2_BATCH = 128;
// half-width of the band around -log(255 o) inside which the forward's exact test decides: covers
// the error of __logf (~1e-6 here) and of the fp32 product o * expf(power) a hundred times over
constexpr float CONTRIB_MARGIN = 1e-4f;

// the forward's contribution test, bit for bit (blend_fwd.cu; forward.cu:336-343)
__device__ __forceinline__ bool contributes_exact(float opac, float power) {
  return !(fminf(0.99f, opac * expf(power)) < 1.0f / 255.0f);
}

// single-instruction hardware approximations (MUFU.EX2 / MUFU.RCP, flush-to-zero): the __expf and
// __fdividef intrinsics wrap them in denormal-range fix-ups that cost more than the operation
__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float rcp_approx(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

struct PixelState {
  float T, T_final, accum_rec[3], accum_depth, accum_alpha, last_alpha, last_color[3], last_depth;
  float dL_dpixel[3], dL_ddepth, dL_dalpha, bg_dot;
  float px, py;
  uint32_t last;
};

__device__ __forceinline__ void pixel_init(PixelState& s, bool inside, uint32_t pix_id, unsigned px, unsigned py,
                                           int W, int H, const float* __restrict__ alphas,
                                           const uint32_t* __restrict__ n_contrib,
                                           const float* __restrict__ dL_dpixels,
                                           const float* __restrict__ dL_dpixel_depths,
                                           const float* __restrict__ dL_dalphas, const float* bg) {
  s.px = (float)px;
  s.py = (float)py;
  s.T_final = inside ? (1 - alphas[pix_id]) : 0;
  s.T = s.T_final;
  s.last = inside ? n_contrib[pix_id] : 0;
  const size_t plane = (size_t)H * W;
#pragma unroll
  for (int ch = 0; ch < 3; ++ch) {
    s.dL_dpixel[ch] = inside ? dL_dpixels[ch * plane + pix_id] : 0.f;
    s.accum_rec[ch] = 0.f;
    s.last_color[ch] = 0.f;
  }
  s.dL_ddepth = inside ? dL_dpixel_depths[pix_id] : 0.f;
  s.dL_dalpha = inside ? dL_dalphas[pix_id] : 0.f;
  s.accum_depth = s.accum_alpha = s.last_alpha = s.last_depth = 0.f;
  s.bg_dot = bg[0] * s.dL_dpixel[0] + bg[1] * s.dL_dpixel[1] + bg[2] * s.dL_dpixel[2];
}

// one contributing fragment: advances the pixel's recurrences and adds its ten partial sums to q.
// COLORS_ONLY (EgsBackwardArgs.colors_only): just the transmittance recurrence and the three colour
// sums; nothing that feeds the geometry or opacity gradients is evaluated.
template <bool COLORS_ONLY>
__device__ __forceinline__ void fragment_backward(PixelState& s, float dx, float dy, float power,
                                                  const float4 a4, const float4 b4, const float4 c4,
                                                  float ddelx_dx, float ddely_dy, float (&q)[10]) {
  const float opac = a4.w;
  const float G = ex2_approx(power * 1.4426950408889634f);
  const float alpha = fminf(0.99f, opac * G);
  const float rcp = rcp_approx(1.f - alpha);
  s.T = s.T * rcp;
  const float w = alpha * s.T;
  if (COLORS_ONLY) {
#pragma unroll
    for (int ch = 0; ch < 3; ++ch) q[7 + ch] += w * s.dL_dpixel[ch];
    return;
  }
  float dL_dopa = 0.0f;
  const float c[3] = {c4.x, c4.y, c4.z};
#pragma unroll
  for (int ch = 0; ch < 3; ++ch) {
    s.accum_rec[ch] = s.last_alpha * s.last_color[ch] + (1.f - s.last_alpha) * s.accum_rec[ch];
    s.last_color[ch] = c[ch];
    dL_dopa += (c[ch] - s.accum_rec[ch]) * s.dL_dpixel[ch];
    q[7 + ch] += w * s.dL_dpixel[ch];
  }
  const float c_d = b4.w;
  s.accum_depth = s.last_alpha * s.last_depth + (1.f - s.last_alpha) * s.accum_depth;
  s.last_depth = c_d;
  dL_dopa += (c_d - s.accum_depth) * s.dL_ddepth;
  q[3] += w * s.dL_ddepth;
  s.accum_alpha = s.last_alpha + (1.f - s.last_alpha) * s.accum_alpha;
  dL_dopa += (1 - s.accum_alpha) * s.dL_dalpha;
  dL_dopa *= s.T;
  s.last_alpha = alpha;
  dL_dopa += (-s.T_final * rcp) * s.bg_dot;

  const float dL_dG = opac * dL_dopa;
  const float gdx = G * dx;
  const float gdy = G * dy;
  const float dG_ddelx = -gdx * b4.x - gdy * b4.y;
  const float dG_ddely = -gdy * b4.z - gdx * b4.y;
  q[0] += dL_dG * dG_ddelx * ddelx_dx;
  q[1] += dL_dG * dG_ddely * ddely_dy;
  q[4] += -0.5f * gdx * dx * dL_dG;
  q[5] += -0.5f * gdx * dy * dL_dG;
  q[6] += -0.5f * gdy * dy * dL_dG;
  q[2] += G * dL_dopa;
}

template <bool COLORS_ONLY>
__global__ void __launch_bounds__(BWD2_THREADS)
blend_backward2_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list,
                       int W, int H, int grid_x, const float* __restrict__ bg_color,
                       const float2* __restrict__ means2D, const float4* __restrict__ conic_opacity,
                       const float4* __restrict__ rgb_depth, const float* __restrict__ alphas,
                       const uint32_t* __restrict__ n_contrib, const float* __restrict__ dL_dpixels,
                       const float* __restrict__ dL_dpixel_depths, const float* __restrict__ dL_dalphas,
                       float4* __restrict__ acc) {
  __shared__ uint32_t s_id[BWD2_BATCH];
  __shared__ float4 s_a[BWD2_BATCH];   // mean.x, mean.y, thr = -log(255 o), opacity
  __shared__ float4 s_b[BWD2_BATCH];   // conic.x, conic.y, conic.z, depth
  __shared__ float4 s_c[BWD2_BATCH];   // r, g, b, -
  // one accumulator row per warp: a warp meets each entry of a batch once, so its total is a plain
  // store (shared-memory float atomicAdd is a compare-and-swap loop on this architecture)
  __shared__ float s_acc[(BWD2_THREADS / 32) * BWD2_BATCH * ACC_PITCH];
  __shared__ uint32_t s_hit[BWD2_BATCH];   // bit w: warp w stored totals for this entry in this batch
  // [pixel warp][32-entry chunk]: entries of the batch that can reach the warp's 8x8 pixels. A warp
  // pops set bits in list order and never issues an instruction for the other entries.
  __shared__ uint32_t s_bits[BWD2_THREADS / 32][BWD2_BATCH / 32];
  __shared__ uint32_t s_max_contrib;

  const unsigned tid = threadIdx.x;
  const unsigned warp = tid >> 5, lane = tid & 31;
  // 4 warps as 2x2 blocks of 8x8 pixels; a lane owns (lx, ly) and (lx, ly + 4) of its block
  const unsigned bx = blockIdx.x * TILE_X + (warp & 1) * 8 + (lane & 7);
  const unsigned by = blockIdx.y * TILE_Y + (warp >> 1) * 8 + (lane >> 3);
  const float bg[3] = {bg_color[0], bg_color[1], bg_color[2]};
  PixelState P0, P1;
  {
    const bool in0 = bx < (unsigned)W && by < (unsigned)H;
    const bool in1 = bx < (unsigned)W && (by + 4) < (unsigned)H;
    pixel_init(P0, in0, (uint32_t)W * by + bx, bx, by, W, H, alphas, n_contrib, dL_dpixels, dL_dpixel_depths,
               dL_dalphas, bg);
    pixel_init(P1, in1, (uint32_t)W * (by + 4) + bx, bx, by + 4, W, H, alphas, n_contrib, dL_dpixels,
               dL_dpixel_depths, dL_dalphas, bg);
  }
  const uint2 range = ranges[blockIdx.y * grid_x + blockIdx.x];

  if (tid == 0) s_max_contrib = 0;
  __syncthreads();
  uint32_t warp_last = max(P0.last, P1.last);
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) warp_last = max(warp_last, __shfl_xor_sync(0xffffffffu, warp_last, d));
  if (lane == 0 && warp_last) atomicMax(&s_max_contrib, warp_last);
  __syncthreads();
  const int walk = (int)s_max_contrib;
  if (walk == 0) return;

  const float ddelx_dx = 0.5 * W;
  const float ddely_dy = 0.5 * H;
  const int my_quantity = quantity_of_lane(lane);
  const int my_slot = my_quantity >= 0 ? record_slot_of_quantity(my_quantity) : 0;
  const bool i_add = my_quantity >= 0 && (lane & 1) == 0;

  // shared-window addresses formed once (see lds_f4 in common.cuh)
  uint32_t a_a = smem_addr(s_a), a_b = smem_addr(s_b), a_c = smem_addr(s_c);
  uint32_t a_bits = smem_addr(&s_bits[warp][0]);
  uint32_t a_acc = smem_addr(&s_acc[warp * BWD2_BATCH * ACC_PITCH + my_slot]);   // this lane's slot, entry 0
  // opaque to the optimiser, or it re-derives the window addresses inside the loop
  asm volatile("" : "+r"(a_a), "+r"(a_b), "+r"(a_c), "+r"(a_bits), "+r"(a_acc));

  const int rounds = (walk + BWD2_BATCH - 1) / BWD2_BATCH;
  for (int r = 0; r < rounds; ++r) {
    __syncthreads();
    const int pos = walk - 1 - (r * BWD2_BATCH + (int)tid);
    unsigned reach = 0u;
    if (pos >= 0) {
      const uint32_t id = point_list[range.x + pos];
      const float2 xy = means2D[id];
      const float4 co = conic_opacity[id];
      const float4 cd = rgb_depth[id];
      s_id[tid] = id;
      // alpha >= 1/255  <=>  power >= -log(255 o), staged with the margin subtracted; o <= 0 can
      // never contribute
      s_a[tid] = make_float4(xy.x, xy.y, co.w > 0.f ? -__logf(255.f * co.w) - CONTRIB_MARGIN : 1.f, co.w);
      reach = footprint_mask<8, 8>(xy, co, (float)(blockIdx.x * TILE_X), (float)(blockIdx.y * TILE_Y));
      s_b[tid] = make_float4(co.x, co.y, co.z, cd.w);
      s_c[tid] = cd;
    }
    s_hit[tid] = 0u;
    {
      uint32_t mine = 0u;   // lane w ends up with pixel warp w's word for this staging warp's chunk
#pragma unroll
      for (unsigned w = 0; w < BWD2_THREADS / 32; ++w) {
        const uint32_t word = __ballot_sync(0xffffffffu, (reach >> w) & 1u);
        if (lane == w) mine = word;
      }
      if (lane < BWD2_THREADS / 32) s_bits[lane][warp] = mine;
    }
    __syncthreads();

    const int n = min(BWD2_BATCH, walk - r * BWD2_BATCH);
    // entries j < jmin lie behind the last contributor of every pixel of this warp
    const int jmin = walk - r * BWD2_BATCH - (int)warp_last;
    for (int c = 0; c < BWD2_BATCH / 32; ++c) {
     uint32_t bits = lds_u32(a_bits + 4u * c);
     const int lo = jmin - 32 * c;
     if (lo >= 32) continue;
     if (lo > 0) bits &= 0xffffffffu << lo;
     while (bits != 0u) {   // warp-uniform: every lane holds the same word
      const int j = c * 32 + __ffs((int)bits) - 1;
      bits &= bits - 1u;
      const uint32_t jpos = (uint32_t)(walk - 1 - (r * BWD2_BATCH + j));
      const float4 a4 = lds_f4(a_a + 16u * j);
      const float4 b4 = lds_f4(a_b + 16u * j);
      const float dx0 = a4.x - P0.px, dy0 = a4.y - P0.py;
      const float dx1 = a4.x - P1.px, dy1 = a4.y - P1.py;
      const float power0 = -0.5f * (b4.x * dx0 * dx0 + b4.z * dy0 * dy0) - b4.y * dx0 * dy0;
      const float power1 = -0.5f * (b4.x * dx1 * dx1 + b4.z * dy1 * dy1) - b4.y * dx1 * dy1;
      bool act0 = jpos < P0.last && power0 <= 0.0f && power0 >= a4.z;
      bool act1 = jpos < P1.last && power1 <= 0.0f && power1 >= a4.z;
      if (!__any_sync(0xffffffffu, act0 || act1)) continue;
      // inside the band around the threshold the forward's own expression decides
      const float band_hi = a4.z + 2.f * CONTRIB_MARGIN;
      if (act0 && power0 < band_hi) act0 = contributes_exact(a4.w, power0);
      if (act1 && power1 < band_hi) act1 = contributes_exact(a4.w, power1);

      float q[10];
#pragma unroll
      for (int k = 0; k < 10; ++k) q[k] = 0.f;
      const float4 c4 = lds_f4(a_c + 16u * j);
      if (act0) fragment_backward<COLORS_ONLY>(P0, dx0, dy0, power0, a4, b4, c4, ddelx_dx, ddely_dy, q);
      if (act1) fragment_backward<COLORS_ONLY>(P1, dx1, dy1, power1, a4, b4, c4, ddelx_dx, ddely_dy, q);
      const float total = warp_packed_reduce10(q, lane);
      if (i_add) sts_f32(a_acc + (uint32_t)(j * ACC_PITCH * 4), total);
      if (lane == 0) atomicOr(&s_hit[j], 1u << warp);
     }
    }
    __syncthreads();

    if ((int)tid < n) {
      float v[ACC_STRIDE];
#pragma unroll
      for (int k = 0; k < ACC_STRIDE; ++k) v[k] = 0.f;
      const uint32_t hit = s_hit[tid];
#pragma unroll
      for (int w = 0; w < BWD2_THREADS / 32; ++w) {
        if (hit & (1u << w)) {
          const float* row = &s_acc[(w * BWD2_BATCH + tid) * ACC_PITCH];
#pragma unroll
          for (int k = 0; k < ACC_STRIDE; ++k)
            if (k != 7 && k != 11) v[k] += row[k];   // slots 7 and 11 are padding, never written
        }
      }
      if (hit) {
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
  if (a.colors_only) {
    blend_backward2_kernel<true><<<grid, BWD2_THREADS, 0, stream>>>(
        im.ranges, point_list, a.width, a.height, (int)grid.x, a.background, g.means2D,
        g.conic_opacity, g.rgb_depth, a.out_alpha, im.n_contrib, a.dL_dout_color, a.dL_dout_depth,
        a.dL_dout_alpha, s.acc);
  } else {
    blend_backward2_kernel<false><<<grid, BWD2_THREADS, 0, stream>>>(
        im.ranges, point_list, a.width, a.height, (int)grid.x, a.background, g.means2D,
        g.conic_opacity, g.rgb_depth, a.out_alpha, im.n_contrib, a.dL_dout_color, a.dL_dout_depth,
        a.dL_dout_alpha, s.acc);
  }
  EGS_AFTER_LAUNCH(stream, a.debug);
  return EGS_OK;
}

}  // namespace egs
