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
//   * the warp totals are stored into a per-warp shared-memory row (12 floats per list entry), and
//     after the batch a thread sums an entry's rows and flushes it with three 16-byte vector
//     reductions (red.global.add.v4.f32) into the 48-byte per-Gaussian record.
#include "common.cuh"

namespace egs {

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
// several pixels per thread, exp only for fragments that pass a log-domain test
// ---------------------------------------------------------------------------------------------
// The per-fragment arithmetic of SURVEY.md App. A.8, organised for throughput (the kernel is bound by
// instruction issue, so everything here is about instructions per fragment):
//   * a warp covers an 8 x (4 PPL) block of the tile, a lane owning PPL pixels of one column (rows y,
//     y + 4, ...): one warp reduction per list entry serves 32 PPL pixels, every staged entry is read
//     from shared memory once per PPL pixels, and the PPL pixel chains give the scheduler independent
//     work. PPL = 4: 64 threads per tile;
//   * whether a fragment contributes is decided without the exponential:
//       alpha = o * exp(power) >= 1/255   <=>   power >= -log(255 * o) =: thr (staged per entry),
//     so exp, the divide and everything after run only on warps where some lane contributes.
//     The forward pass decides with min(0.99, o * expf(power)) >= 1/255 in fp32; a fragment whose
//     power lies within CONTRIB_MARGIN of the threshold is therefore decided here with that very
//     expression, so both passes always agree on a pixel's contributor set (one fragment more or
//     less would shift T = T / (1 - alpha) for every earlier fragment of the pixel);
//   * what lies behind a fragment enters dL/dalpha only through its dot product with the pixel's upstream
//     gradients, so ONE running number per pixel replaces the reference's five accumulated values
//     (PixelState::S): seven instead of fifteen instructions per fragment and sixteen registers per lane
//     less (0.335 -> 0.307 ms at C3); it is advanced right after a fragment has used it;
//   * the geometry gradients are accumulated as moments of v = G dL/dG over the pixels -- sum v, v dx,
//     v dy, v dx^2, v dx dy, v dy^2 -- and turned into dL/dmean2D and dL/dconic once per entry after the
//     reduction (9 instead of 17 instructions per fragment);
//   * backward is tolerance-checked (1e-4 of the tensor maximum), not bit-checked, so the
//     exponential and the reciprocal use the fast hardware approximations (ex2.approx, rcp.approx).
constexpr int BWD_BATCH = 128;
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

// What a pixel carries through the walk: transmittance, its upstream gradients, tb = -T_final *
// (background . dL/dcolour), and S = (what lies behind the current fragment) . (upstream gradients):
// the reference keeps the colour, depth and alpha accumulated behind a fragment (five numbers) only to
// form sum_c (c_c - behind_c) g_c, which is (c . g) - S, and behind' = behind + alpha (c - behind) dotted
// with g is S' = S + alpha ((c . g) - S): one number per pixel does it.
struct PixelState {
  float T, S;
  float gr, gg, gb, gd, ga, tb;
  uint32_t last;
};

__device__ __forceinline__ void pixel_init(PixelState& s, bool inside, uint32_t pix_id, int W, int H,
                                           const float* __restrict__ alphas, const uint32_t* __restrict__ n_contrib,
                                           const float* __restrict__ dL_dpixels,
                                           const float* __restrict__ dL_dpixel_depths,
                                           const float* __restrict__ dL_dalphas, const float* bg) {
  const float T_final = inside ? (1 - alphas[pix_id]) : 0;   // backward.cu:463
  s.T = T_final;
  s.last = inside ? n_contrib[pix_id] : 0;
  const size_t plane = (size_t)H * W;
  s.gr = inside ? dL_dpixels[pix_id] : 0.f;
  s.gg = inside ? dL_dpixels[plane + pix_id] : 0.f;
  s.gb = inside ? dL_dpixels[2 * plane + pix_id] : 0.f;
  s.gd = inside && dL_dpixel_depths != nullptr ? dL_dpixel_depths[pix_id] : 0.f;   // absent = zero
  s.ga = inside && dL_dalphas != nullptr ? dL_dalphas[pix_id] : 0.f;
  s.S = 0.f;
  s.tb = -T_final * (bg[0] * s.gr + bg[1] * s.gg + bg[2] * s.gb);
}

// Quantity order of the reduction and of the per-warp accumulator rows, with u = G dL/dalpha:
//   q0 sum u dx   q1 sum u dy   q2 sum u (= dL/dopacity)   q3 depth
//   q4 sum u dx^2   q5 sum u dx dy   q6 sum u dy^2   q7..q9 colour
// One contributing fragment: advances the pixel's recurrences and adds its partial sums to q.
// COLORS_ONLY (EgsBackwardArgs.colors_only): just the transmittance recurrence and the three colour
// sums; nothing that feeds the geometry or opacity gradients is evaluated.
// DA = false: no upstream gradient for the depth and alpha images (a loss on the colour image only, the
// trainers' usual case): their recurrences and their terms of dL/dalpha are left out.
template <bool COLORS_ONLY, bool DA>
__device__ __forceinline__ void fragment_backward(PixelState& s, float dx, float dy, float power, float opac,
                                                  const float4 c4, float depth, float (&q)[10]) {
  const float G = ex2_approx(power * 1.4426950408889634f);
  const float alpha = fminf(0.99f, opac * G);
  const float rcp = rcp_approx(1.f - alpha);
  s.T = s.T * rcp;
  const float w = alpha * s.T;
  q[7] += w * s.gr;
  q[8] += w * s.gg;
  q[9] += w * s.gb;
  if (COLORS_ONLY) return;
  if (DA) q[3] += w * s.gd;
  // (this fragment's value) - (what lies behind it), dotted with the upstream gradients: feeds dL/dalpha,
  // and advances S to the next fragment of this pixel (one closer to the eye)
  float X = c4.x * s.gr;
  X = fmaf(c4.y, s.gg, X);
  X = fmaf(c4.z, s.gb, X);
  if (DA) {
    X = fmaf(depth, s.gd, X);
    X += s.ga;
  }
  float dL_dopa = X - s.S;
  s.S = fmaf(alpha, dL_dopa, s.S);
  dL_dopa = dL_dopa * s.T + s.tb * rcp;
  const float u = G * dL_dopa;          // the opacity factor of dL/dG = o dL/dalpha is applied once per entry
  const float ux = u * dx, uy = u * dy;
  q[2] += u;
  q[0] += ux;
  q[1] += uy;
  q[4] += ux * dx;
  q[5] += ux * dy;
  q[6] += uy * dy;
}

// (PPL = 4 takes 96 registers, 10 CTAs of 64 threads per SM, since the "behind" state is one number per
// pixel; with the reference's five it took 118 and 8 CTAs, and forcing that form into 96 measured slower)
template <int PPL, bool COLORS_ONLY, bool DA>
__global__ void __launch_bounds__(256 / PPL)
blend_backward_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list,
                      int W, int H, int grid_x, const float* __restrict__ bg_color,
                      const float2* __restrict__ means2D, const float4* __restrict__ conic_opacity,
                      const float4* __restrict__ rgb_depth, const float* __restrict__ alphas,
                      const uint32_t* __restrict__ n_contrib, const float* __restrict__ dL_dpixels,
                      const float* __restrict__ dL_dpixel_depths, const float* __restrict__ dL_dalphas,
                      float4* __restrict__ acc) {
  constexpr int THREADS = 256 / PPL;
  constexpr int WARPS = THREADS / 32;
  constexpr int STAGE = BWD_BATCH / THREADS;     // entries a thread stages per batch
  constexpr int BLOCK_H = 4 * PPL;               // a warp's pixels: 8 wide, BLOCK_H tall
  static_assert(STAGE >= 1 && STAGE * THREADS == BWD_BATCH, "batch");
  __shared__ uint32_t s_id[BWD_BATCH];
  __shared__ float4 s_a[BWD_BATCH];   // mean.x, mean.y, thr = -log(255 o) - margin, opacity
  __shared__ float4 s_b[BWD_BATCH];   // conic.x, conic.y, conic.z, depth
  __shared__ float4 s_c[BWD_BATCH];   // r, g, b, -
  // one accumulator row per warp: a warp meets each entry of a batch once, so its total is a plain
  // store (shared-memory float atomicAdd is a compare-and-swap loop on this architecture)
  __shared__ float s_acc[WARPS * BWD_BATCH * ACC_PITCH];
  __shared__ unsigned char s_hit[WARPS][BWD_BATCH];   // warp w stored totals for this entry in this batch
  // [pixel warp][32-entry chunk]: entries of the batch that can reach the warp's pixels. A warp
  // pops set bits in list order and never issues an instruction for the other entries.
  __shared__ uint32_t s_bits[WARPS][BWD_BATCH / 32];
  __shared__ uint32_t s_max_contrib;

  const unsigned tid = threadIdx.x;
  const unsigned warp = tid >> 5, lane = tid & 31;
  // warps as 2 x (WARPS / 2) blocks of 8 x BLOCK_H pixels; a lane owns (x, y + 4 s), s < PPL
  const unsigned bx = blockIdx.x * TILE_X + (warp & 1) * 8 + (lane & 7);
  const unsigned by = blockIdx.y * TILE_Y + (warp >> 1) * BLOCK_H + (lane >> 3);
  const float bg[3] = {bg_color[0], bg_color[1], bg_color[2]};
  const float pxf = (float)bx, pyf = (float)by;
  PixelState px[PPL];
#pragma unroll
  for (int s = 0; s < PPL; ++s) {
    const unsigned y = by + 4 * s;
    pixel_init(px[s], bx < (unsigned)W && y < (unsigned)H, (uint32_t)W * y + bx, W, H, alphas, n_contrib, dL_dpixels,
               dL_dpixel_depths, dL_dalphas, bg);
  }
  const uint2 range = ranges[blockIdx.y * grid_x + blockIdx.x];

  if (tid == 0) s_max_contrib = 0;
  __syncthreads();
  uint32_t warp_last = 0;
#pragma unroll
  for (int s = 0; s < PPL; ++s) warp_last = max(warp_last, px[s].last);
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
  uint32_t a_hit = smem_addr(&s_hit[warp][0]);
  uint32_t a_acc = smem_addr(&s_acc[warp * BWD_BATCH * ACC_PITCH + my_slot]);   // this lane's slot, entry 0
  // opaque to the optimiser, or it re-derives the window addresses inside the loop
  asm volatile("" : "+r"(a_a), "+r"(a_b), "+r"(a_c), "+r"(a_bits), "+r"(a_acc), "+r"(a_hit));

  const int rounds = (walk + BWD_BATCH - 1) / BWD_BATCH;
  for (int r = 0; r < rounds; ++r) {
    __syncthreads();
#pragma unroll
    for (int h = 0; h < STAGE; ++h) {
      const int slot = h * THREADS + (int)tid;      // staged back to front; its 32-entry chunk is h * WARPS + warp
      const int pos = walk - 1 - (r * BWD_BATCH + slot);
      unsigned reach = 0u;
      if (pos >= 0) {
        const uint32_t id = point_list[range.x + pos];
        const float2 xy = means2D[id];
        const float4 co = conic_opacity[id];
        const float4 cd = rgb_depth[id];
        s_id[slot] = id;
        // alpha >= 1/255  <=>  power >= -log(255 o), staged with the margin subtracted; o <= 0 can
        // never contribute
        s_a[slot] = make_float4(xy.x, xy.y, co.w > 0.f ? -__logf(255.f * co.w) - CONTRIB_MARGIN : 1.f, co.w);
        reach = footprint_mask<8, BLOCK_H>(xy, co, (float)(blockIdx.x * TILE_X), (float)(blockIdx.y * TILE_Y));
        s_b[slot] = make_float4(co.x, co.y, co.z, cd.w);
        s_c[slot] = cd;
      }
#pragma unroll
      for (unsigned w = 0; w < WARPS; ++w) s_hit[w][slot] = 0;
      uint32_t mine = 0u;   // lane w ends up with pixel warp w's word for this staging warp's chunk
#pragma unroll
      for (unsigned w = 0; w < WARPS; ++w) {
        const uint32_t word = __ballot_sync(0xffffffffu, (reach >> w) & 1u);
        if (lane == w) mine = word;
      }
      if (lane < WARPS) s_bits[lane][h * WARPS + warp] = mine;
    }
    __syncthreads();

    const int n = min(BWD_BATCH, walk - r * BWD_BATCH);
    // entries j < jmin lie behind the last contributor of every pixel of this warp
    const int jmin = walk - r * BWD_BATCH - (int)warp_last;
    for (int c = 0; c < BWD_BATCH / 32; ++c) {
     uint32_t bits = lds_u32(a_bits + 4u * c);
     const int lo = jmin - 32 * c;
     if (lo >= 32) continue;
     if (lo > 0) bits &= 0xffffffffu << lo;
     while (bits != 0u) {   // warp-uniform: every lane holds the same word
      const int j = c * 32 + __ffs((int)bits) - 1;
      bits &= bits - 1u;
      const uint32_t jpos = (uint32_t)(walk - 1 - (r * BWD_BATCH + j));
      const float4 a4 = lds_f4(a_a + 16u * j);
      const float4 b4 = lds_f4(a_b + 16u * j);
      const float dx = a4.x - pxf;
      const float ax = __fmul_rn(b4.x, dx), bxm = __fmul_rn(b4.y, dx);   // shared by the lane's pixels (one column)
      float dy[PPL], power[PPL];
      bool act[PPL];
      bool any = false;
#pragma unroll
      for (int s = 0; s < PPL; ++s) {
        dy[s] = a4.y - (pyf + (float)(4 * s));
        power[s] = splat_power_pre(ax, bxm, b4.z, dx, dy[s]);   // the forward's bits
        act[s] = jpos < px[s].last && power[s] <= 0.0f && power[s] >= a4.z;
        any = any || act[s];
      }
      if (!__any_sync(0xffffffffu, any)) continue;
      float q[10];
#pragma unroll
      for (int k = 0; k < 10; ++k) q[k] = 0.f;
      const float4 c4 = lds_f4(a_c + 16u * j);
      const float band_hi = a4.z + 2.f * CONTRIB_MARGIN;
#pragma unroll
      for (int s = 0; s < PPL; ++s) {
        if (act[s]) {
          // inside the band around the threshold the forward's own expression decides
          if (power[s] >= band_hi || contributes_exact(a4.w, power[s]))
            fragment_backward<COLORS_ONLY, DA>(px[s], dx, dy[s], power[s], a4.w, c4, b4.w, q);
        }
      }
      const float total = warp_packed_reduce10(q, lane);
      if (i_add) sts_f32(a_acc + (uint32_t)(j * ACC_PITCH * 4), total);
      if (lane == 0) asm volatile("st.shared.u8 [%0], %1;" ::"r"(a_hit + (uint32_t)j), "r"(1u) : "memory");
     }
    }
    __syncthreads();

    // flush: a thread owns STAGE entries of the batch; moments -> dL/dmean2D, dL/dconic
#pragma unroll
    for (int h = 0; h < STAGE; ++h) {
      const int e = h * THREADS + (int)tid;
      if (e >= n) continue;
      uint32_t hit = 0u;
#pragma unroll
      for (int w = 0; w < WARPS; ++w) hit |= (uint32_t)s_hit[w][e] << w;
      if (!hit) continue;
      float v[ACC_STRIDE];
#pragma unroll
      for (int k = 0; k < ACC_STRIDE; ++k) v[k] = 0.f;
#pragma unroll
      for (int w = 0; w < WARPS; ++w) {
        if (hit & (1u << w)) {
          const float* row = &s_acc[(w * BWD_BATCH + e) * ACC_PITCH];
#pragma unroll
          for (int k = 0; k < ACC_STRIDE; ++k)
            if (k != 7 && k != 11) v[k] += row[k];   // slots 7 and 11 are padding, never written
        }
      }
      float4* dst = acc + (size_t)s_id[e] * 3;
      if (!COLORS_ONLY) {
        const float4 b4 = s_b[e];
        const float o = s_a[e].w;
        const float sx = o * v[0], sy = o * v[1];   // dL/dG = o dL/dalpha
        // dG/ddx = -G (A dx + B dy), dG/ddy = -G (C dy + B dx); d(conic) = -0.5 G (dx^2, dx dy, dy^2)
        atomicAdd(dst + 0, make_float4((-b4.x * sx - b4.y * sy) * ddelx_dx, (-b4.z * sy - b4.y * sx) * ddely_dy, v[2], v[3]));
        const float h = -0.5f * o;
        atomicAdd(dst + 1, make_float4(h * v[4], h * v[5], h * v[6], 0.f));
      }
      atomicAdd(dst + 2, make_float4(v[8], v[9], v[10], 0.f));
    }
  }
}

int launch_blend_backward(const EgsBackwardArgs& a, const GeomPtrs& g, const uint32_t* point_list,
                          const ImagePtrs& im, const BackwardScratch& s, cudaStream_t stream) {
  const dim3 grid(tiles_x(a.width), tiles_y(a.height), 1);
  StageTimer timer(stream, ST_BLEND_BWD);
  // pixels per lane: 4 (two warps per tile), except on images with so few tiles that two warps each would
  // leave most of the GPU's warp slots empty (256 x 256: 0.106 against 0.081 ms);
  // egs_debug_option("blend_backward", "2" | "4") forces either for the A/B
  const int forced = debug_options().bwd_ppl;
  const int ppl = forced ? forced : ((long long)grid.x * grid.y < 148 * 8 ? 2 : 4);
  // no upstream gradient for depth and alpha (both NULL): the kernel without their terms
  const bool da = a.dL_dout_depth != nullptr || a.dL_dout_alpha != nullptr;
#define EGS_BWD_LAUNCH(PPL, CO, DA)                                                                               \
  blend_backward_kernel<PPL, CO, DA><<<grid, 256 / PPL, 0, stream>>>(                                             \
      im.ranges, point_list, a.width, a.height, (int)grid.x, a.background, g.means2D, g.conic_opacity, g.rgb_depth, \
      a.out_alpha, im.n_contrib, a.dL_dout_color, a.dL_dout_depth, a.dL_dout_alpha, s.acc)
  if (a.colors_only) {
    if (ppl == 2) EGS_BWD_LAUNCH(2, true, false); else EGS_BWD_LAUNCH(4, true, false);
  } else if (da) {
    if (ppl == 2) EGS_BWD_LAUNCH(2, false, true); else EGS_BWD_LAUNCH(4, false, true);
  } else {
    if (ppl == 2) EGS_BWD_LAUNCH(2, false, false); else EGS_BWD_LAUNCH(4, false, false);
  }
#undef EGS_BWD_LAUNCH
  EGS_AFTER_LAUNCH(stream, a.debug);
  return EGS_OK;
}

}  // namespace egs
