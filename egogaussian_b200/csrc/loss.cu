// loss.cu -- fused image loss of the trainers: (1 - lambda) * L1 + lambda * (1 - SSIM), value and
// gradient w.r.t. the rendered image in two tiled kernels.
//
// Replaces utils/loss_utils.py:57-106 (l1_loss, ssim, _ssim: five depthwise 11x11 convolutions,
// ~15 elementwise kernels and their autograd graph) and the gradient mask hook of
// trainers/train_static_bg.py:98 (`grad * (1 - int_mask)`).
//
// The 11x11 window is the outer product of a normalised 1-D Gaussian (sigma 1.5), so each 2-D
// correlation is done as a horizontal and a vertical 11-tap pass over a 42x42 shared-memory tile
// (zero padding outside the image, window not renormalised: F.conv2d(padding=5) semantics).
//
// Forward keeps, per pixel, the three partial derivatives of the SSIM map w.r.t. the three
// window sums that depend on img1 (mu1, E[x^2], E[xy]); because the window is symmetric the
// correlation is its own adjoint, so backward is the same tiled correlation applied to those maps:
//   d ssim_sum / d x = corr(dm_dmu1) + 2 x corr(dm_dE11) + y corr(dm_dE12).
#include "common.cuh"

#include <math.h>

namespace egs {

// A CTA of 256 threads produces a 32x32 output tile; every thread owns 4 outputs that are adjacent
// along the pass direction, so an 11-tap window slides over 14 loaded values instead of being
// re-read 11 times per output (3x fewer shared-memory loads than one output per thread), and the
// halo overhead is (42/32)^2 = 1.7 instead of (26/16)^2 = 2.6.
constexpr int LT = 32;             // output tile edge
constexpr int LHALO = 5;           // window_size // 2
constexpr int LS = LT + 2 * LHALO; // staged tile edge (42)
constexpr int LTHREADS = 256;
constexpr int LR = 4;              // outputs per thread along the pass direction
constexpr int LSPAN = LR + 2 * LHALO;   // values a thread loads for them (14)
constexpr int WIN = 11;
constexpr int LHP = LT + 1;        // pitch of the horizontally filtered rows (odd: conflict-free both ways)
static_assert(LTHREADS == LT * LT / LR, "one thread per 4 outputs in the vertical pass");

struct Window {
  float w[WIN];
};

// utils/loss_utils.py:66-68: exp in double, stored as float32, normalised by the float32 sum
static Window make_window() {
  Window win;
  float sum = 0.f;
  for (int x = 0; x < WIN; ++x) {
    win.w[x] = (float)exp(-(double)((x - WIN / 2) * (x - WIN / 2)) / (2.0 * 1.5 * 1.5));
    sum += win.w[x];
  }
  for (int x = 0; x < WIN; ++x) win.w[x] = win.w[x] / sum;
  return win;
}

__device__ __forceinline__ float block_sum(float v, float* s_red) {
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  if (lane == 0) s_red[warp] = v;
  __syncthreads();
  float t = 0.f;
  if (threadIdx.x < LTHREADS / 32) t = s_red[threadIdx.x];
  if (warp == 0) {
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
  }
  __syncthreads();
  return t;   // valid in thread 0
}

// Stage a (LS x LS) tile of one channel plane with zero padding.
__device__ __forceinline__ void stage_tile(const float* __restrict__ plane, int H, int W, int y0, int x0,
                                           float (*dst)[LS + 1]) {
  for (int idx = threadIdx.x; idx < LS * LS; idx += LTHREADS) {
    const int y = idx / LS, x = idx - y * LS;
    const int gy = y0 + y - LHALO, gx = x0 + x - LHALO;
    dst[y][x] = (gy >= 0 && gy < H && gx >= 0 && gx < W) ? plane[(size_t)gy * W + gx] : 0.f;
  }
}

// Horizontal pass work item `item` (0 .. LS * LT / LR): image row `row` of the staged tile and the
// group of LR adjacent output columns starting at c0. Consecutive lanes take consecutive rows, so
// both the reads (pitch LS + 1, odd) and the writes (pitch LHP, odd) are bank-conflict free.
__device__ __forceinline__ void horizontal_item(int item, int& row, int& c0) {
  const int cg = item / LS;
  row = item - cg * LS;
  c0 = cg * LR;
}

__global__ void __launch_bounds__(LTHREADS)
ssim_forward_kernel(int H, int W, const float* __restrict__ img1, const float* __restrict__ img2, Window win,
                    float* __restrict__ dm_dmu1, float* __restrict__ dm_dE11, float* __restrict__ dm_dE12,
                    float* __restrict__ partial, int nblocks) {
  __shared__ float s_x[LS][LS + 1], s_y[LS][LS + 1];
  __shared__ float s_h[5][LS][LHP];   // horizontal sums of x, y, xx, yy, xy
  __shared__ float s_red[LTHREADS / 32];
  const int x0 = blockIdx.x * LT, y0 = blockIdx.y * LT;
  const size_t plane = (size_t)H * W;
  const float* p1 = img1 + blockIdx.z * plane;
  const float* p2 = img2 + blockIdx.z * plane;
  stage_tile(p1, H, W, y0, x0, s_x);
  stage_tile(p2, H, W, y0, x0, s_y);
  __syncthreads();
  for (int item = threadIdx.x; item < LS * (LT / LR); item += LTHREADS) {
    int row, c0;
    horizontal_item(item, row, c0);
    float u[LSPAN], v[LSPAN];
#pragma unroll
    for (int k = 0; k < LSPAN; ++k) {
      u[k] = s_x[row][c0 + k];
      v[k] = s_y[row][c0 + k];
    }
#pragma unroll
    for (int j = 0; j < LR; ++j) {
      float a = 0.f, b = 0.f, aa = 0.f, bb = 0.f, ab = 0.f;
#pragma unroll
      for (int k = 0; k < WIN; ++k) {
        const float w = win.w[k], p = u[j + k], q = v[j + k];
        a = fmaf(w, p, a);
        b = fmaf(w, q, b);
        aa = fmaf(w, p * p, aa);
        bb = fmaf(w, q * q, bb);
        ab = fmaf(w, p * q, ab);
      }
      s_h[0][row][c0 + j] = a; s_h[1][row][c0 + j] = b; s_h[2][row][c0 + j] = aa;
      s_h[3][row][c0 + j] = bb; s_h[4][row][c0 + j] = ab;
    }
  }
  __syncthreads();
  // vertical pass: column tx, output rows r0 .. r0 + LR - 1
  const int tx = threadIdx.x % LT, r0 = (threadIdx.x / LT) * LR;
  float acc[5][LR];
#pragma unroll
  for (int q = 0; q < 5; ++q) {
    float col[LSPAN];
#pragma unroll
    for (int k = 0; k < LSPAN; ++k) col[k] = s_h[q][r0 + k][tx];
#pragma unroll
    for (int j = 0; j < LR; ++j) {
      float t = 0.f;
#pragma unroll
      for (int k = 0; k < WIN; ++k) t = fmaf(win.w[k], col[j + k], t);
      acc[q][j] = t;
    }
  }
  const int gx = x0 + tx;
  float l1 = 0.f, ssim_sum_t = 0.f;
#pragma unroll
  for (int j = 0; j < LR; ++j) {
    const int gy = y0 + r0 + j;
    if (gx < W && gy < H) {
      const float mu1 = acc[0][j], mu2 = acc[1][j], e11 = acc[2][j], e22 = acc[3][j], e12 = acc[4][j];
      const float C1 = 0.01f * 0.01f, C2 = 0.03f * 0.03f;
      const float mu1_sq = mu1 * mu1, mu2_sq = mu2 * mu2, mu12 = mu1 * mu2;
      const float s1 = e11 - mu1_sq, s2 = e22 - mu2_sq, s12 = e12 - mu12;
      const float A = 2.f * mu12 + C1, B = 2.f * s12 + C2;
      const float Cc = mu1_sq + mu2_sq + C1, D = s1 + s2 + C2;
      const float inv = 1.f / (Cc * D);
      const float ssim = A * B * inv;
      // partial derivatives w.r.t. the three window sums that depend on img1
      const float d_s1 = -ssim / D;            // d m / d sigma1_sq  (= d m / d E11)
      const float d_s12 = 2.f * A * inv;       // d m / d sigma12    (= d m / d E12)
      const float d_mu1 = 2.f * mu2 * B * inv - 2.f * mu1 * ssim / Cc - 2.f * mu1 * d_s1 - mu2 * d_s12;
      const size_t o = blockIdx.z * plane + (size_t)gy * W + gx;
      dm_dmu1[o] = d_mu1;
      dm_dE11[o] = d_s1;
      dm_dE12[o] = d_s12;
      ssim_sum_t += ssim;
      l1 += fabsf(s_x[r0 + j + LHALO][tx + LHALO] - s_y[r0 + j + LHALO][tx + LHALO]);
    }
  }
  const int block = (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
  const float l1_sum = block_sum(l1, s_red);
  const float ssim_sum = block_sum(ssim_sum_t, s_red);
  if (threadIdx.x == 0) {
    partial[block] = l1_sum;
    partial[nblocks + block] = ssim_sum;
  }
}

// Deterministic second stage: one CTA adds the per-block partial sums in a fixed order.
__global__ void __launch_bounds__(1024)
loss_finish_kernel(const float* __restrict__ partial, int nblocks, float inv_count, float lambda,
                   float* __restrict__ out_terms) {
  __shared__ double s_l1[32], s_ss[32];
  double l1 = 0.0, ss = 0.0;
  for (int i = threadIdx.x; i < nblocks; i += 1024) {
    l1 += (double)partial[i];
    ss += (double)partial[nblocks + i];
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    l1 += __shfl_xor_sync(0xffffffffu, l1, o);
    ss += __shfl_xor_sync(0xffffffffu, ss, o);
  }
  if ((threadIdx.x & 31) == 0) {
    s_l1[threadIdx.x >> 5] = l1;
    s_ss[threadIdx.x >> 5] = ss;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    double a = 0.0, b = 0.0;
    for (int w = 0; w < 32; ++w) {
      a += s_l1[w];
      b += s_ss[w];
    }
    const float l1_mean = (float)(a * (double)inv_count), ssim_mean = (float)(b * (double)inv_count);
    out_terms[0] = l1_mean;
    out_terms[1] = ssim_mean;
    out_terms[2] = (1.0f - lambda) * l1_mean + lambda * (1.0f - ssim_mean);
  }
}

__global__ void __launch_bounds__(LTHREADS)
ssim_backward_kernel(int H, int W, const float* __restrict__ img1, const float* __restrict__ img2, Window win,
                     const float* __restrict__ dm_dmu1, const float* __restrict__ dm_dE11,
                     const float* __restrict__ dm_dE12, const int* __restrict__ grad_mask, float k_ssim,
                     float k_l1, float* __restrict__ dL_dimg1) {
  __shared__ float s_m[3][LS][LS + 1];
  __shared__ float s_h[3][LS][LHP];
  const int x0 = blockIdx.x * LT, y0 = blockIdx.y * LT;
  const size_t plane = (size_t)H * W;
  const size_t coff = blockIdx.z * plane;
  stage_tile(dm_dmu1 + coff, H, W, y0, x0, s_m[0]);
  stage_tile(dm_dE11 + coff, H, W, y0, x0, s_m[1]);
  stage_tile(dm_dE12 + coff, H, W, y0, x0, s_m[2]);
  __syncthreads();
  for (int item = threadIdx.x; item < LS * (LT / LR); item += LTHREADS) {
    int row, c0;
    horizontal_item(item, row, c0);
#pragma unroll
    for (int q = 0; q < 3; ++q) {
      float u[LSPAN];
#pragma unroll
      for (int k = 0; k < LSPAN; ++k) u[k] = s_m[q][row][c0 + k];
#pragma unroll
      for (int j = 0; j < LR; ++j) {
        float t = 0.f;
#pragma unroll
        for (int k = 0; k < WIN; ++k) t = fmaf(win.w[k], u[j + k], t);
        s_h[q][row][c0 + j] = t;
      }
    }
  }
  __syncthreads();
  const int tx = threadIdx.x % LT, r0 = (threadIdx.x / LT) * LR;
  float g[3][LR];
#pragma unroll
  for (int q = 0; q < 3; ++q) {
    float col[LSPAN];
#pragma unroll
    for (int k = 0; k < LSPAN; ++k) col[k] = s_h[q][r0 + k][tx];
#pragma unroll
    for (int j = 0; j < LR; ++j) {
      float t = 0.f;
#pragma unroll
      for (int k = 0; k < WIN; ++k) t = fmaf(win.w[k], col[j + k], t);
      g[q][j] = t;
    }
  }
  const int gx = x0 + tx;
  if (gx >= W) return;
#pragma unroll
  for (int j = 0; j < LR; ++j) {
    const int gy = y0 + r0 + j;
    if (gy >= H) break;
    const size_t o = coff + (size_t)gy * W + gx;
    const float x = img1[o], y = img2[o];
    const float d_ssim = g[0][j] + 2.f * x * g[1][j] + y * g[2][j];
    const float diff = x - y;
    const float sign = diff > 0.f ? 1.f : (diff < 0.f ? -1.f : 0.f);   // torch.abs backward: sign(0) = 0
    float grad = k_ssim * d_ssim + k_l1 * sign;
    if (grad_mask != nullptr) grad *= (float)(1 - grad_mask[(size_t)gy * W + gx]);
    dL_dimg1[o] = grad;
  }
}

}  // namespace egs

extern "C" size_t egs_image_loss_workspace_bytes(int channels, int height, int width) {
  if (channels <= 0 || height <= 0 || width <= 0) return 0;
  const size_t maps = (size_t)3 * channels * height * width * sizeof(float);
  const size_t blocks = (size_t)channels * ((height + egs::LT - 1) / egs::LT) * ((width + egs::LT - 1) / egs::LT);
  return maps + 2 * blocks * sizeof(float) + 256;
}

extern "C" int egs_image_loss(const EgsLossArgs* a, float* out_terms, float* dL_dimage, void* workspace,
                              size_t workspace_bytes, void* stream) {
  using namespace egs;
  if (a == nullptr || a->channels <= 0 || a->height <= 0 || a->width <= 0 || !a->image || !a->target || !out_terms ||
      !workspace) {
    set_error("egs_image_loss: NULL argument or empty image");
    return EGS_ERR_INVALID_ARGUMENT;
  }
  if (workspace_bytes < egs_image_loss_workspace_bytes(a->channels, a->height, a->width)) {
    set_error("egs_image_loss: workspace too small (%zu bytes given)", workspace_bytes);
    return EGS_ERR_INVALID_ARGUMENT;
  }
  static const Window win = make_window();
  const int C = a->channels, H = a->height, W = a->width;
  const size_t count = (size_t)C * H * W;
  float* maps = static_cast<float*>(workspace);
  float* partial = maps + 3 * count;
  const dim3 grid((W + LT - 1) / LT, (H + LT - 1) / LT, C);
  const int nblocks = (int)(grid.x * grid.y * grid.z);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  StageTimer timer(st, ST_LOSS);
  ssim_forward_kernel<<<grid, LTHREADS, 0, st>>>(H, W, a->image, a->target, win, maps, maps + count,
                                                  maps + 2 * count, partial, nblocks);
  EGS_AFTER_LAUNCH(st, false);
  const float inv_count = (float)(1.0 / (double)count);
  loss_finish_kernel<<<1, 1024, 0, st>>>(partial, nblocks, inv_count, a->lambda_dssim, out_terms);
  EGS_AFTER_LAUNCH(st, false);
  if (dL_dimage != nullptr) {
    // loss = (1 - lambda) * mean|x - y| + lambda * (1 - mean(ssim_map))
    const float k_ssim = -a->lambda_dssim * inv_count, k_l1 = (1.0f - a->lambda_dssim) * inv_count;
    ssim_backward_kernel<<<grid, LTHREADS, 0, st>>>(H, W, a->image, a->target, win, maps, maps + count,
                                                     maps + 2 * count, a->grad_mask, k_ssim, k_l1, dL_dimage);
    EGS_AFTER_LAUNCH(st, false);
  }
  return EGS_OK;
}
