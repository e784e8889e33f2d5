/*
 * cpu_rasterizer.c -- CPU restatement of the reference rasterizer's algorithm.
 *
 * TEST INFRASTRUCTURE ONLY. This file is the parity oracle and the timed CPU baseline. Nothing
 * in the product path (egogaussian_b200/) may import, link or call it; only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs do.
 *
 * The reference (zdwww/EgoGaussian, submodules/diff-gaussian-rasterization, "DGR") has no CPU
 * path; this restates its CUDA kernels in plain C, fp32, one function per kernel:
 *   orc_preprocess           DGR/cuda_rasterizer/forward.cu:155-256 (+ :20-152, auxiliary.h:41-164)
 *   orc_bin                  DGR/cuda_rasterizer/rasterizer_impl.cu:70-138, 301-319
 *   orc_blend_forward        DGR/cuda_rasterizer/forward.cu:261-381
 *   orc_blend_backward       DGR/cuda_rasterizer/backward.cu:415-601
 *   orc_preprocess_backward  DGR/cuda_rasterizer/backward.cu:20-412
 * Tiles are processed in parallel with OpenMP; per-pixel / per-Gaussian arithmetic is scalar
 * fp32 in the reference's order (the GPU contracts a*b+c to FMA, the CPU build does not, so
 * float results agree to rounding, not bit for bit; integer results agree except where a float
 * sits on a rounding boundary). Gradient sums that the GPU forms with float atomics in arbitrary
 * order are accumulated here in double and rounded once.
 *
 * Parity status: pinned against outputs of the reference itself (oracle/_ref, run on a B200)
 * stored under tests/golden/ -- see tests/test_oracle_golden.py.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define TILE 16

static const float SH_C0 = 0.28209479177387814f;
static const float SH_C1 = 0.4886025119029199f;
static const float SH_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                               -1.0925484305920792f, 0.5462742152960396f};
static const float SH_C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                               0.3731763325901154f, -0.4570457994644658f, 1.445305721320277f,
                               -0.5900435899266435f};

typedef struct { float c[3][3]; } m3; /* column-major like glm: c[col][row] */

static m3 m3_mul(const m3* A, const m3* B) { /* glm/detail/type_mat3x3.inl:486-520 order */
  m3 R;
  for (int col = 0; col < 3; ++col)
    for (int row = 0; row < 3; ++row)
      R.c[col][row] = A->c[0][row] * B->c[col][0] + A->c[1][row] * B->c[col][1] + A->c[2][row] * B->c[col][2];
  return R;
}
static m3 m3_t(const m3* A) {
  m3 R;
  for (int c = 0; c < 3; ++c) for (int r = 0; r < 3; ++r) R.c[c][r] = A->c[r][c];
  return R;
}
static void xf4x3(const float* p, const float* m, float* o) { /* auxiliary.h:58-66 */
  o[0] = m[0] * p[0] + m[4] * p[1] + m[8] * p[2] + m[12];
  o[1] = m[1] * p[0] + m[5] * p[1] + m[9] * p[2] + m[13];
  o[2] = m[2] * p[0] + m[6] * p[1] + m[10] * p[2] + m[14];
}
static void xf4x4(const float* p, const float* m, float* o) { /* auxiliary.h:68-76 */
  o[0] = m[0] * p[0] + m[4] * p[1] + m[8] * p[2] + m[12];
  o[1] = m[1] * p[0] + m[5] * p[1] + m[9] * p[2] + m[13];
  o[2] = m[2] * p[0] + m[6] * p[1] + m[10] * p[2] + m[14];
  o[3] = m[3] * p[0] + m[7] * p[1] + m[11] * p[2] + m[15];
}
static float ndc2pix(float v, int S) { return (float)(((v + 1.0) * S - 1.0) * 0.5); } /* auxiliary.h:41-44 */
static int imin(int a, int b) { return a < b ? a : b; }
static int imax(int a, int b) { return a > b ? a : b; }
static void get_rect(float px, float py, int r, int gx, int gy, int* lo, int* hi) { /* auxiliary.h:46-56 */
  lo[0] = imin(gx, imax(0, (int)((px - r) / TILE)));
  lo[1] = imin(gy, imax(0, (int)((py - r) / TILE)));
  hi[0] = imin(gx, imax(0, (int)((px + r + TILE - 1) / TILE)));
  hi[1] = imin(gy, imax(0, (int)((py + r + TILE - 1) / TILE)));
}

static void rot_matrix(const float* q, m3* R) { /* forward.cu:128-138, quaternion NOT normalised */
  const float r = q[0], x = q[1], y = q[2], z = q[3];
  const float v[9] = {1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y),
                      2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x),
                      2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y)};
  for (int c = 0; c < 3; ++c) for (int rr = 0; rr < 3; ++rr) R->c[c][rr] = v[3 * c + rr];
}

static void cov3d(const float* s, float mod, const float* q, float* cov6) { /* forward.cu:118-152 */
  m3 S; memset(&S, 0, sizeof S);
  S.c[0][0] = mod * s[0]; S.c[1][1] = mod * s[1]; S.c[2][2] = mod * s[2];
  m3 R; rot_matrix(q, &R);
  m3 M = m3_mul(&S, &R);
  m3 Mt = m3_t(&M);
  m3 Sg = m3_mul(&Mt, &M);
  cov6[0] = Sg.c[0][0]; cov6[1] = Sg.c[0][1]; cov6[2] = Sg.c[0][2];
  cov6[3] = Sg.c[1][1]; cov6[4] = Sg.c[1][2]; cov6[5] = Sg.c[2][2];
}

/* shared by forward and backward: t (clamped), J, W, T and the 2D covariance (forward.cu:74-113) */
typedef struct { float t[3]; float txtz, tytz, limx, limy; m3 J, W, T, V, cov; } cov2d_ctx;
static void cov2d(const float* mean, float fx, float fy, float tanx, float tany, const float* c6,
                  const float* view, cov2d_ctx* o) {
  xf4x3(mean, view, o->t);
  o->limx = 1.3f * tanx; o->limy = 1.3f * tany;
  o->txtz = o->t[0] / o->t[2]; o->tytz = o->t[1] / o->t[2];
  o->t[0] = fminf(o->limx, fmaxf(-o->limx, o->txtz)) * o->t[2];
  o->t[1] = fminf(o->limy, fmaxf(-o->limy, o->tytz)) * o->t[2];
  const float tz = o->t[2];
  memset(&o->J, 0, sizeof(m3));
  o->J.c[0][0] = fx / tz; o->J.c[0][2] = -(fx * o->t[0]) / (tz * tz);
  o->J.c[1][1] = fy / tz; o->J.c[1][2] = -(fy * o->t[1]) / (tz * tz);
  const float w[9] = {view[0], view[4], view[8], view[1], view[5], view[9], view[2], view[6], view[10]};
  for (int c = 0; c < 3; ++c) for (int r = 0; r < 3; ++r) o->W.c[c][r] = w[3 * c + r];
  o->T = m3_mul(&o->W, &o->J);
  const float v[9] = {c6[0], c6[1], c6[2], c6[1], c6[3], c6[4], c6[2], c6[4], c6[5]};
  for (int c = 0; c < 3; ++c) for (int r = 0; r < 3; ++r) o->V.c[c][r] = v[3 * c + r];
  m3 Tt = m3_t(&o->T), Vt = m3_t(&o->V);
  m3 A = m3_mul(&Tt, &Vt);
  o->cov = m3_mul(&A, &o->T);
  o->cov.c[0][0] += 0.3f;
  o->cov.c[1][1] += 0.3f;
}

/* forward.cu:20-71 */
static void sh_rgb(int deg, const float* pos, const float* cam, const float* sh, float* rgb, uint8_t* clamped) {
  float d[3] = {pos[0] - cam[0], pos[1] - cam[1], pos[2] - cam[2]};
  const float len = sqrtf(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
  const float x = d[0] / len, y = d[1] / len, z = d[2] / len;
  for (int ch = 0; ch < 3; ++ch) {
#define S(k) sh[3 * (k) + ch]
    float r = SH_C0 * S(0);
    if (deg > 0) {
      r = r - SH_C1 * y * S(1) + SH_C1 * z * S(2) - SH_C1 * x * S(3);
      if (deg > 1) {
        const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
        r = r + SH_C2[0] * xy * S(4) + SH_C2[1] * yz * S(5) + SH_C2[2] * (2.0f * zz - xx - yy) * S(6) +
            SH_C2[3] * xz * S(7) + SH_C2[4] * (xx - yy) * S(8);
        if (deg > 2) {
          r = r + SH_C3[0] * y * (3.0f * xx - yy) * S(9) + SH_C3[1] * xy * z * S(10) +
              SH_C3[2] * y * (4.0f * zz - xx - yy) * S(11) +
              SH_C3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * S(12) +
              SH_C3[4] * x * (4.0f * zz - xx - yy) * S(13) + SH_C3[5] * z * (xx - yy) * S(14) +
              SH_C3[6] * x * (xx - 3.0f * yy) * S(15);
        }
      }
    }
#undef S
    r += 0.5f;
    clamped[ch] = r < 0.f;
    rgb[ch] = fmaxf(r, 0.f);
  }
}

/* One pass over all Gaussians. Arrays for culled Gaussians keep their previous contents except
 * radii / tiles_touched (zero), as in the reference. radius_f receives 3*sqrt(lambda_max) before
 * the ceil (0 for culled) so tests can recognise rounding-boundary cases. */
void orc_preprocess(int P, int D, int M, const float* means, const float* scales, float mod,
                    const float* rots, const float* opac, const float* shs, const float* cov_pre,
                    const float* colors_pre, const float* view, const float* proj, const float* campos,
                    int W, int H, float tanx, float tany, int32_t* radii, float* means2D, float* depths,
                    float* cov3D, float* conic_opacity, float* rgb, uint8_t* clamped,
                    uint32_t* tiles_touched, float* radius_f) {
  const float fy = H / (2.0f * tany), fx = W / (2.0f * tanx);
  const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
#pragma omp parallel for schedule(static)
  for (int i = 0; i < P; ++i) {
    radii[i] = 0; tiles_touched[i] = 0;
    if (radius_f) radius_f[i] = 0.f;
    const float* p = means + 3 * i;
    float pv[3]; xf4x3(p, view, pv);
    if (pv[2] <= 0.2f) continue; /* auxiliary.h:150 */
    float ph[4]; xf4x4(p, proj, ph);
    const float pw = 1.0f / (ph[3] + 0.0000001f);
    const float pp[2] = {ph[0] * pw, ph[1] * pw};
    float c6[6];
    if (cov_pre) memcpy(c6, cov_pre + 6 * i, sizeof c6);
    else { cov3d(scales + 3 * i, mod, rots + 4 * i, c6); memcpy(cov3D + 6 * i, c6, sizeof c6); }
    cov2d_ctx cc; cov2d(p, fx, fy, tanx, tany, c6, view, &cc);
    const float a = cc.cov.c[0][0], b = cc.cov.c[0][1], c = cc.cov.c[1][1];
    const float det = a * c - b * b;
    if (det == 0.0f) continue;
    const float det_inv = 1.f / det;
    const float mid = 0.5f * (a + c);
    const float l1 = mid + sqrtf(fmaxf(0.1f, mid * mid - det));
    const float l2 = mid - sqrtf(fmaxf(0.1f, mid * mid - det));
    const float rf = 3.f * sqrtf(fmaxf(l1, l2));
    const float my_radius = ceilf(rf);
    const float px = ndc2pix(pp[0], W), py = ndc2pix(pp[1], H);
    int lo[2], hi[2];
    get_rect(px, py, (int)my_radius, gx, gy, lo, hi);
    if ((hi[0] - lo[0]) * (hi[1] - lo[1]) == 0) continue;
    if (!colors_pre) sh_rgb(D, p, campos, shs + (size_t)i * M * 3, rgb + 3 * i, clamped + 3 * i);
    else memcpy(rgb + 3 * i, colors_pre + 3 * i, 3 * sizeof(float));
    depths[i] = pv[2];
    radii[i] = (int)my_radius;
    if (radius_f) radius_f[i] = rf;
    means2D[2 * i] = px; means2D[2 * i + 1] = py;
    conic_opacity[4 * i] = c * det_inv; conic_opacity[4 * i + 1] = -b * det_inv;
    conic_opacity[4 * i + 2] = a * det_inv; conic_opacity[4 * i + 3] = opac[i];
    tiles_touched[i] = (uint32_t)((hi[1] - lo[1]) * (hi[0] - lo[0]));
  }
}

/* Duplicate (rasterizer_impl.cu:70-111), stable sort by key (:301-309), tile ranges (:116-138).
 * keys/vals must hold sum(tiles_touched) entries; tmp_keys/tmp_vals the same (scratch). */
int64_t orc_bin(int P, const float* means2D, const float* depths, const int32_t* radii, int W, int H,
                uint64_t* keys, uint32_t* vals, uint64_t* tmp_keys, uint32_t* tmp_vals, uint32_t* ranges) {
  const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
  int64_t n = 0;
  for (int i = 0; i < P; ++i) {
    if (radii[i] <= 0) continue;
    int lo[2], hi[2];
    get_rect(means2D[2 * i], means2D[2 * i + 1], radii[i], gx, gy, lo, hi);
    uint32_t dbits; memcpy(&dbits, depths + i, 4);
    for (int y = lo[1]; y < hi[1]; ++y)
      for (int x = lo[0]; x < hi[0]; ++x) {
        keys[n] = ((uint64_t)(uint32_t)(y * gx + x) << 32) | dbits;
        vals[n] = (uint32_t)i;
        ++n;
      }
  }
  /* stable LSD radix sort, 16-bit digits over all 64 bits (the bits above 32+bit are zero) */
  size_t* cnt = (size_t*)malloc(65536 * sizeof(size_t));
  uint64_t *ka = keys, *kb = tmp_keys; uint32_t *va = vals, *vb = tmp_vals;
  for (int pass = 0; pass < 4; ++pass) {
    const int sh = 16 * pass;
    memset(cnt, 0, 65536 * sizeof(size_t));
    for (int64_t i = 0; i < n; ++i) cnt[(ka[i] >> sh) & 0xffff]++;
    size_t run = 0;
    for (int d = 0; d < 65536; ++d) { size_t c = cnt[d]; cnt[d] = run; run += c; }
    for (int64_t i = 0; i < n; ++i) { size_t dst = cnt[(ka[i] >> sh) & 0xffff]++; kb[dst] = ka[i]; vb[dst] = va[i]; }
    uint64_t* tk = ka; ka = kb; kb = tk; uint32_t* tv = va; va = vb; vb = tv;
  }
  free(cnt); /* 4 passes: result is back in keys / vals */
  memset(ranges, 0, (size_t)gx * gy * 2 * sizeof(uint32_t));
  for (int64_t i = 0; i < n; ++i) {
    const uint32_t cur = (uint32_t)(keys[i] >> 32);
    if (i == 0) ranges[2 * cur] = 0;
    else {
      const uint32_t prev = (uint32_t)(keys[i - 1] >> 32);
      if (cur != prev) { ranges[2 * prev + 1] = (uint32_t)i; ranges[2 * cur] = (uint32_t)i; }
    }
    if (i == n - 1) ranges[2 * cur + 1] = (uint32_t)n;
  }
  return n;
}

/* forward.cu:261-381. colors: [P,3]; outputs planar [3,H,W], [H,W], [H,W]; returns the number of
 * list entries visited (sum of n_contrib) through *fragments when non-NULL. */
void orc_blend_forward(int W, int H, const uint32_t* ranges, const uint32_t* point_list,
                       const float* means2D, const float* colors, const float* depths,
                       const float* conic_opacity, const float* bg, float* out_color, float* out_depth,
                       float* out_alpha, uint32_t* n_contrib, int64_t* fragments) {
  const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
  int64_t frag = 0;
#pragma omp parallel for schedule(dynamic, 4) reduction(+ : frag)
  for (int tile = 0; tile < gx * gy; ++tile) {
    const int tx = tile % gx, ty = tile / gx;
    const uint32_t r0 = ranges[2 * tile], r1 = ranges[2 * tile + 1];
    for (int ly = 0; ly < TILE; ++ly) for (int lx = 0; lx < TILE; ++lx) {
      const int px = tx * TILE + lx, py = ty * TILE + ly;
      if (px >= W || py >= H) continue;
      const float pxf = (float)px, pyf = (float)py;
      float T = 1.0f, C[3] = {0, 0, 0}, weight = 0, Dp = 0;
      uint32_t contributor = 0, last = 0;
      for (uint32_t k = r0; k < r1; ++k) {
        contributor++;
        const uint32_t id = point_list[k];
        const float dx = means2D[2 * id] - pxf, dy = means2D[2 * id + 1] - pyf;
        const float* co = conic_opacity + 4 * id;
        const float power = -0.5f * (co[0] * dx * dx + co[2] * dy * dy) - co[1] * dx * dy;
        if (power > 0.0f) continue;
        const float alpha = fminf(0.99f, co[3] * expf(power));
        if (alpha < 1.0f / 255.0f) continue;
        const float test_T = T * (1 - alpha);
        if (test_T < 0.0001f) break; /* done */
        for (int ch = 0; ch < 3; ++ch) C[ch] += colors[3 * id + ch] * alpha * T;
        weight += alpha * T;
        Dp += depths[id] * alpha * T;
        T = test_T;
        last = contributor;
      }
      const size_t pix = (size_t)W * py + px;
      n_contrib[pix] = last;
      frag += last;
      for (int ch = 0; ch < 3; ++ch) out_color[(size_t)ch * H * W + pix] = C[ch] + T * bg[ch];
      out_alpha[pix] = weight;
      out_depth[pix] = Dp;
    }
  }
  if (fragments) *fragments = frag;
}

/* backward.cu:415-601. Accumulators (double): dmean2D [P,2], dconic [P,3] (x,y,w), dopacity [P],
 * dcolors [P,3], ddepths [P]; the caller zero-fills them. */
void orc_blend_backward(int W, int H, const uint32_t* ranges, const uint32_t* point_list, const float* bg,
                        const float* means2D, const float* conic_opacity, const float* colors,
                        const float* depths, const float* alphas, const uint32_t* n_contrib,
                        const float* dL_dpix, const float* dL_ddepthpix, const float* dL_dalphapix,
                        double* dmean2D, double* dconic, double* dopacity, double* dcolors, double* ddepths) {
  const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
  const float ddelx_dx = (float)(0.5 * W), ddely_dy = (float)(0.5 * H);
#pragma omp parallel for schedule(dynamic, 4)
  for (int tile = 0; tile < gx * gy; ++tile) {
    const int tx = tile % gx, ty = tile / gx;
    const uint32_t r0 = ranges[2 * tile], r1 = ranges[2 * tile + 1];
    const uint32_t len = r1 - r0;
    if (len == 0) continue;
    uint32_t maxc = 0;
    for (int ly = 0; ly < TILE; ++ly) for (int lx = 0; lx < TILE; ++lx) {
      const int px = tx * TILE + lx, py = ty * TILE + ly;
      if (px < W && py < H && n_contrib[(size_t)W * py + px] > maxc) maxc = n_contrib[(size_t)W * py + px];
    }
    if (maxc == 0) continue;
    double* loc = (double*)calloc((size_t)maxc * 10, sizeof(double)); /* per list position */
    for (int ly = 0; ly < TILE; ++ly) for (int lx = 0; lx < TILE; ++lx) {
      const int px = tx * TILE + lx, py = ty * TILE + ly;
      if (px >= W || py >= H) continue;
      const size_t pix = (size_t)W * py + px;
      const float pxf = (float)px, pyf = (float)py;
      const float T_final = 1 - alphas[pix];
      float T = T_final;
      const uint32_t last = n_contrib[pix];
      float accum_rec[3] = {0, 0, 0}, dLp[3], accum_d = 0, accum_a = 0;
      for (int ch = 0; ch < 3; ++ch) dLp[ch] = dL_dpix[(size_t)ch * H * W + pix];
      const float dLd = dL_ddepthpix[pix], dLa = dL_dalphapix[pix];
      float last_alpha = 0, last_color[3] = {0, 0, 0}, last_depth = 0;
      float bg_dot = 0;
      for (int ch = 0; ch < 3; ++ch) bg_dot += bg[ch] * dLp[ch];
      for (int64_t pos = (int64_t)last - 1; pos >= 0; --pos) { /* entries >= last are skipped (:511-513) */
        const uint32_t id = point_list[r0 + pos];
        const float dx = means2D[2 * id] - pxf, dy = means2D[2 * id + 1] - pyf;
        const float* co = conic_opacity + 4 * id;
        const float power = -0.5f * (co[0] * dx * dx + co[2] * dy * dy) - co[1] * dx * dy;
        if (power > 0.0f) continue;
        const float G = expf(power);
        const float alpha = fminf(0.99f, co[3] * G);
        if (alpha < 1.0f / 255.0f) continue;
        T = T / (1.f - alpha);
        const float w = alpha * T;
        float dL_dopa = 0.0f;
        double* L = loc + (size_t)pos * 10;
        for (int ch = 0; ch < 3; ++ch) {
          const float c = colors[3 * id + ch];
          accum_rec[ch] = last_alpha * last_color[ch] + (1.f - last_alpha) * accum_rec[ch];
          last_color[ch] = c;
          dL_dopa += (c - accum_rec[ch]) * dLp[ch];
          L[7 + ch] += w * dLp[ch];
        }
        const float cd = depths[id];
        accum_d = last_alpha * last_depth + (1.f - last_alpha) * accum_d;
        last_depth = cd;
        dL_dopa += (cd - accum_d) * dLd;
        L[6] += w * dLd;
        accum_a = last_alpha + (1.f - last_alpha) * accum_a;
        dL_dopa += (1 - accum_a) * dLa;
        dL_dopa *= T;
        last_alpha = alpha;
        dL_dopa += (-T_final / (1.f - alpha)) * bg_dot;
        const float dL_dG = co[3] * dL_dopa;
        const float gdx = G * dx, gdy = G * dy;
        const float dG_ddelx = -gdx * co[0] - gdy * co[1];
        const float dG_ddely = -gdy * co[2] - gdx * co[1];
        L[0] += dL_dG * dG_ddelx * ddelx_dx;
        L[1] += dL_dG * dG_ddely * ddely_dy;
        L[2] += -0.5f * gdx * dx * dL_dG;
        L[3] += -0.5f * gdx * dy * dL_dG;
        L[4] += -0.5f * gdy * dy * dL_dG;
        L[5] += G * dL_dopa;
      }
    }
    for (uint32_t pos = 0; pos < maxc; ++pos) {
      const uint32_t id = point_list[r0 + pos];
      const double* L = loc + (size_t)pos * 10;
#define ADD(dst, v) if ((v) != 0.0) { _Pragma("omp atomic") dst += (v); }
      ADD(dmean2D[2 * id], L[0]) ADD(dmean2D[2 * id + 1], L[1])
      ADD(dconic[3 * id], L[2]) ADD(dconic[3 * id + 1], L[3]) ADD(dconic[3 * id + 2], L[4])
      ADD(dopacity[id], L[5]) ADD(ddepths[id], L[6])
      ADD(dcolors[3 * id], L[7]) ADD(dcolors[3 * id + 1], L[8]) ADD(dcolors[3 * id + 2], L[9])
#undef ADD
    }
    free(loc);
  }
}

typedef struct { float x, y, z; } v3;
static v3 v3s(float s, v3 a) { v3 r = {s * a.x, s * a.y, s * a.z}; return r; }
static v3 v3a(v3 a, v3 b) { v3 r = {a.x + b.x, a.y + b.y, a.z + b.z}; return r; }
static float v3d(v3 a, v3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }

/* backward.cu:20-139 */
static void sh_backward(int deg, int M, const float* pos, const float* cam, const float* shp,
                        const uint8_t* clamped, const float* dL_dcolor, float* dL_dsh, float* dmean_add) {
  const float d0[3] = {pos[0] - cam[0], pos[1] - cam[1], pos[2] - cam[2]};
  const float len = sqrtf(d0[0] * d0[0] + d0[1] * d0[1] + d0[2] * d0[2]);
  const float x = d0[0] / len, y = d0[1] / len, z = d0[2] / len;
  v3 g = {dL_dcolor[0] * (clamped[0] ? 0.f : 1.f), dL_dcolor[1] * (clamped[1] ? 0.f : 1.f),
          dL_dcolor[2] * (clamped[2] ? 0.f : 1.f)};
#define SHV(k) ((v3){shp[3 * (k)], shp[3 * (k) + 1], shp[3 * (k) + 2]})
#define PUT(k, w) do { dL_dsh[3 * (k)] = (w) * g.x; dL_dsh[3 * (k) + 1] = (w) * g.y; dL_dsh[3 * (k) + 2] = (w) * g.z; } while (0)
  for (int k = 0; k < 3 * M; ++k) dL_dsh[k] = 0.f;
  v3 ddx = {0, 0, 0}, ddy = {0, 0, 0}, ddz = {0, 0, 0};
  PUT(0, SH_C0);
  if (deg > 0) {
    PUT(1, -SH_C1 * y); PUT(2, SH_C1 * z); PUT(3, -SH_C1 * x);
    ddx = v3s(-SH_C1, SHV(3)); ddy = v3s(-SH_C1, SHV(1)); ddz = v3s(SH_C1, SHV(2));
    if (deg > 1) {
      const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
      PUT(4, SH_C2[0] * xy); PUT(5, SH_C2[1] * yz); PUT(6, SH_C2[2] * (2.f * zz - xx - yy));
      PUT(7, SH_C2[3] * xz); PUT(8, SH_C2[4] * (xx - yy));
      ddx = v3a(ddx, v3a(v3a(v3s(SH_C2[0] * y, SHV(4)), v3s(SH_C2[2] * 2.f * -x, SHV(6))),
                         v3a(v3s(SH_C2[3] * z, SHV(7)), v3s(SH_C2[4] * 2.f * x, SHV(8)))));
      ddy = v3a(ddy, v3a(v3a(v3s(SH_C2[0] * x, SHV(4)), v3s(SH_C2[1] * z, SHV(5))),
                         v3a(v3s(SH_C2[2] * 2.f * -y, SHV(6)), v3s(SH_C2[4] * 2.f * -y, SHV(8)))));
      ddz = v3a(ddz, v3a(v3a(v3s(SH_C2[1] * y, SHV(5)), v3s(SH_C2[2] * 2.f * 2.f * z, SHV(6))),
                         v3s(SH_C2[3] * x, SHV(7))));
      if (deg > 2) {
        PUT(9, SH_C3[0] * y * (3.f * xx - yy)); PUT(10, SH_C3[1] * xy * z);
        PUT(11, SH_C3[2] * y * (4.f * zz - xx - yy)); PUT(12, SH_C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy));
        PUT(13, SH_C3[4] * x * (4.f * zz - xx - yy)); PUT(14, SH_C3[5] * z * (xx - yy));
        PUT(15, SH_C3[6] * x * (xx - 3.f * yy));
        const float wx[7] = {SH_C3[0] * 3.f * 2.f * xy, SH_C3[1] * yz, SH_C3[2] * -2.f * xy, SH_C3[3] * -3.f * 2.f * xz,
                             SH_C3[4] * (-3.f * xx + 4.f * zz - yy), SH_C3[5] * 2.f * xz, SH_C3[6] * 3.f * (xx - yy)};
        const float wy[7] = {SH_C3[0] * 3.f * (xx - yy), SH_C3[1] * xz, SH_C3[2] * (-3.f * yy + 4.f * zz - xx),
                             SH_C3[3] * -3.f * 2.f * yz, SH_C3[4] * -2.f * xy, SH_C3[5] * -2.f * yz, SH_C3[6] * -3.f * 2.f * xy};
        const float wz[7] = {0.f, SH_C3[1] * xy, SH_C3[2] * 4.f * 2.f * yz, SH_C3[3] * 3.f * (2.f * zz - xx - yy),
                             SH_C3[4] * 4.f * 2.f * xz, SH_C3[5] * (xx - yy), 0.f};
        for (int k = 0; k < 7; ++k) {
          ddx = v3a(ddx, v3s(wx[k], SHV(9 + k)));
          ddy = v3a(ddy, v3s(wy[k], SHV(9 + k)));
          ddz = v3a(ddz, v3s(wz[k], SHV(9 + k)));
        }
      }
    }
  }
#undef SHV
#undef PUT
  const float dd[3] = {v3d(ddx, g), v3d(ddy, g), v3d(ddz, g)};
  /* through the normalisation: auxiliary.h:107-117 */
  const float s2 = d0[0] * d0[0] + d0[1] * d0[1] + d0[2] * d0[2];
  const float inv = 1.0f / sqrtf(s2 * s2 * s2);
  dmean_add[0] = ((+s2 - d0[0] * d0[0]) * dd[0] - d0[1] * d0[0] * dd[1] - d0[2] * d0[0] * dd[2]) * inv;
  dmean_add[1] = (-d0[0] * d0[1] * dd[0] + (s2 - d0[1] * d0[1]) * dd[1] - d0[2] * d0[1] * dd[2]) * inv;
  dmean_add[2] = (-d0[0] * d0[2] * dd[0] - d0[1] * d0[2] * dd[1] + (s2 - d0[2] * d0[2]) * dd[2]) * inv;
}

/* backward.cu:144-274 + :346-412 + :278-341. Inputs are fp32 gradients per Gaussian:
 * dmean2D [P,2], dconic [P,3] (x,y,w), dcolor [P,3], ddepth [P]. cov3D [P,6] is the covariance
 * used in forward (precomputed or from orc_preprocess). Outputs: dmeans3D [P,3], dcov3D [P,6],
 * dsh [P,M,3] (if shs), dscales [P,3] / drots [P,4] (if scales). All outputs zero for culled. */
void orc_preprocess_backward(int P, int D, int M, const float* means, const int32_t* radii,
                             const float* shs, const uint8_t* clamped, const float* scales,
                             const float* rots, float mod, const float* cov3D, const float* view,
                             const float* proj, const float* campos, int W, int H, float tanx, float tany,
                             const float* dmean2D, const float* dconic, const float* dcolor,
                             const float* ddepth, float* dmeans3D, float* dcov3D, float* dsh,
                             float* dscales, float* drots) {
  const float hy = H / (2.0f * tany), hx = W / (2.0f * tanx);
#pragma omp parallel for schedule(static)
  for (int i = 0; i < P; ++i) {
    for (int k = 0; k < 3; ++k) dmeans3D[3 * i + k] = 0.f;
    for (int k = 0; k < 6; ++k) dcov3D[6 * i + k] = 0.f;
    if (shs) for (int k = 0; k < 3 * M; ++k) dsh[(size_t)i * M * 3 + k] = 0.f;
    if (scales) { for (int k = 0; k < 3; ++k) dscales[3 * i + k] = 0.f; for (int k = 0; k < 4; ++k) drots[4 * i + k] = 0.f; }
    if (!(radii[i] > 0)) continue;
    const float* mean = means + 3 * i;
    const float* c6 = cov3D + 6 * i;
    cov2d_ctx cc; cov2d(mean, hx, hy, tanx, tany, c6, view, &cc);
    const float xgm = (cc.txtz < -cc.limx || cc.txtz > cc.limx) ? 0.f : 1.f;
    const float ygm = (cc.tytz < -cc.limy || cc.tytz > cc.limy) ? 0.f : 1.f;
    const float a = cc.cov.c[0][0], b = cc.cov.c[0][1], c = cc.cov.c[1][1];
    const float dA = dconic[3 * i], dB = dconic[3 * i + 1], dC = dconic[3 * i + 2];
    const float denom = a * c - b * b;
    float dL_da = 0, dL_db = 0, dL_dc = 0;
    const float d2i = 1.0f / ((denom * denom) + 0.0000001f);
    const m3* T = &cc.T; const m3* V = &cc.V; const m3* Wm = &cc.W;
    float* dc = dcov3D + 6 * i;
    if (d2i != 0) {
      dL_da = d2i * (-c * c * dA + 2 * b * c * dB + (denom - a * c) * dC);
      dL_dc = d2i * (-a * a * dC + 2 * a * b * dB + (denom - a * c) * dA);
      dL_db = d2i * 2 * (b * c * dA - (denom + 2 * b * b) * dB + a * b * dC);
      dc[0] = (T->c[0][0] * T->c[0][0] * dL_da + T->c[0][0] * T->c[1][0] * dL_db + T->c[1][0] * T->c[1][0] * dL_dc);
      dc[3] = (T->c[0][1] * T->c[0][1] * dL_da + T->c[0][1] * T->c[1][1] * dL_db + T->c[1][1] * T->c[1][1] * dL_dc);
      dc[5] = (T->c[0][2] * T->c[0][2] * dL_da + T->c[0][2] * T->c[1][2] * dL_db + T->c[1][2] * T->c[1][2] * dL_dc);
      dc[1] = 2 * T->c[0][0] * T->c[0][1] * dL_da + (T->c[0][0] * T->c[1][1] + T->c[0][1] * T->c[1][0]) * dL_db + 2 * T->c[1][0] * T->c[1][1] * dL_dc;
      dc[2] = 2 * T->c[0][0] * T->c[0][2] * dL_da + (T->c[0][0] * T->c[1][2] + T->c[0][2] * T->c[1][0]) * dL_db + 2 * T->c[1][0] * T->c[1][2] * dL_dc;
      dc[4] = 2 * T->c[0][2] * T->c[0][1] * dL_da + (T->c[0][1] * T->c[1][2] + T->c[0][2] * T->c[1][1]) * dL_db + 2 * T->c[1][1] * T->c[1][2] * dL_dc;
    }
    float dT[2][3];
    for (int k = 0; k < 3; ++k) {
      const float r0 = T->c[0][0] * V->c[k][0] + T->c[0][1] * V->c[k][1] + T->c[0][2] * V->c[k][2];
      const float r1 = T->c[1][0] * V->c[k][0] + T->c[1][1] * V->c[k][1] + T->c[1][2] * V->c[k][2];
      dT[0][k] = 2 * r0 * dL_da + r1 * dL_db;
      dT[1][k] = 2 * r1 * dL_dc + r0 * dL_db;
    }
    const float dJ00 = Wm->c[0][0] * dT[0][0] + Wm->c[0][1] * dT[0][1] + Wm->c[0][2] * dT[0][2];
    const float dJ02 = Wm->c[2][0] * dT[0][0] + Wm->c[2][1] * dT[0][1] + Wm->c[2][2] * dT[0][2];
    const float dJ11 = Wm->c[1][0] * dT[1][0] + Wm->c[1][1] * dT[1][1] + Wm->c[1][2] * dT[1][2];
    const float dJ12 = Wm->c[2][0] * dT[1][0] + Wm->c[2][1] * dT[1][1] + Wm->c[2][2] * dT[1][2];
    const float tz = 1.f / cc.t[2], tz2 = tz * tz, tz3 = tz2 * tz;
    const float dtx = xgm * -hx * tz2 * dJ02;
    const float dty = ygm * -hy * tz2 * dJ12;
    const float dtz = -hx * tz2 * dJ00 - hy * tz2 * dJ11 + (2 * hx * cc.t[0]) * tz3 * dJ02 + (2 * hy * cc.t[1]) * tz3 * dJ12;
    float dm[3] = {view[0] * dtx + view[1] * dty + view[2] * dtz, view[4] * dtx + view[5] * dty + view[6] * dtz,
                   view[8] * dtx + view[9] * dty + view[10] * dtz};
    /* projection part (:368-379) */
    float mh[4]; xf4x4(mean, proj, mh);
    const float mw = 1.0f / (mh[3] + 0.0000001f);
    const float mul1 = (proj[0] * mean[0] + proj[4] * mean[1] + proj[8] * mean[2] + proj[12]) * mw * mw;
    const float mul2 = (proj[1] * mean[0] + proj[5] * mean[1] + proj[9] * mean[2] + proj[13]) * mw * mw;
    const float gx2 = dmean2D[2 * i], gy2 = dmean2D[2 * i + 1];
    dm[0] += (proj[0] * mw - proj[3] * mul1) * gx2 + (proj[1] * mw - proj[3] * mul2) * gy2;
    dm[1] += (proj[4] * mw - proj[7] * mul1) * gx2 + (proj[5] * mw - proj[7] * mul2) * gy2;
    dm[2] += (proj[8] * mw - proj[11] * mul1) * gx2 + (proj[9] * mw - proj[11] * mul2) * gy2;
    /* depth part (:384-392) */
    const float mul3 = view[2] * mean[0] + view[6] * mean[1] + view[10] * mean[2] + view[14];
    dm[0] += (view[2] - view[3] * mul3) * ddepth[i];
    dm[1] += (view[6] - view[7] * mul3) * ddepth[i];
    dm[2] += (view[10] - view[11] * mul3) * ddepth[i];
    if (shs) {
      float add[3];
      sh_backward(D, M, mean, campos, shs + (size_t)i * M * 3, clamped + 3 * i, dcolor + 3 * i,
                  dsh + (size_t)i * M * 3, add);
      dm[0] += add[0]; dm[1] += add[1]; dm[2] += add[2];
    }
    for (int k = 0; k < 3; ++k) dmeans3D[3 * i + k] = dm[k];
    if (scales) { /* :278-341 */
      const float* q = rots + 4 * i;
      const float r = q[0], x = q[1], y = q[2], z = q[3];
      m3 R; rot_matrix(q, &R);
      const float s[3] = {mod * scales[3 * i], mod * scales[3 * i + 1], mod * scales[3 * i + 2]};
      m3 S; memset(&S, 0, sizeof S); S.c[0][0] = s[0]; S.c[1][1] = s[1]; S.c[2][2] = s[2];
      m3 Mm = m3_mul(&S, &R);
      m3 dS; const float ds[9] = {dc[0], 0.5f * dc[1], 0.5f * dc[2], 0.5f * dc[1], dc[3], 0.5f * dc[4],
                                  0.5f * dc[2], 0.5f * dc[4], dc[5]};
      for (int cc2 = 0; cc2 < 3; ++cc2) for (int rr = 0; rr < 3; ++rr) { dS.c[cc2][rr] = ds[3 * cc2 + rr]; Mm.c[cc2][rr] *= 2.0f; }
      m3 dM = m3_mul(&Mm, &dS);
      m3 Rt = m3_t(&R), dMt = m3_t(&dM);
      for (int k = 0; k < 3; ++k)
        dscales[3 * i + k] = Rt.c[k][0] * dMt.c[k][0] + Rt.c[k][1] * dMt.c[k][1] + Rt.c[k][2] * dMt.c[k][2];
      for (int k = 0; k < 3; ++k) for (int e = 0; e < 3; ++e) dMt.c[k][e] *= s[k];
      float* dq = drots + 4 * i;
      dq[0] = 2 * z * (dMt.c[0][1] - dMt.c[1][0]) + 2 * y * (dMt.c[2][0] - dMt.c[0][2]) + 2 * x * (dMt.c[1][2] - dMt.c[2][1]);
      dq[1] = 2 * y * (dMt.c[1][0] + dMt.c[0][1]) + 2 * z * (dMt.c[2][0] + dMt.c[0][2]) + 2 * r * (dMt.c[1][2] - dMt.c[2][1]) - 4 * x * (dMt.c[2][2] + dMt.c[1][1]);
      dq[2] = 2 * x * (dMt.c[1][0] + dMt.c[0][1]) + 2 * r * (dMt.c[2][0] - dMt.c[0][2]) + 2 * z * (dMt.c[1][2] + dMt.c[2][1]) - 4 * y * (dMt.c[2][2] + dMt.c[0][0]);
      dq[3] = 2 * r * (dMt.c[0][1] - dMt.c[1][0]) + 2 * x * (dMt.c[2][0] + dMt.c[0][2]) + 2 * y * (dMt.c[1][2] + dMt.c[2][1]) - 4 * z * (dMt.c[1][1] + dMt.c[0][0]);
    }
  }
}

/* rasterizer_impl.cu:54-66 */
void orc_mark_visible(int P, const float* means, const float* view, uint8_t* present) {
  for (int i = 0; i < P; ++i) {
    float pv[3]; xf4x3(means + 3 * i, view, pv);
    present[i] = pv[2] > 0.2f;
  }
}

int orc_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}
