// torch_binding.cpp -- thin PyTorch C++ extension over the C-ABI (include/egs_rasterizer.h).
//
// Exposes the three functions the reference's Python layer calls, with the same argument order
// and return tuples, so diff_gaussian_rasterization/__init__.py-style callers work unchanged:
//   rasterize_gaussians           <- RasterizeGaussiansCUDA          DGR/rasterize_points.cu:35-119
//   rasterize_gaussians_backward  <- RasterizeGaussiansBackwardCUDA  DGR/rasterize_points.cu:121-208
//   mark_visible                  <- markVisible                     DGR/rasterize_points.cu:210-229
// (pybind names as in DGR/ext.cpp:15-18). torch is plumbing only: tensor allocation, the current
// stream and a device guard. All compute happens behind the C-ABI in libegs_rasterizer.so.
#include <torch/extension.h>
#include <ATen/cuda/CUDAContext.h>
#include <c10/cuda/CUDAGuard.h>

#include <tuple>
#include <vector>

#include "../../include/egs_rasterizer.h"

namespace {

void check(int rc, const char* what) {
  TORCH_CHECK(rc == EGS_OK, what, " failed (", rc, "): ", egs_last_error());
}

// "absent" inputs are empty tensors (the reference passes torch.Tensor([]) on the CPU)
const float* opt_ptr(const torch::Tensor& t, torch::Tensor& keep, const torch::Device& dev, const char* name) {
  if (!t.defined() || t.numel() == 0) return nullptr;
  TORCH_CHECK(t.device() == dev, name, " must be on ", dev);
  TORCH_CHECK(t.scalar_type() == torch::kFloat32, name, " must be float32");
  keep = t.contiguous();
  return keep.data_ptr<float>();
}

const float* req_ptr(const torch::Tensor& t, torch::Tensor& keep, const torch::Device& dev, const char* name) {
  TORCH_CHECK(t.defined() && t.numel() > 0, name, " is required");
  return opt_ptr(t, keep, dev, name);
}

struct BlobCtx {
  torch::Tensor blob;
  torch::Device dev;
};

// Recent num_rendered per problem size (a handful of entries per host thread is plenty: a trainer
// renders one or two resolutions).
struct CountGuess {
  int P = -1, W = 0, H = 0;
  double recent_max = 0;   // decays slowly so that a past outlier does not inflate the workspace for ever
};
CountGuess& guess_for(int P, int W, int H) {
  static thread_local CountGuess table[4];
  static thread_local int next = 0;
  for (auto& g : table)
    if (g.P == P && g.W == W && g.H == H) return g;
  CountGuess& g = table[next];
  next = (next + 1) % 4;
  g = CountGuess{P, W, H, 0};
  return g;
}

void* blob_alloc(void* ctx, size_t bytes) {
  auto* c = static_cast<BlobCtx*>(ctx);
  c->blob = torch::empty({(int64_t)bytes}, torch::TensorOptions().dtype(torch::kUInt8).device(c->dev));
  return c->blob.data_ptr();
}

std::tuple<int, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor,
           torch::Tensor, torch::Tensor>
rasterize_gaussians(const torch::Tensor& background, const torch::Tensor& means3D,
                    const torch::Tensor& colors, const torch::Tensor& opacity,
                    const torch::Tensor& scales, const torch::Tensor& rotations,
                    const float scale_modifier, const torch::Tensor& cov3D_precomp,
                    const torch::Tensor& viewmatrix, const torch::Tensor& projmatrix,
                    const float tan_fovx, const float tan_fovy, const int image_height,
                    const int image_width, const torch::Tensor& sh, const int degree,
                    const torch::Tensor& campos, const bool prefiltered, const bool debug) {
  if (means3D.ndimension() != 2 || means3D.size(1) != 3) {
    AT_ERROR("means3D must have dimensions (num_points, 3)");
  }
  TORCH_CHECK(means3D.is_cuda(), "means3D must be a CUDA tensor (there is no CPU path)");
  const torch::Device dev = means3D.device();
  c10::cuda::CUDAGuard guard(dev);
  const int P = (int)means3D.size(0);
  const int H = image_height, W = image_width;
  auto fopt = means3D.options().dtype(torch::kFloat32);

  // every pixel / every Gaussian is written by the kernels, so no zero-fill (the reference
  // fills all four, rasterize_points.cu:68-71); P == 0 is zeroed inside egs_forward.
  torch::Tensor out_color = torch::empty({EGS_NUM_CHANNELS, H, W}, fopt);
  torch::Tensor out_depth = torch::empty({1, H, W}, fopt);
  torch::Tensor out_alpha = torch::empty({1, H, W}, fopt);
  torch::Tensor radii = torch::empty({P}, means3D.options().dtype(torch::kInt32));

  BlobCtx geom{torch::empty({0}, torch::TensorOptions().dtype(torch::kUInt8).device(dev)), dev};
  BlobCtx binning = geom, image = geom;

  torch::Tensor k[12];
  EgsForwardArgs a{};
  a.P = P;
  a.sh_degree = degree;
  a.sh_coeffs = (sh.defined() && sh.numel() != 0) ? (int)sh.size(1) : 0;
  a.width = W;
  a.height = H;
  a.background = req_ptr(background, k[0], dev, "background");
  a.means3D = P ? req_ptr(means3D, k[1], dev, "means3D") : nullptr;
  a.shs = opt_ptr(sh, k[2], dev, "sh");
  a.colors_precomp = opt_ptr(colors, k[3], dev, "colors_precomp");
  a.opacities = opt_ptr(opacity, k[4], dev, "opacity");
  a.scales = opt_ptr(scales, k[5], dev, "scales");
  a.rotations = opt_ptr(rotations, k[6], dev, "rotations");
  a.cov3D_precomp = opt_ptr(cov3D_precomp, k[7], dev, "cov3D_precomp");
  a.scale_modifier = scale_modifier;
  a.viewmatrix = req_ptr(viewmatrix, k[8], dev, "viewmatrix");
  a.projmatrix = req_ptr(projmatrix, k[9], dev, "projmatrix");
  a.campos = req_ptr(campos, k[10], dev, "campos");
  a.tan_fovx = tan_fovx;
  a.tan_fovy = tan_fovy;
  a.prefiltered = prefiltered;
  a.debug = debug;
  a.out_color = out_color.data_ptr<float>();
  a.out_depth = out_depth.data_ptr<float>();
  a.out_alpha = out_alpha.data_ptr<float>();
  a.radii = radii.data_ptr<int>();
  a.geom_alloc = blob_alloc;    a.geom_ctx = &geom;
  a.binning_alloc = blob_alloc; a.binning_ctx = &binning;
  a.image_alloc = blob_alloc;   a.image_ctx = &image;
  // guess of num_rendered from recent calls with the same problem size: lets egs_forward enqueue
  // everything before the count is read back (it re-runs by itself when the guess is too small)
  CountGuess& guess = guess_for(P, W, H);
  a.num_rendered_hint = guess.recent_max > 0 ? (int)(guess.recent_max * 1.25) + 8192 : 0;

  int rendered = 0;
  check(egs_forward(&a, at::cuda::getCurrentCUDAStream().stream(), &rendered), "egs_forward");
  guess.recent_max = std::max<double>(rendered, guess.recent_max * (63.0 / 64.0));
  return std::make_tuple(rendered, out_color, out_depth, out_alpha, radii, geom.blob, binning.blob,
                         image.blob);
}

std::tuple<torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor,
           torch::Tensor, torch::Tensor>
rasterize_backward_impl(bool colors_only, const torch::Tensor& background, const torch::Tensor& means3D,
                             const torch::Tensor& radii, const torch::Tensor& colors,
                             const torch::Tensor& scales, const torch::Tensor& rotations,
                             const float scale_modifier, const torch::Tensor& cov3D_precomp,
                             const torch::Tensor& viewmatrix, const torch::Tensor& projmatrix,
                             const float tan_fovx, const float tan_fovy,
                             const torch::Tensor& dL_dout_color, const torch::Tensor& dL_dout_depth,
                             const torch::Tensor& dL_dout_alpha, const torch::Tensor& sh,
                             const int degree, const torch::Tensor& campos,
                             const torch::Tensor& geomBuffer, const int R,
                             const torch::Tensor& binningBuffer, const torch::Tensor& imageBuffer,
                             const torch::Tensor& alphas, const bool debug) {
  TORCH_CHECK(means3D.is_cuda(), "means3D must be a CUDA tensor (there is no CPU path)");
  const torch::Device dev = means3D.device();
  c10::cuda::CUDAGuard guard(dev);
  const int P = (int)means3D.size(0);
  const int H = (int)dL_dout_color.size(1);
  const int W = (int)dL_dout_color.size(2);
  const int M = (sh.defined() && sh.numel() != 0) ? (int)sh.size(1) : 0;
  auto fopt = means3D.options().dtype(torch::kFloat32);

  const bool has_sh = M != 0;
  const bool has_colors = colors.defined() && colors.numel() != 0;
  const bool has_cov = cov3D_precomp.defined() && cov3D_precomp.numel() != 0;
  const bool has_sr = scales.defined() && scales.numel() != 0;

  // outputs are fully written by the kernels: empty, not zeros (the reference zero-fills ten
  // tensors, rasterize_points.cu:158-167). Gradients of absent inputs are zero-size tensors.
  TORCH_CHECK(!colors_only || has_colors, "colors-only backward needs colors_precomp");
  const int64_t Pg = colors_only ? 0 : P;   // colors-only: no other gradient is produced
  torch::Tensor dL_dmeans3D = torch::empty({Pg, 3}, fopt);
  torch::Tensor dL_dmeans2D = torch::empty({Pg, 3}, fopt);
  torch::Tensor dL_dopacity = torch::empty({Pg, 1}, fopt);
  torch::Tensor dL_dcolors = torch::empty({has_colors ? P : 0, EGS_NUM_CHANNELS}, fopt);
  torch::Tensor dL_dcov3D = torch::empty({has_cov ? Pg : 0, 6}, fopt);
  torch::Tensor dL_dsh = torch::empty({has_sh ? Pg : 0, M, 3}, fopt);
  torch::Tensor dL_dscales = torch::empty({has_sr ? Pg : 0, 3}, fopt);
  torch::Tensor dL_drotations = torch::empty({has_sr ? Pg : 0, 4}, fopt);

  if (P != 0) {
    torch::Tensor scratch = torch::empty({(int64_t)egs_backward_scratch_bytes(P)},
                                         torch::TensorOptions().dtype(torch::kUInt8).device(dev));
    torch::Tensor k[16];
    EgsBackwardArgs a{};
    a.P = P;
    a.sh_degree = degree;
    a.sh_coeffs = M;
    a.num_rendered = R;
    a.width = W;
    a.height = H;
    a.background = req_ptr(background, k[0], dev, "background");
    a.means3D = req_ptr(means3D, k[1], dev, "means3D");
    a.shs = opt_ptr(sh, k[2], dev, "sh");
    a.colors_precomp = opt_ptr(colors, k[3], dev, "colors_precomp");
    a.scales = opt_ptr(scales, k[4], dev, "scales");
    a.rotations = opt_ptr(rotations, k[5], dev, "rotations");
    a.cov3D_precomp = opt_ptr(cov3D_precomp, k[6], dev, "cov3D_precomp");
    a.scale_modifier = scale_modifier;
    a.viewmatrix = req_ptr(viewmatrix, k[7], dev, "viewmatrix");
    a.projmatrix = req_ptr(projmatrix, k[8], dev, "projmatrix");
    a.campos = req_ptr(campos, k[9], dev, "campos");
    a.tan_fovx = tan_fovx;
    a.tan_fovy = tan_fovy;
    TORCH_CHECK(radii.scalar_type() == torch::kInt32 && radii.device() == dev, "radii must be int32 on ", dev);
    k[10] = radii.contiguous();
    a.radii = k[10].data_ptr<int>();
    a.out_alpha = req_ptr(alphas, k[11], dev, "alphas");
    TORCH_CHECK(geomBuffer.device() == dev && imageBuffer.device() == dev, "workspace blobs on wrong device");
    a.geom_buffer = geomBuffer.data_ptr();
    a.binning_buffer = binningBuffer.numel() ? binningBuffer.data_ptr() : nullptr;
    a.image_buffer = imageBuffer.data_ptr();
    a.dL_dout_color = req_ptr(dL_dout_color, k[12], dev, "dL_dout_color");
    a.dL_dout_depth = opt_ptr(dL_dout_depth, k[13], dev, "dL_dout_depth");   // empty: no gradient flows into it
    a.dL_dout_alpha = opt_ptr(dL_dout_alpha, k[14], dev, "dL_dout_alpha");
    a.scratch = scratch.data_ptr();
    auto out = [](torch::Tensor& t) { return t.numel() ? t.data_ptr<float>() : nullptr; };
    a.dL_dmeans2D = out(dL_dmeans2D);
    a.dL_dcolors = out(dL_dcolors);
    a.dL_dopacity = out(dL_dopacity);
    a.dL_dmeans3D = out(dL_dmeans3D);
    a.dL_dcov3D = out(dL_dcov3D);
    a.dL_dsh = out(dL_dsh);
    a.dL_dscales = out(dL_dscales);
    a.dL_drotations = out(dL_drotations);
    a.debug = debug;
    a.colors_only = colors_only ? 1 : 0;
    a.binning_bytes = (size_t)binningBuffer.numel();   // the blob may be laid out for a guessed capacity
    check(egs_backward(&a, at::cuda::getCurrentCUDAStream().stream()), "egs_backward");
  }
  return std::make_tuple(dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh,
                         dL_dscales, dL_drotations);
}

// the reference's signature (rasterize_points.h:41-65), and the same with only dL_dcolors produced
#define EGS_BWD_PARAMS                                                                                      \
  const torch::Tensor &background, const torch::Tensor &means3D, const torch::Tensor &radii,                \
      const torch::Tensor &colors, const torch::Tensor &scales, const torch::Tensor &rotations,             \
      const float scale_modifier, const torch::Tensor &cov3D_precomp, const torch::Tensor &viewmatrix,      \
      const torch::Tensor &projmatrix, const float tan_fovx, const float tan_fovy,                          \
      const torch::Tensor &dL_dout_color, const torch::Tensor &dL_dout_depth,                               \
      const torch::Tensor &dL_dout_alpha, const torch::Tensor &sh, const int degree,                        \
      const torch::Tensor &campos, const torch::Tensor &geomBuffer, const int R,                            \
      const torch::Tensor &binningBuffer, const torch::Tensor &imageBuffer, const torch::Tensor &alphas,    \
      const bool debug
#define EGS_BWD_ARGS                                                                                        \
  background, means3D, radii, colors, scales, rotations, scale_modifier, cov3D_precomp, viewmatrix,         \
      projmatrix, tan_fovx, tan_fovy, dL_dout_color, dL_dout_depth, dL_dout_alpha, sh, degree, campos,      \
      geomBuffer, R, binningBuffer, imageBuffer, alphas, debug
auto rasterize_gaussians_backward(EGS_BWD_PARAMS) { return rasterize_backward_impl(false, EGS_BWD_ARGS); }
auto rasterize_gaussians_backward_colors(EGS_BWD_PARAMS) { return rasterize_backward_impl(true, EGS_BWD_ARGS); }

torch::Tensor mark_visible(torch::Tensor& means3D, torch::Tensor& viewmatrix, torch::Tensor& projmatrix) {
  TORCH_CHECK(means3D.is_cuda(), "means3D must be a CUDA tensor (there is no CPU path)");
  const torch::Device dev = means3D.device();
  c10::cuda::CUDAGuard guard(dev);
  const int P = (int)means3D.size(0);
  torch::Tensor present = torch::empty({P}, means3D.options().dtype(at::kBool));
  if (P != 0) {
    torch::Tensor k[3];
    check(egs_mark_visible(P, req_ptr(means3D, k[0], dev, "means3D"),
                           req_ptr(viewmatrix, k[1], dev, "viewmatrix"),
                           req_ptr(projmatrix, k[2], dev, "projmatrix"),
                           reinterpret_cast<unsigned char*>(present.data_ptr<bool>()),
                           at::cuda::getCurrentCUDAStream().stream()),
          "egs_mark_visible");
  }
  return present;
}

// ---- optimizer step -------------------------------------------------------------------------
void adam_step(std::vector<torch::Tensor> params, std::vector<torch::Tensor> grads,
               std::vector<torch::Tensor> exp_avgs, std::vector<torch::Tensor> exp_avg_sqs,
               std::vector<double> lrs, std::vector<int64_t> steps, double beta1, double beta2, double eps) {
  const size_t n = params.size();
  TORCH_CHECK(n <= EGS_ADAM_MAX_GROUPS, "at most ", EGS_ADAM_MAX_GROUPS, " groups per call");
  TORCH_CHECK(grads.size() == n && exp_avgs.size() == n && exp_avg_sqs.size() == n && lrs.size() == n && steps.size() == n,
              "adam_step: list lengths differ");
  if (n == 0) return;
  c10::cuda::CUDAGuard guard(params[0].device());
  EgsAdamGroup g[EGS_ADAM_MAX_GROUPS];
  for (size_t i = 0; i < n; ++i) {
    for (const torch::Tensor* t : {&params[i], &grads[i], &exp_avgs[i], &exp_avg_sqs[i]}) {
      TORCH_CHECK(t->is_cuda() && t->scalar_type() == torch::kFloat32 && t->is_contiguous(),
                  "adam_step: tensors must be contiguous float32 CUDA");
      TORCH_CHECK(t->numel() == params[i].numel(), "adam_step: size mismatch in group ", i);
    }
    g[i].param = params[i].data_ptr<float>();
    g[i].grad = grads[i].data_ptr<float>();
    g[i].exp_avg = exp_avgs[i].data_ptr<float>();
    g[i].exp_avg_sq = exp_avg_sqs[i].data_ptr<float>();
    g[i].n = params[i].numel();
    g[i].lr = lrs[i];
    g[i].step = (int)steps[i];
  }
  check(egs_adam_step(g, (int)n, beta1, beta2, eps,
                      at::cuda::getCurrentCUDAStream().stream()),
        "egs_adam_step");
}

std::tuple<torch::Tensor, torch::Tensor> se3_apply(const torch::Tensor& xyz, const torch::Tensor& is_object,
                                                   int which, const torch::Tensor& Rt,
                                                   const torch::Tensor& cov) {
  TORCH_CHECK(xyz.is_cuda() && xyz.scalar_type() == torch::kFloat32, "xyz must be float32 CUDA");
  TORCH_CHECK(is_object.scalar_type() == torch::kInt32 && is_object.is_cuda(), "is_object must be int32 CUDA");
  TORCH_CHECK(Rt.numel() == 12 && Rt.is_cuda() && Rt.scalar_type() == torch::kFloat32, "Rt must be 12 float32 CUDA");
  c10::cuda::CUDAGuard guard(xyz.device());
  auto x = xyz.contiguous();
  auto io = is_object.contiguous();
  auto rt = Rt.contiguous();
  const int P = (int)x.size(0);
  torch::Tensor out = torch::empty_like(x);
  torch::Tensor cov_c, cov_out;
  const bool has_cov = cov.defined() && cov.numel() != 0;
  if (has_cov) {
    cov_c = cov.contiguous();
    cov_out = torch::empty_like(cov_c);
  } else {
    cov_out = torch::empty({0}, x.options());
  }
  check(egs_se3_apply(P, x.data_ptr<float>(), io.data_ptr<int>(), which, rt.data_ptr<float>(),
                      out.data_ptr<float>(), has_cov ? cov_c.data_ptr<float>() : nullptr,
                      has_cov ? cov_out.data_ptr<float>() : nullptr,
                      at::cuda::getCurrentCUDAStream().stream()),
        "egs_se3_apply");
  return std::make_tuple(out, cov_out);
}

void adam_se3_step(torch::Tensor xyz, const torch::Tensor& grad, torch::Tensor exp_avg,
                   torch::Tensor exp_avg_sq, const torch::Tensor& is_object, int which,
                   const torch::Tensor& Rt, double lr, int64_t step, double beta1, double beta2, double eps) {
  for (const torch::Tensor* t : {(const torch::Tensor*)&xyz, &grad, (const torch::Tensor*)&exp_avg, (const torch::Tensor*)&exp_avg_sq})
    TORCH_CHECK(t->is_cuda() && t->scalar_type() == torch::kFloat32 && t->is_contiguous(),
                "adam_se3_step: tensors must be contiguous float32 CUDA");
  TORCH_CHECK(is_object.scalar_type() == torch::kInt32 && is_object.is_cuda() && is_object.is_contiguous(),
              "is_object must be contiguous int32 CUDA");
  c10::cuda::CUDAGuard guard(xyz.device());
  auto rt = Rt.contiguous();
  check(egs_adam_se3_step((int)xyz.size(0), xyz.data_ptr<float>(), grad.data_ptr<float>(),
                          exp_avg.data_ptr<float>(), exp_avg_sq.data_ptr<float>(), is_object.data_ptr<int>(),
                          which, rt.data_ptr<float>(), lr, (int)step, beta1, beta2,
                          eps, at::cuda::getCurrentCUDAStream().stream()),
        "egs_adam_se3_step");
}

// ---- GaussianModel activations (egs_model_forward / egs_model_backward) --------------------------
// Tensors that are absent are passed as empty tensors, like the rasterizer's optional inputs.
struct ModelHold {
  torch::Tensor s, q, o, dc, rest, obj, A, T;
  EgsModelArgs args{};
};

void fill_model(ModelHold& h, const torch::Device& dev, const torch::Tensor& scaling_raw,
                const torch::Tensor& rotation_raw, const torch::Tensor& opacity_raw,
                const torch::Tensor& features_dc, const torch::Tensor& features_rest, double scale_modifier,
                const torch::Tensor& is_object, int64_t which_object, const torch::Tensor& accum_R,
                const torch::Tensor& train_R, int64_t row0_copies, int64_t P) {
  EgsModelArgs& m = h.args;
  m.P = (int)P;
  m.scaling_raw = opt_ptr(scaling_raw, h.s, dev, "scaling_raw");
  m.rotation_raw = opt_ptr(rotation_raw, h.q, dev, "rotation_raw");
  m.opacity_raw = opt_ptr(opacity_raw, h.o, dev, "opacity_raw");
  m.features_dc = opt_ptr(features_dc, h.dc, dev, "features_dc");
  m.features_rest = opt_ptr(features_rest, h.rest, dev, "features_rest");
  m.sh_coeffs = m.features_dc ? 1 + (features_rest.numel() ? (int)features_rest.size(1) : 0) : 0;
  m.scale_modifier = (float)scale_modifier;
  m.is_object = nullptr;
  if (is_object.defined() && is_object.numel() != 0) {
    TORCH_CHECK(is_object.device() == dev && is_object.scalar_type() == torch::kInt32 && is_object.numel() == P,
                "is_object must be int32 [P] or [P,1] on ", dev);
    h.obj = is_object.contiguous();
    m.is_object = h.obj.data_ptr<int>();
  }
  m.which_object = (int)which_object;
  m.accum_R = opt_ptr(accum_R, h.A, dev, "accum_R");
  m.train_R = opt_ptr(train_R, h.T, dev, "train_R");
  TORCH_CHECK((!m.accum_R || accum_R.numel() == 9) && (!m.train_R || train_R.numel() == 9), "rotations must be 3x3");
  m.row0_copies = (int)row0_copies;
}

std::tuple<torch::Tensor, torch::Tensor, torch::Tensor> model_forward(
    const torch::Tensor& scaling_raw, const torch::Tensor& rotation_raw, const torch::Tensor& opacity_raw,
    const torch::Tensor& features_dc, const torch::Tensor& features_rest, double scale_modifier,
    const torch::Tensor& is_object, int64_t which_object, const torch::Tensor& accum_R,
    const torch::Tensor& train_R, int64_t row0_copies) {
  const torch::Tensor& first = scaling_raw.numel() ? scaling_raw : (opacity_raw.numel() ? opacity_raw : features_dc);
  TORCH_CHECK(first.defined() && first.is_cuda(), "model_forward: parameters must be CUDA tensors");
  const torch::Device dev = first.device();
  c10::cuda::CUDAGuard guard(dev);
  const int64_t P = first.size(0);
  ModelHold h;
  fill_model(h, dev, scaling_raw, rotation_raw, opacity_raw, features_dc, features_rest, scale_modifier, is_object,
             which_object, accum_R, train_R, row0_copies, P);
  auto opt = torch::TensorOptions().dtype(torch::kFloat32).device(dev);
  torch::Tensor cov = h.args.scaling_raw ? torch::empty({P, 6}, opt) : torch::empty({0}, opt);
  torch::Tensor opacity = h.args.opacity_raw ? torch::empty({P, 1}, opt) : torch::empty({0}, opt);
  torch::Tensor shs = h.args.features_dc ? torch::empty({P, (int64_t)h.args.sh_coeffs, 3}, opt) : torch::empty({0}, opt);
  check(egs_model_forward(&h.args, cov.numel() ? cov.data_ptr<float>() : nullptr,
                          opacity.numel() ? opacity.data_ptr<float>() : nullptr,
                          shs.numel() ? shs.data_ptr<float>() : nullptr,
                          at::cuda::getCurrentCUDAStream().stream()),
        "egs_model_forward");
  return std::make_tuple(cov, opacity, shs);
}

// returns (d_scaling_raw, d_rotation_raw, d_opacity_raw, d_features_dc, d_features_rest, d_train_R);
// entries whose upstream gradient is absent come back as empty tensors
std::vector<torch::Tensor> model_backward(
    const torch::Tensor& scaling_raw, const torch::Tensor& rotation_raw, const torch::Tensor& opacity_raw,
    int64_t sh_coeffs, double scale_modifier, const torch::Tensor& is_object, int64_t which_object,
    const torch::Tensor& accum_R, const torch::Tensor& train_R, int64_t row0_copies,
    const torch::Tensor& dL_dcov3D, const torch::Tensor& dL_dopacity, const torch::Tensor& dL_dshs) {
  const torch::Tensor& first = dL_dcov3D.numel() ? dL_dcov3D : (dL_dopacity.numel() ? dL_dopacity : dL_dshs);
  TORCH_CHECK(first.defined() && first.is_cuda(), "model_backward: gradients must be CUDA tensors");
  const torch::Device dev = first.device();
  c10::cuda::CUDAGuard guard(dev);
  const int64_t P = first.size(0);
  ModelHold h;
  fill_model(h, dev, scaling_raw, rotation_raw, opacity_raw, torch::Tensor(), torch::Tensor(), scale_modifier,
             is_object, which_object, accum_R, train_R, row0_copies, P);
  h.args.sh_coeffs = (int)sh_coeffs;
  torch::Tensor kc, ko, ks;
  const float* gc = opt_ptr(dL_dcov3D, kc, dev, "dL_dcov3D");
  const float* go = opt_ptr(dL_dopacity, ko, dev, "dL_dopacity");
  const float* gs = opt_ptr(dL_dshs, ks, dev, "dL_dshs");
  auto opt = torch::TensorOptions().dtype(torch::kFloat32).device(dev);
  auto none = [&]() { return torch::empty({0}, opt); };
  torch::Tensor ds = gc ? torch::empty({P, 3}, opt) : none();
  torch::Tensor dq = gc ? torch::empty({P, 4}, opt) : none();
  torch::Tensor dT = (gc && h.args.train_R) ? torch::empty({3, 3}, opt) : none();
  torch::Tensor dop = go ? torch::empty({P, 1}, opt) : none();
  torch::Tensor ddc = gs ? torch::empty({P, 1, 3}, opt) : none();
  torch::Tensor drest = gs ? torch::empty({P, sh_coeffs - 1, 3}, opt) : none();
  auto p = [](torch::Tensor& t) { return t.numel() ? t.data_ptr<float>() : nullptr; };
  check(egs_model_backward(&h.args, gc, go, gs, p(ds), p(dq), p(dop), p(ddc), p(drest), p(dT),
                           at::cuda::getCurrentCUDAStream().stream()),
        "egs_model_backward");
  return {ds, dq, dop, ddc, drest, dT};
}

// ---- rasterizer on raw parameters (egs_forward_raw / egs_backward_raw) -------------------------------
// returns (num_rendered, color, depth, alpha, radii, geomBuffer, binningBuffer, imgBuffer)
std::tuple<int, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor,
           torch::Tensor>
rasterize_gaussians_raw(const torch::Tensor& background, const torch::Tensor& means3D, const torch::Tensor& colors,
                        const torch::Tensor& scaling_raw, const torch::Tensor& rotation_raw,
                        const torch::Tensor& opacity_raw, const torch::Tensor& features_dc,
                        const torch::Tensor& features_rest, double scale_modifier, const torch::Tensor& is_object,
                        int64_t which_object, const torch::Tensor& accum_R, const torch::Tensor& train_R,
                        int64_t row0_copies, const torch::Tensor& viewmatrix, const torch::Tensor& projmatrix,
                        double tan_fovx, double tan_fovy, int64_t image_height, int64_t image_width, int64_t degree,
                        const torch::Tensor& campos, bool prefiltered, bool debug) {
  TORCH_CHECK(means3D.is_cuda() && means3D.ndimension() == 2 && means3D.size(1) == 3,
              "means3D must be a CUDA tensor (num_points, 3); there is no CPU path");
  const torch::Device dev = means3D.device();
  c10::cuda::CUDAGuard guard(dev);
  const int P = (int)means3D.size(0);
  const int H = (int)image_height, W = (int)image_width;
  auto fopt = means3D.options().dtype(torch::kFloat32);
  torch::Tensor out_color = torch::empty({EGS_NUM_CHANNELS, H, W}, fopt);
  torch::Tensor out_depth = torch::empty({1, H, W}, fopt);
  torch::Tensor out_alpha = torch::empty({1, H, W}, fopt);
  torch::Tensor radii = torch::empty({P}, means3D.options().dtype(torch::kInt32));
  BlobCtx geom{torch::empty({0}, torch::TensorOptions().dtype(torch::kUInt8).device(dev)), dev};
  BlobCtx binning = geom, image = geom;

  const bool has_colors = colors.defined() && colors.numel() != 0;
  ModelHold h;
  fill_model(h, dev, scaling_raw, rotation_raw, opacity_raw, has_colors ? torch::Tensor() : features_dc,
             has_colors ? torch::Tensor() : features_rest, scale_modifier, is_object, which_object, accum_R, train_R,
             row0_copies, P);
  torch::Tensor k[6];
  EgsForwardArgs a{};
  a.P = P;
  a.sh_degree = (int)degree;
  a.sh_coeffs = h.args.sh_coeffs;
  a.width = W;
  a.height = H;
  a.background = req_ptr(background, k[0], dev, "background");
  a.means3D = P ? req_ptr(means3D, k[1], dev, "means3D") : nullptr;
  a.colors_precomp = opt_ptr(colors, k[2], dev, "colors_precomp");
  a.scale_modifier = (float)scale_modifier;
  a.viewmatrix = req_ptr(viewmatrix, k[3], dev, "viewmatrix");
  a.projmatrix = req_ptr(projmatrix, k[4], dev, "projmatrix");
  a.campos = req_ptr(campos, k[5], dev, "campos");
  a.tan_fovx = (float)tan_fovx;
  a.tan_fovy = (float)tan_fovy;
  a.prefiltered = prefiltered;
  a.debug = debug;
  a.out_color = out_color.data_ptr<float>();
  a.out_depth = out_depth.data_ptr<float>();
  a.out_alpha = out_alpha.data_ptr<float>();
  a.radii = radii.data_ptr<int>();
  a.geom_alloc = blob_alloc;    a.geom_ctx = &geom;
  a.binning_alloc = blob_alloc; a.binning_ctx = &binning;
  a.image_alloc = blob_alloc;   a.image_ctx = &image;
  CountGuess& guess = guess_for(P, W, H);
  a.num_rendered_hint = guess.recent_max > 0 ? (int)(guess.recent_max * 1.25) + 8192 : 0;
  int rendered = 0;
  check(egs_forward_raw(&a, &h.args, at::cuda::getCurrentCUDAStream().stream(), &rendered), "egs_forward_raw");
  guess.recent_max = std::max<double>(rendered, guess.recent_max * (63.0 / 64.0));
  return std::make_tuple(rendered, out_color, out_depth, out_alpha, radii, geom.blob, binning.blob, image.blob);
}

// returns (dL_dmeans2D, dL_dmeans3D, dL_dcolors, d_scaling_raw, d_rotation_raw, d_opacity_raw,
//          d_features_dc, d_features_rest, d_train_R); absent ones are empty tensors
std::vector<torch::Tensor> rasterize_gaussians_raw_backward(
    const torch::Tensor& background, const torch::Tensor& means3D, const torch::Tensor& radii,
    const torch::Tensor& colors, const torch::Tensor& scaling_raw, const torch::Tensor& rotation_raw,
    const torch::Tensor& opacity_raw, const torch::Tensor& features_dc, const torch::Tensor& features_rest,
    double scale_modifier, const torch::Tensor& is_object, int64_t which_object, const torch::Tensor& accum_R,
    const torch::Tensor& train_R, int64_t row0_copies, const torch::Tensor& viewmatrix,
    const torch::Tensor& projmatrix, double tan_fovx, double tan_fovy, const torch::Tensor& dL_dout_color,
    const torch::Tensor& dL_dout_depth, const torch::Tensor& dL_dout_alpha, int64_t degree,
    const torch::Tensor& campos, const torch::Tensor& geomBuffer, int64_t R, const torch::Tensor& binningBuffer,
    const torch::Tensor& imageBuffer, const torch::Tensor& alphas, bool want_train_R, bool debug) {
  TORCH_CHECK(means3D.is_cuda(), "means3D must be a CUDA tensor (there is no CPU path)");
  const torch::Device dev = means3D.device();
  c10::cuda::CUDAGuard guard(dev);
  const int P = (int)means3D.size(0);
  const int H = (int)dL_dout_color.size(1), W = (int)dL_dout_color.size(2);
  auto fopt = means3D.options().dtype(torch::kFloat32);
  const bool has_colors = colors.defined() && colors.numel() != 0;
  ModelHold h;
  fill_model(h, dev, scaling_raw, rotation_raw, opacity_raw, has_colors ? torch::Tensor() : features_dc,
             has_colors ? torch::Tensor() : features_rest, scale_modifier, is_object, which_object, accum_R, train_R,
             row0_copies, P);
  const int M = h.args.sh_coeffs;
  auto none = [&]() { return torch::empty({0}, fopt); };
  torch::Tensor dL_dmeans3D = torch::empty({P, 3}, fopt), dL_dmeans2D = torch::empty({P, 3}, fopt);
  torch::Tensor dL_dcolors = has_colors ? torch::empty({P, EGS_NUM_CHANNELS}, fopt) : none();
  torch::Tensor ds = torch::empty({P, 3}, fopt), dq = torch::empty({P, 4}, fopt), dop = torch::empty({P, 1}, fopt);
  torch::Tensor ddc = (!has_colors && M > 0) ? torch::empty({P, 1, 3}, fopt) : none();
  torch::Tensor drest = (!has_colors && M > 1) ? torch::empty({P, (int64_t)M - 1, 3}, fopt) : none();
  torch::Tensor dT = (want_train_R && h.args.train_R) ? torch::zeros({3, 3}, fopt) : none();
  if (P != 0) {
    torch::Tensor scratch = torch::empty({(int64_t)egs_backward_scratch_bytes(P)},
                                         torch::TensorOptions().dtype(torch::kUInt8).device(dev));
    torch::Tensor k[12];
    EgsBackwardArgs a{};
    a.P = P;
    a.sh_degree = (int)degree;
    a.sh_coeffs = M;
    a.num_rendered = (int)R;
    a.width = W;
    a.height = H;
    a.background = req_ptr(background, k[0], dev, "background");
    a.means3D = req_ptr(means3D, k[1], dev, "means3D");
    a.colors_precomp = opt_ptr(colors, k[2], dev, "colors_precomp");
    a.scale_modifier = (float)scale_modifier;
    a.viewmatrix = req_ptr(viewmatrix, k[3], dev, "viewmatrix");
    a.projmatrix = req_ptr(projmatrix, k[4], dev, "projmatrix");
    a.campos = req_ptr(campos, k[5], dev, "campos");
    a.tan_fovx = (float)tan_fovx;
    a.tan_fovy = (float)tan_fovy;
    TORCH_CHECK(radii.scalar_type() == torch::kInt32 && radii.device() == dev, "radii must be int32 on ", dev);
    k[6] = radii.contiguous();
    a.radii = k[6].data_ptr<int>();
    a.out_alpha = req_ptr(alphas, k[7], dev, "alphas");
    a.geom_buffer = geomBuffer.data_ptr();
    a.binning_buffer = binningBuffer.numel() ? binningBuffer.data_ptr() : nullptr;
    a.image_buffer = imageBuffer.data_ptr();
    a.dL_dout_color = req_ptr(dL_dout_color, k[8], dev, "dL_dout_color");
    a.dL_dout_depth = opt_ptr(dL_dout_depth, k[9], dev, "dL_dout_depth");
    a.dL_dout_alpha = opt_ptr(dL_dout_alpha, k[10], dev, "dL_dout_alpha");
    a.scratch = scratch.data_ptr();
    a.dL_dmeans2D = dL_dmeans2D.data_ptr<float>();
    a.dL_dmeans3D = dL_dmeans3D.data_ptr<float>();
    a.dL_dcolors = has_colors ? dL_dcolors.data_ptr<float>() : nullptr;
    a.debug = debug;
    a.binning_bytes = (size_t)binningBuffer.numel();
    auto p = [](torch::Tensor& t) { return t.numel() ? t.data_ptr<float>() : nullptr; };
    EgsModelGrads g{p(ds), p(dq), p(dop), p(ddc), p(drest), p(dT)};
    check(egs_backward_raw(&a, &h.args, &g, at::cuda::getCurrentCUDAStream().stream()), "egs_backward_raw");
  }
  return {dL_dmeans2D, dL_dmeans3D, dL_dcolors, ds, dq, dop, ddc, drest, dT};
}

// ---- image loss (egs_image_loss) ---------------------------------------------------------------------
// returns (terms [3] = L1 mean, SSIM mean, loss; d loss / d image, empty when want_grad is false)
std::tuple<torch::Tensor, torch::Tensor> image_loss(const torch::Tensor& image, const torch::Tensor& target,
                                                    const torch::Tensor& grad_mask, double lambda_dssim,
                                                    bool want_grad) {
  TORCH_CHECK(image.is_cuda() && image.dim() == 3 && image.scalar_type() == torch::kFloat32,
              "image_loss: image must be a float32 CUDA tensor [C,H,W]");
  TORCH_CHECK(target.sizes() == image.sizes(), "image_loss: image and target must have the same shape");
  const torch::Device dev = image.device();
  c10::cuda::CUDAGuard guard(dev);
  torch::Tensor ki, kt, km;
  EgsLossArgs a{};
  a.channels = (int)image.size(0);
  a.height = (int)image.size(1);
  a.width = (int)image.size(2);
  a.image = req_ptr(image, ki, dev, "image");
  a.target = req_ptr(target, kt, dev, "target");
  a.grad_mask = nullptr;
  if (grad_mask.defined() && grad_mask.numel() != 0) {
    TORCH_CHECK(grad_mask.device() == dev && grad_mask.numel() == image.size(1) * image.size(2),
                "image_loss: grad_mask must hold H*W elements on ", dev);
    km = grad_mask.to(torch::kInt32).contiguous();
    a.grad_mask = km.data_ptr<int>();
  }
  a.lambda_dssim = (float)lambda_dssim;
  const size_t bytes = egs_image_loss_workspace_bytes(a.channels, a.height, a.width);
  auto f32 = torch::TensorOptions().dtype(torch::kFloat32).device(dev);
  torch::Tensor ws = torch::empty({(int64_t)bytes}, f32.dtype(torch::kByte));
  torch::Tensor terms = torch::empty({3}, f32);
  torch::Tensor grad = want_grad ? torch::empty_like(ki) : torch::empty({0}, f32);
  check(egs_image_loss(&a, terms.data_ptr<float>(), want_grad ? grad.data_ptr<float>() : nullptr, ws.data_ptr(),
                       bytes, at::cuda::getCurrentCUDAStream().stream()),
        "egs_image_loss");
  return std::make_tuple(terms, grad);
}

void densify_stats(const torch::Tensor& radii, const torch::Tensor& dL_dmeans2D, torch::Tensor max_radii2D,
                   torch::Tensor xyz_gradient_accum, torch::Tensor denom) {
  TORCH_CHECK(radii.is_cuda() && radii.scalar_type() == torch::kInt32 && radii.is_contiguous(),
              "densify_stats: radii must be contiguous int32 CUDA");
  const int64_t P = radii.numel();
  for (const torch::Tensor* t : {&dL_dmeans2D, (const torch::Tensor*)&max_radii2D,
                                 (const torch::Tensor*)&xyz_gradient_accum, (const torch::Tensor*)&denom})
    TORCH_CHECK(t->device() == radii.device() && t->scalar_type() == torch::kFloat32 && t->is_contiguous(),
                "densify_stats: tensors must be contiguous float32 on the device of radii");
  TORCH_CHECK(dL_dmeans2D.numel() == 3 * P && max_radii2D.numel() == P && xyz_gradient_accum.numel() == P &&
                  denom.numel() == P, "densify_stats: shape mismatch");
  c10::cuda::CUDAGuard guard(radii.device());
  check(egs_densify_stats((int)P, radii.data_ptr<int>(), dL_dmeans2D.data_ptr<float>(), max_radii2D.data_ptr<float>(),
                          xyz_gradient_accum.data_ptr<float>(), denom.data_ptr<float>(),
                          at::cuda::getCurrentCUDAStream().stream()),
        "egs_densify_stats");
}

}  // namespace

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  m.def("densify_stats", &densify_stats);
  m.def("image_loss", &image_loss);
  m.def("rasterize_gaussians_raw", &rasterize_gaussians_raw);
  m.def("rasterize_gaussians_raw_backward", &rasterize_gaussians_raw_backward);
  m.def("model_forward", &model_forward);
  m.def("model_backward", &model_backward);
  m.def("rasterize_gaussians", &rasterize_gaussians);
  m.def("rasterize_gaussians_backward", &rasterize_gaussians_backward);
  m.def("rasterize_gaussians_backward_colors", &rasterize_gaussians_backward_colors);
  m.def("mark_visible", &mark_visible);
  m.def("adam_step", &adam_step);
  m.def("se3_apply", &se3_apply);
  m.def("adam_se3_step", &adam_se3_step);
  m.def("launch_count", []() { return egs_launch_count(); });
  m.def("abi_version", []() { return egs_abi_version(); });
}
