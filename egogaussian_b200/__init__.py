"""egogaussian_b200 -- B200-native differentiable Gaussian-splatting rasterizer.

Layout (only what the hot path needs):
  csrc/                         CUDA kernels (sm_100a), the C-ABI and the torch binding
  diff_gaussian_rasterization/  host-side mirror of the reference's Python interface
  gaussian_renderer.py          render() mirror; activations / covariance and the raw-parameter rasterizer
  render_helper.py              get_render_label mirror (the trainers' per-iteration label render)
  loss_utils.py                 l1_loss / ssim / image_loss mirrors on the fused loss kernels
  simple_knn.py                 distCUDA2 mirror (init-time mean squared 3-NN distance)
  cabi.py                       ctypes view of include/egs_rasterizer.h (tests, bench)
  parallel.py                   view-data-parallel gradient all-reduce (torch.distributed / NCCL)
  scenes.py                     seeded synthetic scenes and cameras
  build.py                      in-tree sm_100a build
"""
__version__ = "0.1.0"
