"""ctypes view of the C-ABI declared in include/egs_rasterizer.h.

This is the binding a non-torch host would write (INTEGRATION.md shows the same stub). The parity
tests and bench.py call the library through it with raw device pointers, so what they exercise is
exactly the `extern "C"` surface, not the torch extension. torch is used only to own device
memory and to name the stream.

There is no CPU path: `load()` raises if libegs_rasterizer.so is missing.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path
from typing import Optional

_LIB_PATH = Path(__file__).resolve().parent / "libegs_rasterizer.so"
_lib = None

EGS_OK = 0
ALLOC_FN = C.CFUNCTYPE(C.c_void_p, C.c_void_p, C.c_size_t)

c_float_p = C.c_void_p  # device pointers are passed as opaque addresses


class EgsForwardArgs(C.Structure):
    _fields_ = [
        ("P", C.c_int), ("sh_degree", C.c_int), ("sh_coeffs", C.c_int),
        ("width", C.c_int), ("height", C.c_int),
        ("background", c_float_p), ("means3D", c_float_p), ("shs", c_float_p),
        ("colors_precomp", c_float_p), ("opacities", c_float_p), ("scales", c_float_p),
        ("rotations", c_float_p), ("cov3D_precomp", c_float_p), ("scale_modifier", C.c_float),
        ("viewmatrix", c_float_p), ("projmatrix", c_float_p), ("campos", c_float_p),
        ("tan_fovx", C.c_float), ("tan_fovy", C.c_float), ("prefiltered", C.c_int), ("debug", C.c_int),
        ("out_color", c_float_p), ("out_depth", c_float_p), ("out_alpha", c_float_p), ("radii", C.c_void_p),
        ("geom_alloc", ALLOC_FN), ("geom_ctx", C.c_void_p),
        ("binning_alloc", ALLOC_FN), ("binning_ctx", C.c_void_p),
        ("image_alloc", ALLOC_FN), ("image_ctx", C.c_void_p),
        ("num_rendered_hint", C.c_int),
    ]


class EgsBackwardArgs(C.Structure):
    _fields_ = [
        ("P", C.c_int), ("sh_degree", C.c_int), ("sh_coeffs", C.c_int), ("num_rendered", C.c_int),
        ("width", C.c_int), ("height", C.c_int),
        ("background", c_float_p), ("means3D", c_float_p), ("shs", c_float_p),
        ("colors_precomp", c_float_p), ("scales", c_float_p), ("rotations", c_float_p),
        ("cov3D_precomp", c_float_p), ("scale_modifier", C.c_float),
        ("viewmatrix", c_float_p), ("projmatrix", c_float_p), ("campos", c_float_p),
        ("tan_fovx", C.c_float), ("tan_fovy", C.c_float),
        ("radii", C.c_void_p), ("out_alpha", c_float_p),
        ("geom_buffer", C.c_void_p), ("binning_buffer", C.c_void_p), ("image_buffer", C.c_void_p),
        ("dL_dout_color", c_float_p), ("dL_dout_depth", c_float_p), ("dL_dout_alpha", c_float_p),
        ("scratch", C.c_void_p),
        ("dL_dmeans2D", c_float_p), ("dL_dcolors", c_float_p), ("dL_dopacity", c_float_p),
        ("dL_dmeans3D", c_float_p), ("dL_dcov3D", c_float_p), ("dL_dsh", c_float_p),
        ("dL_dscales", c_float_p), ("dL_drotations", c_float_p), ("debug", C.c_int),
        ("binning_bytes", C.c_size_t), ("colors_only", C.c_int), ("dL_drgb", C.c_void_p),
        ("drgb_n_remote", C.c_int), ("drgb_remote", C.c_void_p * 8), ("drgb_multicast", C.c_void_p),
    ]


class EgsAdamGroup(C.Structure):
    _fields_ = [("param", c_float_p), ("grad", c_float_p), ("exp_avg", c_float_p),
                ("exp_avg_sq", c_float_p), ("n", C.c_longlong), ("lr", C.c_double), ("step", C.c_int)]


class EgsModelArgs(C.Structure):
    _fields_ = [("P", C.c_int), ("sh_coeffs", C.c_int),
                ("scaling_raw", c_float_p), ("rotation_raw", c_float_p), ("opacity_raw", c_float_p),
                ("features_dc", c_float_p), ("features_rest", c_float_p), ("scale_modifier", C.c_float),
                ("is_object", C.c_void_p), ("which_object", C.c_int),
                ("accum_R", c_float_p), ("train_R", c_float_p), ("row0_copies", C.c_int)]


class EgsModelGrads(C.Structure):
    _fields_ = [("d_scaling_raw", c_float_p), ("d_rotation_raw", c_float_p), ("d_opacity_raw", c_float_p),
                ("d_features_dc", c_float_p), ("d_features_rest", c_float_p), ("d_train_R", c_float_p)]


class EgsRowTensor(C.Structure):
    _fields_ = [("src", C.c_void_p), ("dst", C.c_void_p), ("row_bytes", C.c_int)]


class EgsDensifySelect(C.Structure):
    _fields_ = [("P", C.c_int), ("n_graded", C.c_int), ("grad_accum", C.c_void_p), ("denom", C.c_void_p),
                ("scaling_raw", C.c_void_p), ("grad_threshold", C.c_float), ("size_threshold", C.c_float),
                ("split", C.c_int), ("is_object", C.c_void_p), ("which_object", C.c_int),
                ("generation", C.c_void_p), ("only_generation", C.c_int)]


class EgsAppendTensor(C.Structure):
    _fields_ = [("src", C.c_void_p), ("dst", C.c_void_p), ("row_bytes", C.c_int), ("fill", C.c_int), ("value", C.c_int)]


class EgsLossArgs(C.Structure):
    _fields_ = [("channels", C.c_int), ("height", C.c_int), ("width", C.c_int),
                ("image", c_float_p), ("target", c_float_p), ("grad_mask", C.c_void_p),
                ("lambda_dssim", C.c_float)]


EGS_MAX_PEERS = 8
EGS_EXCHANGE_SIGNAL_WORDS = 64
EGS_EXCHANGE_MAX_SEGMENTS = 8


class EgsPeerGroup(C.Structure):
    _fields_ = [("world", C.c_int), ("rank", C.c_int), ("grads", C.c_void_p * EGS_MAX_PEERS),
                ("signals", C.c_void_p * EGS_MAX_PEERS), ("grads_multicast", C.c_void_p)]


class EgsShFactors(C.Structure):
    _fields_ = [("P", C.c_int), ("sh_degree", C.c_int), ("sh_coeffs", C.c_int), ("means3D", C.c_void_p),
                ("drgb", C.c_void_p * EGS_MAX_PEERS), ("dL_dsh", C.c_void_p),
                ("n_views", C.c_int), ("view_flat", C.c_void_p * EGS_MAX_PEERS)]


class EgsExchangeSegment(C.Structure):
    _fields_ = [("begin", C.c_longlong), ("end", C.c_longlong), ("lr_a", C.c_double), ("lr_b", C.c_double),
                ("period", C.c_int), ("split", C.c_int)]


class EgsExchangeAdamArgs(C.Structure):
    _fields_ = [("params", C.c_void_p * EGS_MAX_PEERS), ("params_multicast", C.c_void_p),
                ("exp_avg", C.c_void_p), ("exp_avg_sq", C.c_void_p),
                ("segments", C.POINTER(EgsExchangeSegment)), ("n_segments", C.c_int), ("step", C.c_int),
                ("beta1", C.c_double), ("beta2", C.c_double), ("eps", C.c_double)]


class EgsPlyColumns(C.Structure):   # also EgsPlyColumnsOut: same layout, without the consts
    _fields_ = [(n, C.c_void_p) for n in ("xyz", "features_dc", "features_rest", "opacity", "scaling", "rotation",
                                          "label", "generation", "is_object")]


class EgsGeomView(C.Structure):
    _fields_ = [("depths", C.c_void_p), ("means2D", C.c_void_p), ("conic_opacity", C.c_void_p),
                ("rgb_depth", C.c_void_p), ("clamped", C.c_void_p), ("tiles_touched", C.c_void_p),
                ("point_offsets", C.c_void_p)]


class EgsBinningView(C.Structure):
    _fields_ = [("keys_sorted", C.c_void_p), ("values_sorted", C.c_void_p)]


class EgsImageView(C.Structure):
    _fields_ = [("n_contrib", C.c_void_p), ("ranges", C.c_void_p)]


# every symbol include/egs_rasterizer.h declares (the CPU test suite checks they all resolve)
EXPORTED_SYMBOLS = [
    "egs_geom_bytes", "egs_binning_bytes", "egs_binning_capacity", "egs_image_bytes", "egs_forward",
    "egs_backward_scratch_bytes", "egs_backward", "egs_mark_visible", "egs_adam_step",
    "egs_se3_apply", "egs_adam_se3_step", "egs_model_forward", "egs_model_backward",
    "egs_forward_raw", "egs_backward_raw",
    "egs_image_loss_workspace_bytes", "egs_image_loss", "egs_densify_stats",
    "egs_compact_scratch_bytes", "egs_compact_plan", "egs_compact_rows", "egs_knn_workspace_bytes", "egs_knn_mean_dist2", "egs_ply_record_floats", "egs_ply_pack", "egs_ply_unpack",
    "egs_exchange_sum", "egs_exchange_shard_floats", "egs_exchange_adam_step", "egs_geom_view", "egs_binning_view", "egs_image_view",
    "egs_launch_count", "egs_last_error", "egs_abi_version",
    "egs_profile_enable", "egs_profile_read", "egs_profile_stage_name",
]
PROFILE_STAGES = 16


def load() -> C.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not _LIB_PATH.exists():
        raise ImportError(f"{_LIB_PATH} not found: build it with `python -m egogaussian_b200.build` "
                          "(there is no CPU fallback)")
    lib = C.CDLL(str(_LIB_PATH))
    lib.egs_geom_bytes.restype = C.c_size_t
    lib.egs_geom_bytes.argtypes = [C.c_int]
    lib.egs_binning_bytes.restype = C.c_size_t
    lib.egs_binning_bytes.argtypes = [C.c_int, C.c_int, C.c_int]
    lib.egs_binning_capacity.restype = C.c_int
    lib.egs_binning_capacity.argtypes = [C.c_size_t]
    lib.egs_image_bytes.restype = C.c_size_t
    lib.egs_image_bytes.argtypes = [C.c_int, C.c_int]
    lib.egs_backward_scratch_bytes.restype = C.c_size_t
    lib.egs_backward_scratch_bytes.argtypes = [C.c_int]
    lib.egs_forward.restype = C.c_int
    lib.egs_forward.argtypes = [C.POINTER(EgsForwardArgs), C.c_void_p, C.POINTER(C.c_int)]
    lib.egs_backward.restype = C.c_int
    lib.egs_backward.argtypes = [C.POINTER(EgsBackwardArgs), C.c_void_p]
    lib.egs_mark_visible.restype = C.c_int
    lib.egs_mark_visible.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.egs_adam_step.restype = C.c_int
    lib.egs_adam_step.argtypes = [C.POINTER(EgsAdamGroup), C.c_int, C.c_double, C.c_double, C.c_double, C.c_void_p]
    lib.egs_se3_apply.restype = C.c_int
    lib.egs_se3_apply.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p,
                                  C.c_void_p, C.c_void_p, C.c_void_p]
    lib.egs_adam_se3_step.restype = C.c_int
    lib.egs_adam_se3_step.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                      C.c_int, C.c_void_p, C.c_double, C.c_int, C.c_double, C.c_double,
                                      C.c_double, C.c_void_p]
    lib.egs_model_forward.restype = C.c_int
    lib.egs_model_forward.argtypes = [C.POINTER(EgsModelArgs), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.egs_model_backward.restype = C.c_int
    lib.egs_model_backward.argtypes = [C.POINTER(EgsModelArgs)] + [C.c_void_p] * 10
    lib.egs_forward_raw.restype = C.c_int
    lib.egs_forward_raw.argtypes = [C.POINTER(EgsForwardArgs), C.POINTER(EgsModelArgs), C.c_void_p, C.POINTER(C.c_int)]
    lib.egs_backward_raw.restype = C.c_int
    lib.egs_backward_raw.argtypes = [C.POINTER(EgsBackwardArgs), C.POINTER(EgsModelArgs), C.POINTER(EgsModelGrads),
                                     C.c_void_p]
    lib.egs_image_loss_workspace_bytes.restype = C.c_size_t
    lib.egs_image_loss_workspace_bytes.argtypes = [C.c_int, C.c_int, C.c_int]
    lib.egs_image_loss.restype = C.c_int
    lib.egs_image_loss.argtypes = [C.POINTER(EgsLossArgs), C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
    lib.egs_compact_scratch_bytes.restype = C.c_size_t
    lib.egs_compact_scratch_bytes.argtypes = [C.c_int]
    lib.egs_compact_plan.restype = C.c_int
    lib.egs_compact_plan.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_size_t, C.POINTER(C.c_int), C.c_void_p]
    lib.egs_compact_rows.restype = C.c_int
    lib.egs_compact_rows.argtypes = [C.c_int, C.c_void_p, C.POINTER(EgsRowTensor), C.c_int, C.c_void_p]
    lib.egs_densify_select.restype = C.c_int
    lib.egs_densify_select.argtypes = [C.POINTER(EgsDensifySelect), C.c_void_p, C.c_void_p]
    lib.egs_append_rows.restype = C.c_int
    lib.egs_append_rows.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_int, C.POINTER(EgsAppendTensor), C.c_int, C.c_void_p]
    lib.egs_split_samples.restype = C.c_int
    lib.egs_split_samples.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_int] + [C.c_void_p] * 7
    lib.egs_knn_mean_dist2.restype = C.c_int
    lib.egs_knn_mean_dist2.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
    lib.egs_exchange_sum.restype = C.c_int
    lib.egs_exchange_sum.argtypes = [C.POINTER(EgsPeerGroup), C.c_longlong, C.c_float, C.c_uint32, C.c_int, C.c_void_p]
    lib.egs_exchange_sum_factored.restype = C.c_int
    lib.egs_exchange_sum_factored.argtypes = [C.POINTER(EgsPeerGroup), C.c_longlong, C.POINTER(EgsShFactors), C.c_float,
                                              C.c_uint32, C.c_int, C.c_void_p]
    lib.egs_exchange_shard_floats.restype = C.c_longlong
    lib.egs_exchange_shard_floats.argtypes = [C.c_longlong, C.c_int]
    lib.egs_exchange_adam_step.restype = C.c_int
    lib.egs_exchange_adam_step.argtypes = [C.POINTER(EgsPeerGroup), C.POINTER(EgsExchangeAdamArgs), C.c_longlong,
                                           C.c_float, C.c_uint32, C.c_int, C.c_void_p]
    lib.egs_ply_record_floats.restype = C.c_int
    lib.egs_ply_record_floats.argtypes = [C.c_int, C.c_int]
    lib.egs_ply_pack.restype = C.c_int
    lib.egs_ply_pack.argtypes = [C.c_int, C.c_int, C.POINTER(EgsPlyColumns), C.c_int, C.c_void_p, C.c_void_p]
    lib.egs_ply_unpack.restype = C.c_int
    lib.egs_ply_unpack.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_int, C.POINTER(C.c_int), C.POINTER(EgsPlyColumns),
                                   C.c_float, C.c_int, C.c_void_p]
    lib.egs_knn_workspace_bytes.restype = C.c_size_t
    lib.egs_knn_workspace_bytes.argtypes = [C.c_int]
    lib.egs_densify_stats.restype = C.c_int
    lib.egs_densify_stats.argtypes = [C.c_int] + [C.c_void_p] * 6
    lib.egs_geom_view.restype = C.c_int
    lib.egs_geom_view.argtypes = [C.c_void_p, C.c_int, C.POINTER(EgsGeomView)]
    lib.egs_binning_view.restype = C.c_int
    lib.egs_binning_view.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(EgsBinningView)]
    lib.egs_image_view.restype = C.c_int
    lib.egs_image_view.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(EgsImageView)]
    lib.egs_binning_keys.restype = C.c_int
    lib.egs_binning_keys.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_void_p]
    lib.egs_debug_option.restype = C.c_int
    lib.egs_debug_option.argtypes = [C.c_char_p, C.c_char_p]
    lib.egs_exchange_set_timeout_ms.restype = C.c_int
    lib.egs_exchange_set_timeout_ms.argtypes = [C.c_ulonglong]
    lib.egs_launch_count.restype = C.c_ulonglong
    lib.egs_profile_enable.restype = C.c_int
    lib.egs_profile_enable.argtypes = [C.c_int]
    lib.egs_profile_read.restype = C.c_int
    lib.egs_profile_read.argtypes = [C.POINTER(C.c_double), C.POINTER(C.c_ulonglong)]
    lib.egs_profile_stage_name.restype = C.c_char_p
    lib.egs_profile_stage_name.argtypes = [C.c_int]
    lib.egs_last_error.restype = C.c_char_p
    lib.egs_abi_version.restype = C.c_int
    _lib = lib
    return lib


class EgsError(RuntimeError):
    pass


def debug_option(name: str, value: Optional[str] = None) -> None:
    """egs_debug_option: force a binning path / kernel variant for tests and A/B timing (None = default)."""
    _check(load().egs_debug_option(name.encode(), None if value is None else str(value).encode()), "egs_debug_option")


def exchange_set_timeout_ms(ms: int) -> None:
    _check(load().egs_exchange_set_timeout_ms(int(ms)), "egs_exchange_set_timeout_ms")


def _check(rc: int, what: str) -> None:
    if rc != EGS_OK:
        raise EgsError(f"{what} failed ({rc}): {load().egs_last_error().decode()}")


def _ptr(t) -> Optional[int]:
    """Device address of a tensor, or None (NULL) for an absent / empty one."""
    if t is None or t.numel() == 0:
        return None
    assert t.is_cuda and t.is_contiguous(), "C-ABI inputs must be contiguous CUDA tensors"
    return t.data_ptr()


def _view(addr, shape, dtype, device):
    """Zero-copy torch view of device memory at `addr` (owned by a workspace blob)."""
    import numpy as np
    import torch
    n = int(np.prod(shape))
    if n == 0:
        return torch.empty(shape, dtype=dtype, device=device)
    itemsize = torch.empty((), dtype=dtype).element_size()

    class _Holder:  # __cuda_array_interface__ provider
        pass
    h = _Holder()
    typestr = {torch.float32: "<f4", torch.int32: "<i4", torch.uint8: "|u1", torch.int64: "<i8"}[dtype]
    h.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(addr), False),
                                  "version": 2, "strides": None}
    del itemsize
    return torch.as_tensor(h, device=device)


class ForwardResult:
    """Outputs and workspaces of one egs_forward call (keeps the blobs alive for backward)."""

    def __init__(self):
        self.num_rendered = 0
        self.color = self.depth = self.alpha = self.radii = None
        self.geom = self.binning = self.image = None
        self.P = self.W = self.H = 0

    # --- introspection of the opaque blobs (tests only) -------------------------------------
    def geom_view(self):
        import torch
        lib = load()
        v = EgsGeomView()
        _check(lib.egs_geom_view(self.geom.data_ptr(), self.P, C.byref(v)), "egs_geom_view")
        d, P = self.geom.device, self.P
        return {
            "depths": _view(v.depths, (P,), torch.float32, d),
            "means2D": _view(v.means2D, (P, 2), torch.float32, d),
            "conic_opacity": _view(v.conic_opacity, (P, 4), torch.float32, d),
            "rgb_depth": _view(v.rgb_depth, (P, 4), torch.float32, d),
            "clamped": _view(v.clamped, (P,), torch.uint8, d),
            "tiles_touched": _view(v.tiles_touched, (P,), torch.int32, d),
            "point_offsets": _view(v.point_offsets, (P,), torch.int32, d),
        }

    def binning_view(self):
        import torch
        lib = load()
        R = self.num_rendered
        if R == 0:
            e = torch.empty((0,), dtype=torch.int64, device=self.geom.device)
            return {"keys": e, "values": e.to(torch.int32)}
        v = EgsBinningView()
        cap = lib.egs_binning_capacity(int(self.binning.numel()))   # the blob may be laid out for a guessed capacity
        # forward does not write the sorted keys (nothing on the path reads them): rebuild them
        with torch.cuda.device(self.binning.device):
            _check(lib.egs_binning_keys(self.geom.data_ptr(), self.P, self.binning.data_ptr(), cap,
                                        self.image.data_ptr(), self.W, self.H, _stream_handle()), "egs_binning_keys")
        _check(lib.egs_binning_view(self.binning.data_ptr(), cap, self.W, self.H, C.byref(v)), "egs_binning_view")
        d = self.binning.device
        return {"keys": _view(v.keys_sorted, (R,), torch.int64, d),
                "values": _view(v.values_sorted, (R,), torch.int32, d)}

    def image_view(self):
        import torch
        lib = load()
        v = EgsImageView()
        _check(lib.egs_image_view(self.image.data_ptr(), self.W, self.H, C.byref(v)), "egs_image_view")
        d = self.image.device
        tiles = ((self.W + 15) // 16) * ((self.H + 15) // 16)
        return {"n_contrib": _view(v.n_contrib, (self.H, self.W), torch.int32, d),
                "ranges": _view(v.ranges, (tiles, 2), torch.int32, d)}


def _stream_handle(stream=None) -> int:
    import torch
    s = stream if stream is not None else torch.cuda.current_stream()
    return s.cuda_stream


def forward(*, means3D, opacities, bg, viewmatrix, projmatrix, campos, tanfovx, tanfovy,
            image_height, image_width, sh_degree=0, shs=None, colors_precomp=None, scales=None,
            rotations=None, cov3D_precomp=None, scale_modifier=1.0, prefiltered=False, debug=False,
            stream=None, out=None, num_rendered_hint=0) -> ForwardResult:
    """egs_forward on torch-owned device memory. `out` may carry preallocated output tensors;
    `num_rendered_hint` > 0 lets the library enqueue everything before the instance count is read
    back (results never depend on it, see the header)."""
    import torch
    lib = load()
    dev = means3D.device
    P = int(means3D.shape[0])
    H, W = int(image_height), int(image_width)
    res = ForwardResult()
    res.P, res.W, res.H = P, W, H
    if out is None:
        res.color = torch.empty((3, H, W), dtype=torch.float32, device=dev)
        res.depth = torch.empty((1, H, W), dtype=torch.float32, device=dev)
        res.alpha = torch.empty((1, H, W), dtype=torch.float32, device=dev)
        res.radii = torch.empty((P,), dtype=torch.int32, device=dev)
    else:
        res.color, res.depth, res.alpha, res.radii = out

    blobs = {}

    def make_alloc(name):
        def alloc(_ctx, nbytes):
            blobs[name] = torch.empty((int(nbytes),), dtype=torch.uint8, device=dev)
            return blobs[name].data_ptr()
        return ALLOC_FN(alloc)

    cb = {n: make_alloc(n) for n in ("geom", "binning", "image")}
    a = EgsForwardArgs()
    a.P, a.sh_degree, a.width, a.height = P, int(sh_degree), W, H
    a.sh_coeffs = int(shs.shape[1]) if shs is not None and shs.numel() else 0
    a.background = _ptr(bg)
    a.means3D = _ptr(means3D)
    a.shs = _ptr(shs)
    a.colors_precomp = _ptr(colors_precomp)
    a.opacities = _ptr(opacities)
    a.scales = _ptr(scales)
    a.rotations = _ptr(rotations)
    a.cov3D_precomp = _ptr(cov3D_precomp)
    a.scale_modifier = float(scale_modifier)
    a.viewmatrix, a.projmatrix, a.campos = _ptr(viewmatrix), _ptr(projmatrix), _ptr(campos)
    a.tan_fovx, a.tan_fovy = float(tanfovx), float(tanfovy)
    a.prefiltered, a.debug = int(prefiltered), int(debug)
    a.out_color, a.out_depth, a.out_alpha = res.color.data_ptr(), res.depth.data_ptr(), res.alpha.data_ptr()
    a.radii = res.radii.data_ptr() if P else None
    a.geom_alloc, a.binning_alloc, a.image_alloc = cb["geom"], cb["binning"], cb["image"]
    a.num_rendered_hint = int(num_rendered_hint)
    n = C.c_int(0)
    with torch.cuda.device(dev):
        _check(lib.egs_forward(C.byref(a), _stream_handle(stream), C.byref(n)), "egs_forward")
    res.num_rendered = n.value
    res.geom, res.binning, res.image = blobs.get("geom"), blobs.get("binning"), blobs.get("image")
    return res


def backward(fwd: ForwardResult, *, means3D, bg, viewmatrix, projmatrix, campos, tanfovx,
             tanfovy, dL_dcolor, dL_ddepth, dL_dalpha, sh_degree=0, shs=None, colors_precomp=None,
             scales=None, rotations=None, cov3D_precomp=None, scale_modifier=1.0, debug=False,
             stream=None, scratch=None, out=None, factored=False, rgb_remote=(), rgb_multicast=0):
    """egs_backward; returns a dict of gradient tensors (absent inputs -> None). `out` may carry
    preallocated gradient tensors under the same keys (e.g. views of one flat all-reduce buffer).
    `factored` (or an "rgb" entry in `out`): the SH gradient comes back as "rgb", a slot of
    drgb_slot_floats(P) floats holding the clamp-masked dL/dRGB rows [P,3] and the camera centre
    (EgsBackwardArgs.dL_drgb), and "sh" is not written. `rgb_remote` / `rgb_multicast`: addresses of the
    same slot in other ranks' gather buffers, written by the same kernel."""
    import torch
    lib = load()
    dev = means3D.device
    P, W, H = fwd.P, fwd.W, fwd.H
    M = int(shs.shape[1]) if shs is not None and shs.numel() else 0
    f = dict(dtype=torch.float32, device=dev)
    g = {
        "means2D": torch.empty((P, 3), **f), "opacity": torch.empty((P, 1), **f),
        "means3D": torch.empty((P, 3), **f),
        "colors": torch.empty((P, 3), **f) if colors_precomp is not None else None,
        "cov3D": torch.empty((P, 6), **f) if cov3D_precomp is not None else None,
        "sh": torch.empty((P, M, 3), **f) if M and not factored else None,
        "rgb": torch.empty((drgb_slot_floats(P),), **f) if M and factored else None,
        "scales": torch.empty((P, 3), **f) if scales is not None else None,
        "rotations": torch.empty((P, 4), **f) if rotations is not None else None,
    } if out is None else dict(out)
    for key in ("colors", "cov3D", "sh", "rgb", "scales", "rotations"):
        g.setdefault(key, None)
    if P == 0:
        return g
    if scratch is None:
        scratch = torch.empty((lib.egs_backward_scratch_bytes(P),), dtype=torch.uint8, device=dev)
    a = EgsBackwardArgs()
    a.P, a.sh_degree, a.sh_coeffs, a.num_rendered, a.width, a.height = P, int(sh_degree), M, fwd.num_rendered, W, H
    a.background, a.means3D, a.shs = _ptr(bg), _ptr(means3D), _ptr(shs)
    a.colors_precomp, a.scales, a.rotations = _ptr(colors_precomp), _ptr(scales), _ptr(rotations)
    a.cov3D_precomp = _ptr(cov3D_precomp)
    a.scale_modifier = float(scale_modifier)
    a.viewmatrix, a.projmatrix, a.campos = _ptr(viewmatrix), _ptr(projmatrix), _ptr(campos)
    a.tan_fovx, a.tan_fovy = float(tanfovx), float(tanfovy)
    a.radii, a.out_alpha = fwd.radii.data_ptr(), fwd.alpha.data_ptr()
    a.geom_buffer = fwd.geom.data_ptr()
    a.binning_buffer = fwd.binning.data_ptr() if fwd.binning is not None and fwd.num_rendered else None
    a.binning_bytes = int(fwd.binning.numel()) if fwd.binning is not None else 0
    a.image_buffer = fwd.image.data_ptr()
    a.dL_dout_color, a.dL_dout_depth, a.dL_dout_alpha = _ptr(dL_dcolor), _ptr(dL_ddepth), _ptr(dL_dalpha)
    a.scratch = scratch.data_ptr()
    a.dL_dmeans2D, a.dL_dopacity, a.dL_dmeans3D = g["means2D"].data_ptr(), g["opacity"].data_ptr(), g["means3D"].data_ptr()
    for key, field in (("colors", "dL_dcolors"), ("cov3D", "dL_dcov3D"), ("sh", "dL_dsh"), ("rgb", "dL_drgb"),
                       ("scales", "dL_dscales"), ("rotations", "dL_drotations")):
        setattr(a, field, g[key].data_ptr() if g[key] is not None else None)
    a.drgb_n_remote = len(rgb_remote)
    for q, ptr in enumerate(rgb_remote):
        a.drgb_remote[q] = int(ptr)
    a.drgb_multicast = int(rgb_multicast) or None
    a.debug = int(debug)
    with torch.cuda.device(dev):
        _check(lib.egs_backward(C.byref(a), _stream_handle(stream)), "egs_backward")
    return g


STATIC_SCRATCH_MAX_GAUSSIANS = 262144   # EGS_STATIC_SCRATCH_MAX_GAUSSIANS in the header


def backward_static_scratch(fwd: ForwardResult, *, means3D, bg, viewmatrix, projmatrix, campos, tanfovx,
                            tanfovy, dL_dcolor, dL_ddepth, dL_dalpha, out, sh_degree=0, shs=None,
                            colors_precomp=None, scales=None, rotations=None, cov3D_precomp=None,
                            scale_modifier=1.0, debug=False, stream=None):
    """egs_backward_static_scratch: backward for callers that must not allocate on the device.

    Meant for hosts whose GPU memory belongs to a pool they do not control (a serving process that
    reserved the device up front, a plugin inside one): they pass only buffers they already own.
    `out` must therefore hold ALL gradient tensors, preallocated by the caller, under the keys
    cabi.backward returns ("means2D", "opacity", "means3D" and, for the inputs that are present,
    "colors", "cov3D", "sh", "scales", "rotations"); this wrapper creates no tensor and passes no
    scratch pointer -- the accumulator records are an array inside libegs_rasterizer.so, limited
    to STATIC_SCRATCH_MAX_GAUSSIANS Gaussians and shared by every call on the device, so calls
    must not overlap on the GPU. The first call on a device does a one-time address look-up and
    should happen before any stream capture begins."""
    import torch
    lib = load()
    fn = lib.egs_backward_static_scratch
    fn.restype = C.c_int
    fn.argtypes = [C.POINTER(EgsBackwardArgs), C.c_void_p]
    P, W, H = fwd.P, fwd.W, fwd.H
    M = int(shs.shape[1]) if shs is not None and shs.numel() else 0
    a = EgsBackwardArgs()
    a.P, a.sh_degree, a.sh_coeffs, a.num_rendered, a.width, a.height = P, int(sh_degree), M, fwd.num_rendered, W, H
    a.background, a.means3D, a.shs = _ptr(bg), _ptr(means3D), _ptr(shs)
    a.colors_precomp, a.scales, a.rotations = _ptr(colors_precomp), _ptr(scales), _ptr(rotations)
    a.cov3D_precomp = _ptr(cov3D_precomp)
    a.scale_modifier = float(scale_modifier)
    a.viewmatrix, a.projmatrix, a.campos = _ptr(viewmatrix), _ptr(projmatrix), _ptr(campos)
    a.tan_fovx, a.tan_fovy = float(tanfovx), float(tanfovy)
    a.radii = fwd.radii.data_ptr() if P else None
    a.out_alpha = fwd.alpha.data_ptr()
    a.geom_buffer = fwd.geom.data_ptr() if fwd.geom is not None else None
    a.binning_buffer = fwd.binning.data_ptr() if fwd.binning is not None and fwd.num_rendered else None
    a.binning_bytes = int(fwd.binning.numel()) if fwd.binning is not None else 0
    a.image_buffer = fwd.image.data_ptr() if fwd.image is not None else None
    a.dL_dout_color, a.dL_dout_depth, a.dL_dout_alpha = _ptr(dL_dcolor), _ptr(dL_ddepth), _ptr(dL_dalpha)
    a.scratch = None   # ignored by this entry point
    for key, field in (("means2D", "dL_dmeans2D"), ("opacity", "dL_dopacity"), ("means3D", "dL_dmeans3D"),
                       ("colors", "dL_dcolors"), ("cov3D", "dL_dcov3D"), ("sh", "dL_dsh"),
                       ("scales", "dL_dscales"), ("rotations", "dL_drotations")):
        t = out.get(key)
        setattr(a, field, t.data_ptr() if t is not None and t.numel() else None)
    a.debug = int(debug)
    with torch.cuda.device(means3D.device):
        _check(fn(C.byref(a), _stream_handle(stream)), "egs_backward_static_scratch")
    return out


def mark_visible(means3D, viewmatrix, projmatrix, stream=None):
    import torch
    lib = load()
    P = int(means3D.shape[0])
    out = torch.empty((P,), dtype=torch.bool, device=means3D.device)
    if P:
        with torch.cuda.device(means3D.device):
            _check(lib.egs_mark_visible(P, _ptr(means3D), _ptr(viewmatrix), _ptr(projmatrix), out.data_ptr(),
                                        _stream_handle(stream)), "egs_mark_visible")
    return out


def _model_args(*, scaling_raw=None, rotation_raw=None, opacity_raw=None, features_dc=None, features_rest=None,
                scale_modifier=1.0, is_object=None, which_object=None, accum_R=None, train_R=None,
                row0_copies=0) -> EgsModelArgs:
    m = EgsModelArgs()
    first = next(t for t in (scaling_raw, opacity_raw, features_dc) if t is not None)
    m.P = int(first.shape[0])
    m.sh_coeffs = 0 if features_dc is None else 1 + (0 if features_rest is None else int(features_rest.shape[1]))
    m.scaling_raw, m.rotation_raw, m.opacity_raw = _ptr(scaling_raw), _ptr(rotation_raw), _ptr(opacity_raw)
    m.features_dc, m.features_rest = _ptr(features_dc), _ptr(features_rest)
    m.scale_modifier = float(scale_modifier)
    # which_object=None selects every Gaussian (gaussian_model.py:50-51): no mask at all
    m.is_object = _ptr(is_object) if which_object is not None else None
    m.which_object = int(which_object) if which_object is not None else 0
    m.accum_R, m.train_R = _ptr(accum_R), _ptr(train_R)
    m.row0_copies = int(row0_copies)
    return m


def model_forward(*, want=("cov3D", "opacity", "shs"), stream=None, **model):
    """egs_model_forward: raw GaussianModel parameters -> {cov3D [P,6], opacity [P,1], shs [P,M,3]}."""
    import torch
    lib = load()
    m = _model_args(**model)
    dev = next(t for t in model.values() if hasattr(t, "device")).device
    out = {}
    if "cov3D" in want:
        out["cov3D"] = torch.empty((m.P, 6), dtype=torch.float32, device=dev)
    if "opacity" in want:
        out["opacity"] = torch.empty((m.P, 1), dtype=torch.float32, device=dev)
    if "shs" in want:
        out["shs"] = torch.empty((m.P, m.sh_coeffs, 3), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        _check(lib.egs_model_forward(C.byref(m), _ptr(out.get("cov3D")), _ptr(out.get("opacity")),
                                     _ptr(out.get("shs")), _stream_handle(stream)), "egs_model_forward")
    return out


def model_backward(*, dL_dcov3D=None, dL_dopacity=None, dL_dshs=None, stream=None, **model):
    """egs_model_backward: gradients w.r.t. the raw parameters (and train_R when it is given)."""
    import torch
    lib = load()
    m = _model_args(**model)
    dev = next(t for t in model.values() if hasattr(t, "device")).device
    out = {}
    if dL_dcov3D is not None:
        out["scaling_raw"] = torch.empty((m.P, 3), dtype=torch.float32, device=dev)
        out["rotation_raw"] = torch.empty((m.P, 4), dtype=torch.float32, device=dev)
        if model.get("train_R") is not None:
            out["train_R"] = torch.empty((3, 3), dtype=torch.float32, device=dev)
    if dL_dopacity is not None:
        out["opacity_raw"] = torch.empty((m.P, 1), dtype=torch.float32, device=dev)
    if dL_dshs is not None:
        out["features_dc"] = torch.empty((m.P, 1, 3), dtype=torch.float32, device=dev)
        out["features_rest"] = torch.empty((m.P, m.sh_coeffs - 1, 3), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        _check(lib.egs_model_backward(C.byref(m), _ptr(dL_dcov3D), _ptr(dL_dopacity), _ptr(dL_dshs),
                                      _ptr(out.get("scaling_raw")), _ptr(out.get("rotation_raw")),
                                      _ptr(out.get("opacity_raw")), _ptr(out.get("features_dc")),
                                      _ptr(out.get("features_rest")), _ptr(out.get("train_R")),
                                      _stream_handle(stream)), "egs_model_backward")
    return out


def image_loss(image, target, lambda_dssim, grad_mask=None, want_grad=True, stream=None):
    """egs_image_loss: returns (terms [3] = L1 mean, SSIM mean, loss; dL/dimage or None)."""
    import torch
    lib = load()
    a = EgsLossArgs()
    a.channels, a.height, a.width = (int(v) for v in image.shape)
    a.image, a.target = _ptr(image), _ptr(target)
    a.grad_mask = _ptr(grad_mask)
    a.lambda_dssim = float(lambda_dssim)
    nbytes = lib.egs_image_loss_workspace_bytes(a.channels, a.height, a.width)
    ws = torch.empty((nbytes,), dtype=torch.uint8, device=image.device)
    terms = torch.empty((3,), dtype=torch.float32, device=image.device)
    grad = torch.empty_like(image) if want_grad else None
    with torch.cuda.device(image.device):
        _check(lib.egs_image_loss(C.byref(a), terms.data_ptr(), _ptr(grad), ws.data_ptr(), nbytes,
                                  _stream_handle(stream)), "egs_image_loss")
    return terms, grad


def compact_rows(keep, tensors, stream=None):
    """egs_compact_plan + egs_compact_rows: `[t[keep] for t in tensors]` for contiguous tensors whose
    first dimension is len(keep) and whose rows are a multiple of 4 bytes; one scan, one gather."""
    import torch
    lib = load()
    P = int(keep.shape[0])
    keep_u8 = keep.to(torch.uint8).contiguous() if keep.dtype != torch.uint8 else keep.contiguous()
    dev = keep.device
    nbytes = lib.egs_compact_scratch_bytes(P)
    scratch = torch.empty((nbytes,), dtype=torch.uint8, device=dev)
    count = C.c_int()
    with torch.cuda.device(dev):
        _check(lib.egs_compact_plan(P, _ptr(keep_u8), scratch.data_ptr(), nbytes, C.byref(count),
                                    _stream_handle(stream)), "egs_compact_plan")
        n = count.value
        outs = [torch.empty((n,) + tuple(t.shape[1:]), dtype=t.dtype, device=dev) for t in tensors]
        work = [(t, o) for t, o in zip(tensors, outs) if t.numel()]   # zero-width rows (f_rest at SH degree 0)
        for first in range(0, len(work), 32):
            chunk = work[first:first + 32]
            arr = (EgsRowTensor * len(chunk))()
            for k, (t, o) in enumerate(chunk):
                assert t.is_contiguous() and t.shape[0] == P
                row_bytes = t.element_size() * (t.numel() // max(P, 1))
                arr[k].src, arr[k].dst, arr[k].row_bytes = t.data_ptr(), o.data_ptr() if n else None, row_bytes
            if n and P:
                _check(lib.egs_compact_rows(P, scratch.data_ptr(), arr, len(chunk), _stream_handle(stream)),
                       "egs_compact_rows")
    return outs


def densify_select(*, grad_accum, denom, scaling_raw, grad_threshold, size_threshold, split, is_object=None,
                   which_object=None, generation=None, only_generation=None, stream=None):
    """egs_densify_select: uint8 [P] mask of the rows densify_and_clone (split=False) / densify_and_split
    (split=True) would pick. grad_accum / denom may be shorter than scaling_raw (rows appended since)."""
    import torch
    P = int(scaling_raw.shape[0])
    a = EgsDensifySelect()
    a.P, a.n_graded = P, int(grad_accum.shape[0])
    a.grad_accum, a.denom, a.scaling_raw = _ptr(grad_accum), _ptr(denom), _ptr(scaling_raw)
    a.grad_threshold, a.size_threshold, a.split = float(grad_threshold), float(size_threshold), int(bool(split))
    a.is_object = _ptr(is_object) if which_object is not None else None
    a.which_object = int(which_object) if which_object is not None else 0
    a.generation = _ptr(generation) if only_generation is not None else None
    a.only_generation = int(only_generation) if only_generation is not None else 0
    out = torch.empty((P,), dtype=torch.uint8, device=scaling_raw.device)
    if P:
        with torch.cuda.device(out.device):
            _check(load().egs_densify_select(C.byref(a), out.data_ptr(), _stream_handle(stream)), "egs_densify_select")
    return out


def append_selected_rows(selected, tensors, fills, copies=1, stream=None):
    """egs_compact_plan + egs_append_rows: for every tensor t (first dimension P = len(selected)),
    torch.cat((t, t[selected].repeat(copies, ...))) -- or zeros / a constant in the new rows: `fills[i]` is
    "copy", "zero" or an int (4-byte rows only). Returns (outputs, count, plan) with `plan` reusable by
    split_samples."""
    import torch
    lib = load()
    P = int(selected.shape[0])
    dev = selected.device
    sel = selected.to(torch.uint8).contiguous() if selected.dtype != torch.uint8 else selected.contiguous()
    nbytes = lib.egs_compact_scratch_bytes(P)
    plan = torch.empty((nbytes,), dtype=torch.uint8, device=dev)
    count = C.c_int()
    outs = []
    with torch.cuda.device(dev):
        _check(lib.egs_compact_plan(P, _ptr(sel), plan.data_ptr(), nbytes, C.byref(count), _stream_handle(stream)),
               "egs_compact_plan")
        n = count.value
        outs = [torch.empty((P + copies * n,) + tuple(t.shape[1:]), dtype=t.dtype, device=dev) for t in tensors]
        work = [(t, o, f) for t, o, f in zip(tensors, outs, fills) if o.numel()]   # zero-width rows (f_rest at SH degree 0)
        for first in range(0, len(work), 32):
            chunk = work[first:first + 32]
            arr = (EgsAppendTensor * len(chunk))()
            for k, (t, o, f) in enumerate(chunk):
                assert t.is_contiguous() and t.shape[0] == P
                arr[k].src, arr[k].dst = _ptr(t), o.data_ptr() if o.numel() else None
                arr[k].row_bytes = t.element_size() * (o.numel() // max(o.shape[0], 1))
                arr[k].fill = 0 if f == "copy" else 1 if f == "zero" else 2
                arr[k].value = int(f) if arr[k].fill == 2 else 0
            if P + copies * n:
                _check(lib.egs_append_rows(P, plan.data_ptr(), n, copies, arr, len(chunk), _stream_handle(stream)),
                       "egs_append_rows")
    return outs, n, plan


def split_samples(P, plan, count, copies, xyz, scaling_raw, rotation_raw, noise, xyz_new, scaling_new, stream=None):
    """egs_split_samples: positions and log-scales of the rows a split adds, written into xyz_new /
    scaling_new ([copies * count, 3] views, e.g. the tails of the appended tensors)."""
    import torch
    if count == 0:
        return
    assert xyz_new.is_contiguous() and scaling_new.is_contiguous() and noise.is_contiguous()
    with torch.cuda.device(xyz.device):
        _check(load().egs_split_samples(P, plan.data_ptr(), count, copies, _ptr(xyz), _ptr(scaling_raw), _ptr(rotation_raw),
                                        _ptr(noise), xyz_new.data_ptr(), scaling_new.data_ptr(), _stream_handle(stream)),
               "egs_split_samples")


def knn_mean_dist2(points, stream=None, exhaustive=False):
    """egs_knn_mean_dist2: [P] mean squared distance to the three nearest neighbours (distCUDA2).
    `exhaustive` withholds the workspace, which forces the all-pairs sweep at any size."""
    import torch
    P = int(points.shape[0])
    pts = points.contiguous().float()
    out = torch.empty((P,), dtype=torch.float32, device=points.device)
    if P:
        lib = load()
        nbytes = 0 if exhaustive else int(lib.egs_knn_workspace_bytes(P))
        with torch.cuda.device(points.device):
            ws = torch.empty((nbytes,), dtype=torch.uint8, device=points.device) if nbytes else None
            _check(lib.egs_knn_mean_dist2(P, _ptr(pts), out.data_ptr(), ws.data_ptr() if nbytes else None, nbytes,
                                          _stream_handle(stream)), "egs_knn_mean_dist2")
    return out


def peer_group(world: int, rank: int, grad_ptrs, signal_ptrs, grads_multicast: int = 0) -> EgsPeerGroup:
    """EgsPeerGroup from the per-rank device addresses (ints) of the gradient buffers and signal words."""
    g = EgsPeerGroup()
    g.world, g.rank = world, rank
    g.grads_multicast = int(grads_multicast) or None
    for q in range(world):
        g.grads[q] = int(grad_ptrs[q])
        g.signals[q] = int(signal_ptrs[q])
    return g


def exchange_sum(group: EgsPeerGroup, n_floats: int, epoch: int, scale: float = 1.0, max_ctas: int = 0,
                 stream=None) -> None:
    """egs_exchange_sum: every rank's flat gradient buffer <- scale * the sum over ranks."""
    _check(load().egs_exchange_sum(C.byref(group), n_floats, scale, epoch, max_ctas, _stream_handle(stream)),
           "egs_exchange_sum")


def drgb_slot_floats(P: int) -> int:
    """EGS_DRGB_SLOT_FLOATS: [P,3] rows padded to a multiple of 4, then the camera centre and a spare float."""
    return (3 * P + 3) // 4 * 4 + 4


def exchange_sum_factored(group: EgsPeerGroup, n_floats: int, epoch: int, *, means3D, drgb_ptrs, dL_dsh,
                          sh_degree: int, scale: float = 1.0, max_ctas: int = 0, stream=None, view_flat_ptrs=None) -> None:
    """egs_exchange_sum_factored: the flat buffer of the small gradients is summed over the ranks, and
    dL_dsh [P,M,3] (local) <- sum over views of basis(view's camera) x view's dL/dRGB, read from the
    slots `drgb_ptrs` in this rank's memory. With `view_flat_ptrs` the sums run over a batch of views
    (one entry per view in both lists)."""
    f = EgsShFactors()
    f.P, f.sh_degree, f.sh_coeffs = int(dL_dsh.shape[0]), int(sh_degree), int(dL_dsh.shape[1])
    f.means3D, f.dL_dsh = _ptr(means3D), dL_dsh.data_ptr()
    n = group.world if view_flat_ptrs is None else len(view_flat_ptrs)
    f.n_views = 0 if view_flat_ptrs is None else n
    for q in range(n):
        if view_flat_ptrs is not None:
            f.view_flat[q] = int(view_flat_ptrs[q])
        f.drgb[q] = int(drgb_ptrs[q])
    _check(load().egs_exchange_sum_factored(C.byref(group), n_floats, C.byref(f), scale, epoch, max_ctas,
                                            _stream_handle(stream)), "egs_exchange_sum_factored")


def exchange_adam_step(group: EgsPeerGroup, param_ptrs, exp_avg, exp_avg_sq, segments, n_floats: int, epoch: int,
                       step: int, betas=(0.9, 0.999), eps: float = 1e-15, scale: float = 1.0, max_ctas: int = 0,
                       stream=None, params_multicast: int = 0) -> None:
    """egs_exchange_adam_step. `segments`: [(begin, end, lr_a, lr_b, period, split)] in floats;
    exp_avg / exp_avg_sq: this rank's moment shards (egs_exchange_shard_floats floats each)."""
    a = EgsExchangeAdamArgs()
    for q in range(group.world):
        a.params[q] = int(param_ptrs[q])
    a.params_multicast = int(params_multicast) or None
    a.exp_avg, a.exp_avg_sq = exp_avg.data_ptr(), exp_avg_sq.data_ptr()
    segs = (EgsExchangeSegment * max(len(segments), 1))()
    for k, (b, e, la, lb, per, sp) in enumerate(segments):
        segs[k] = EgsExchangeSegment(b, e, la, lb, per, sp)
    a.segments, a.n_segments, a.step = segs, len(segments), step
    a.beta1, a.beta2, a.eps = betas[0], betas[1], eps
    _check(load().egs_exchange_adam_step(C.byref(group), C.byref(a), n_floats, scale, epoch, max_ctas,
                                         _stream_handle(stream)), "egs_exchange_adam_step")


def densify_stats(radii, dL_dmeans2D, max_radii2D, xyz_gradient_accum, denom, stream=None) -> None:
    """egs_densify_stats: in-place update of the three bookkeeping tensors."""
    import torch
    with torch.cuda.device(radii.device):
        _check(load().egs_densify_stats(int(radii.shape[0]), _ptr(radii), _ptr(dL_dmeans2D), _ptr(max_radii2D),
                                        _ptr(xyz_gradient_accum), _ptr(denom), _stream_handle(stream)),
               "egs_densify_stats")


def launch_count() -> int:
    return int(load().egs_launch_count())


def profile_enable(on: bool) -> None:
    load().egs_profile_enable(int(bool(on)))


def profile_read() -> dict:
    """{stage name: (total milliseconds, kernel launches)} since the last read; waits for the GPU."""
    lib = load()
    ms = (C.c_double * PROFILE_STAGES)()
    cnt = (C.c_ulonglong * PROFILE_STAGES)()
    _check(lib.egs_profile_read(ms, cnt), "egs_profile_read")
    out = {}
    for i in range(PROFILE_STAGES):
        name = lib.egs_profile_stage_name(i).decode()
        if name:
            out[name] = (float(ms[i]), int(cnt[i]))
    return out
