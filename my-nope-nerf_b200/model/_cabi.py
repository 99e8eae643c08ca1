"""ctypes binding of libnope_nerf_b200.so (include/nope_nerf_b200.h).

There is NO CPU or PyTorch fallback: if the library is missing, or a call is made with tensors that are
not on a CUDA device, this module raises.  PyTorch only supplies device memory and the current stream.
"""
import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(_HERE), "libnope_nerf_b200.so")

NUM_PARAMS = 24
PREC_FP32, PREC_BF16 = 0, 1

PARAM_ORDER = (
    ["layers0.%d.%s" % (i, k) for i in (0, 2, 4, 6) for k in ("weight", "bias")]
    + ["layers1.%d.%s" % (i, k) for i in (0, 2, 4, 6) for k in ("weight", "bias")]
    + ["fc_density.weight", "fc_density.bias", "fc_feature.weight", "fc_feature.bias",
       "rgb_layers.0.weight", "rgb_layers.0.bias", "fc_rgb.weight", "fc_rgb.bias"]
)


class RenderCfg(C.Structure):
    _fields_ = [("n_rays", C.c_int), ("n_samples", C.c_int), ("near_", C.c_float), ("far_", C.c_float),
                ("occ_softplus", C.c_int), ("dist_alpha", C.c_int), ("white_bkgd", C.c_int), ("use_dir", C.c_int),
                ("precision", C.c_int)]


class RayExtras(C.Structure):
    """nnb_ray_extras: per-ray origins / view directions of the NDC sampling branch and their backward outputs."""
    _fields_ = [("ray_origins", C.c_void_p), ("view_dirs", C.c_void_p), ("d_ray_origins", C.c_void_p), ("d_view_dirs", C.c_void_p),
                ("wgrad_stream", C.c_void_p)]


_p, _i, _f, _sz = C.c_void_p, C.c_int, C.c_float, C.c_size_t
_exp = C.POINTER(RayExtras)
_cfgp = C.POINTER(RenderCfg)
_pp = C.POINTER(C.c_void_p)
COPY_MAX = 16   # NNB_COPY_MAX

# name -> argtypes (return type is int unless listed in _RESTYPE)
SIGNATURES = {
    "nnb_version": [],
    "nnb_launch_count": [],
    "nnb_error_string": [_i],
    "nnb_device_info": [C.POINTER(_i), C.POINTER(_i), C.POINTER(_i)],
    "nnb_pose_fwd": [_p, _p, _p, _i, _p, _p],
    "nnb_pose_bwd": [_p, _p, _p, _i, _p, _p, _p, _p],
    "nnb_mat4_inv_fwd": [_p, _p, _p],
    "nnb_mat4_inv_bwd": [_p, _p, _p, _p],
    "nnb_mat4_mul_fwd": [_p, _p, _p, _p],
    "nnb_mat4_mul_bwd": [_p, _p, _p, _p, _p, _p],
    "nnb_distortion_fwd": [_p, _p, _i, _i, _i, _p, _p],
    "nnb_distortion_bwd": [_p, _i, _i, _i, _p, _p, _p, _p],
    "nnb_pixel_grid": [_i, _i, _p, _p, _p],
    "nnb_ray_choice": [_p, C.c_longlong, _i, _p, _p],
    "nnb_copy_many": [_p, _p, _p, _i, _p, _p, _p],
    "nnb_gather_rays_fwd": [_p, _i, _i, _i, _p, _p, _p, _p, _p, _p, _p],
    "nnb_gather_rays_bwd": [_p, _i, _p, _p, _p, _p],
    "nnb_raygen_fwd": [_p, _p, _p, _p, _p, _i, _i, _p, _p, _p, _p, _p, _p, _p],
    "nnb_raygen_bwd": [_p, _p, _p, _p, _p, _p, _i, _i, _p, _p, _p, _p, _p, _p, _p, _p, _p],
    "nnb_render_workspace_bytes": [_cfgp, _i],
    "nnb_render_fwd": [_cfgp, _pp, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _sz, _i, _p],
    "nnb_render_bwd": [_cfgp, _pp, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _pp, _p, _p, _p, _sz, _p],
    "nnb_render_fwd_ex": [_cfgp, _pp, _p, _p, _p, _exp, _p, _p, _p, _p, _p, _p, _p, _sz, _i, _p],
    "nnb_render_bwd_ex": [_cfgp, _pp, _p, _p, _p, _exp, _p, _p, _p, _p, _p, _p, _p, _pp, _p, _p, _p, _sz, _p],
    "nnb_ndc_rays_fwd": [_p, _f, _p, _p, _i, _p, _p, _p],
    "nnb_ndc_rays_bwd": [_p, _f, _p, _p, _i, _p, _p, _p, _p, _p, _p],
    "nnb_composite_fwd": [_p, _p, _p, _i, _i, _i, _i, _p, _p, _p, _p, _p],
    "nnb_composite_bwd": [_p, _p, _p, _p, _i, _i, _i, _i, _p, _p, _p, _p, _p, _p],
    "nnb_mlp_points_workspace_bytes": [C.c_longlong],
    "nnb_mlp_points_fwd": [_pp, _p, _p, C.c_longlong, _i, _i, _p, _p, _p, _sz, _p],
    "nnb_mlp_points_grad_workspace_bytes": [C.c_longlong],
    "nnb_mlp_points_density_grad": [_pp, _p, C.c_longlong, _p, _p, _p, _sz, _p],
    "nnb_pack_weights_bytes": [],
    "nnb_pack_weights": [_pp, _p, _p],
    "nnb_debug_set_trace": [_p],
    "nnb_debug_bwd_timing": [_i, C.POINTER(C.c_float)],
    "nnb_rgb_loss_fwd": [_p, _p, _i, _p, _p],
    "nnb_rgb_loss_bwd": [_p, _p, _i, _p, _p, _p],
    "nnb_depth_l1_fwd": [_p, _p, _p, _i, _p, _p],
    "nnb_depth_l1_bwd": [_p, _p, _p, _i, _p, _p, _p, _p, _p],
    "nnb_median_select": [_p, _p, _i, _p, _p, _p],
    "nnb_depth_invariant_fwd": [_p, _p, _p, _i, _p, _p, _p, _p],
    "nnb_depth_invariant_bwd": [_p, _p, _p, _i, _p, _p, _p, _p, _p, _p],
    "nnb_backproject_fwd": [_p, _i, _i, _i, _i, _p, _p, _f, _p, _p, _p],
    "nnb_backproject_bwd": [_p, _i, _i, _i, _i, _p, _p, _p, _p, _p],
    "nnb_resize_bilinear": [_p, _i, _i, _i, _i, _i, _p, _p],
    "nnb_reproj_fwd": [_p, _p, _i, _i, _p, _p, _p, _p, _f, _p, _p, _p, _p],
    "nnb_reproj_bwd": [_p, _i, _i, _p, _p, _p, _p, _f, _p, _p, _p, _p, _p],
    "nnb_photometric_fwd": [_p, _p, _p, _i, _p, _p],
    "nnb_photometric_bwd": [_p, _p, _p, _i, _p, _p, _p, _p, _p],
    "nnb_chamfer_workspace_bytes": [_i, _i],
    "nnb_debug_chamfer_pass1": [_i],
    "nnb_chamfer_nn": [_p, _i, _p, _i, _p, _p, _p, C.c_size_t, _p],
    "nnb_chamfer_nn_static": [_p, _i, _p, _i, _p, _p, _p],
    "nnb_chamfer_err_fwd": [_p, _i, _p, _p, _p, _p],
    "nnb_chamfer_err_bwd": [_p, _i, _p, _p, _p, _p, _p, _p],
    "nnb_align_clouds_fwd": [_p, _i, _p, _i, _p, _p, _p, _p, _p],
    "nnb_align_clouds_bwd": [_p, _i, _p, _i, _p, _p, _p, _p, _p, _p, _p, _p, _p],
    "nnb_adam_step": [_p, _p, _p, _p, _sz, _f, _f, _f, _f, _i, _p],
}
_RESTYPE = {"nnb_error_string": C.c_char_p, "nnb_launch_count": C.c_ulonglong, "nnb_render_workspace_bytes": C.c_size_t, "nnb_pack_weights_bytes": C.c_size_t, "nnb_chamfer_workspace_bytes": C.c_size_t,
            "nnb_mlp_points_workspace_bytes": C.c_size_t, "nnb_mlp_points_grad_workspace_bytes": C.c_size_t}

_lib = None
launch_count = 0  # number of C-ABI calls that launch kernels (bench.py reports it)


def lib():
    """Load the shared library (once).  Raises if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                "libnope_nerf_b200.so not found at %s -- run `python __graft_entry__.py build`; "
                "there is no CPU fallback" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        for name, args in SIGNATURES.items():
            fn = getattr(L, name)
            fn.argtypes = args
            fn.restype = _RESTYPE.get(name, C.c_int)
        _lib = L
    return _lib


def stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def ptr(t):
    """Device pointer of a contiguous CUDA tensor (None -> NULL)."""
    if t is None:
        return None
    if not t.is_cuda:
        raise RuntimeError("nope_nerf_b200 kernels need CUDA tensors (got %s); there is no CPU fallback" % t.device)
    if not t.is_contiguous():
        raise RuntimeError("nope_nerf_b200 kernels need contiguous tensors")
    return C.c_void_p(t.data_ptr())


def check(rc, what=""):
    if rc != 0:
        msg = lib().nnb_error_string(rc)
        raise RuntimeError("%s failed: %s (code %d)" % (what or "nope_nerf_b200 call", msg.decode() if msg else "?", rc))


profile = None  # set to a dict to collect per-entry-point device time: name -> [ms_total, calls]
_pending = []


def call(name, *args):
    global launch_count
    launch_count += 1
    if profile is None:
        check(getattr(lib(), name)(*args), name)
        return
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    check(getattr(lib(), name)(*args), name)
    b.record()
    _pending.append((name, a, b))


def profile_flush():
    """Resolve the recorded event pairs into `profile` (call after a synchronize)."""
    torch.cuda.synchronize()
    for name, a, b in _pending:
        rec = profile.setdefault(name, [0.0, 0])
        rec[0] += a.elapsed_time(b)
        rec[1] += 1
    del _pending[:]


def kernel_launches():
    return int(lib().nnb_launch_count())


def f32(t):
    """Contiguous fp32 view/copy of `t`."""
    if t.dtype != torch.float32:
        t = t.float()
    return t.contiguous()


def param_ptr_array(tensors):
    arr = (C.c_void_p * NUM_PARAMS)()
    for i, t in enumerate(tensors):
        arr[i] = ptr(t).value
    return arr
