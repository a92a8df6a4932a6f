"""ctypes binding of the C ABI declared in include/isogs_rast.h.

The product path has NO CPU fallback: if libisogs_rast.so is missing or a call fails this
module raises.  torch is used only for device memory and streams.
"""
import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libisogs_rast.so")

REC = 12
TILE = 16
F_LOG_SCALES, F_LOGIT_OPAC, F_NO_OPAC_EXTENT = 1, 2, 4
E_ARG, E_CUDA, E_CAPACITY = -1, -2, -3


class Camera(C.Structure):
    """isogs_camera (include/isogs_rast.h)."""
    _fields_ = [("width", C.c_int32), ("height", C.c_int32),
                ("fx", C.c_float), ("fy", C.c_float), ("cx", C.c_float), ("cy", C.c_float),
                ("R", C.c_float * 9), ("t", C.c_float * 3),
                ("near_plane", C.c_float), ("far_plane", C.c_float), ("eps2d", C.c_float),
                ("bg", C.c_float * 3)]

    @property
    def n_tiles(self):
        return ((self.width + TILE - 1) // TILE) * ((self.height + TILE - 1) // TILE)


class Pose(C.Structure):
    _fields_ = [("q", C.c_float * 4), ("t", C.c_float * 3)]


class AdamGroup(C.Structure):
    _fields_ = [("param", C.c_void_p), ("grad", C.c_void_p), ("exp_avg", C.c_void_p), ("exp_avg_sq", C.c_void_p),
                ("numel", C.c_int64), ("lr", C.c_float)]


class Rows(C.Structure):
    _fields_ = [("inp", C.c_void_p), ("out", C.c_void_p), ("width", C.c_int32)]


class AdamState(C.Structure):
    """isogs_adam_state: DEVICE-resident; this mirror only gives its size / layout (32 bytes)."""
    _fields_ = [("step", C.c_int64), ("inv_bc1", C.c_double), ("inv_bc2_sqrt", C.c_double), ("reserved", C.c_double)]


class Frame(C.Structure):
    """isogs_frame: the DEVICE-side current-frame slot (48 bytes); built on the host, copied to the device."""
    _fields_ = [("gt_rgb", C.c_void_p), ("gt_depth", C.c_void_p), ("pose", C.c_float * 7), ("inv_n_depth", C.c_float)]


class PeerAdamGroup(C.Structure):
    _fields_ = [("offset", C.c_int64), ("numel", C.c_int64), ("param", C.c_void_p * 16), ("exp_avg", C.c_void_p),
                ("exp_avg_sq", C.c_void_p), ("lr", C.c_float)]


class Gaussians(C.Structure):
    _fields_ = [("N", C.c_int64), ("scale_dim", C.c_int32), ("reserved", C.c_int32),
                ("p", C.c_void_p * 5), ("g", C.c_void_p * 5), ("m", C.c_void_p * 5), ("v", C.c_void_p * 5)]


class Workspace(C.Structure):
    _fields_ = [("rec", C.c_void_p), ("v_rec", C.c_void_p), ("radii", C.c_void_p), ("bucket", C.c_void_p),
                ("vals", C.c_void_p), ("ranges", C.c_void_p), ("totals", C.c_void_p), ("sort_temp", C.c_void_p),
                ("sort_temp_bytes", C.c_size_t), ("isect_cap", C.c_int64), ("render", C.c_void_p),
                ("v_render", C.c_void_p), ("alpha", C.c_void_p), ("last", C.c_void_p), ("small", C.c_void_p),
                ("adam_state", C.c_void_p), ("dmaps", C.c_void_p), ("median_scratch", C.c_void_p),
                ("median_temp", C.c_void_p), ("median_temp_bytes", C.c_size_t), ("tile_rects", C.c_void_p)]


class StepParams(C.Structure):
    _fields_ = [("flags", C.c_int32), ("phases", C.c_int32), ("w_im", C.c_float), ("w_depth", C.c_float),
                ("sil_thres", C.c_float), ("lr", C.c_float * 5), ("beta1", C.c_float), ("beta2", C.c_float),
                ("eps", C.c_float), ("lr_pose", C.c_float * 2)]


STEP_SSIM, STEP_OUTLIER, STEP_NO_SIL = 1, 2, 4
PHASE_RENDER, PHASE_ADAM = 1, 2
STEP_EVENTS = 9
STEP_EVENT_NAMES = ("begin", "preprocess_fwd", "tile_bin", "blend_fwd", "loss", "ssim", "blend_bwd", "preprocess_bwd",
                    "adam")


def build(force=False):
    """Compile the CUDA library for sm_100a (nvcc cross-compiles without a GPU)."""
    src_dir = os.path.join(_HERE, "csrc")
    srcs = [os.path.join(src_dir, f) for f in os.listdir(src_dir)] + \
           [os.path.join(_HERE, "..", "include", "isogs_rast.h"), os.path.join(_HERE, "Makefile")]
    if (not force and os.path.exists(LIB_PATH)
            and all(os.path.getmtime(LIB_PATH) >= os.path.getmtime(s) for s in srcs)):
        return LIB_PATH
    subprocess.check_call(["make", "-C", _HERE, "-s"])
    return LIB_PATH


_lib = None

_VP, _I64, _I32, _F, _SZ = C.c_void_p, C.c_int64, C.c_int, C.c_float, C.c_size_t

_SIGS = {
    "isogs_last_error": (C.c_char_p, []),
    "isogs_version": (C.c_int, []),
    "isogs_temp_bytes": (C.c_int, [_I64, _I64, C.POINTER(_SZ), C.POINTER(_SZ)]),
    "isogs_preprocess_fwd": (C.c_int, [C.POINTER(Camera), C.POINTER(Pose), _VP, _I64, _VP, _VP, _I32, _VP, _VP, _VP,
                                       _I32, _I32, _VP, _VP, _VP, _VP, _VP, _SZ, C.POINTER(_I64), _VP, _VP, _VP, _VP]),
    "isogs_tile_bin": (C.c_int, [C.POINTER(Camera), _I64, _I64, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _SZ, _VP]),
    "isogs_tile_bin_cub": (C.c_int, [C.POINTER(Camera), _I64, _I64, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _SZ, _VP]),
    "isogs_tile_bin_async": (C.c_int, [C.POINTER(Camera), _I64, _I64, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _SZ, _VP,
                                       _I32, _VP]),
    "isogs_blend_fwd": (C.c_int, [C.POINTER(Camera), _I32, _VP, _VP, _VP, _VP, _VP, _VP, _VP]),
    "isogs_blend_bwd": (C.c_int, [C.POINTER(Camera), _I32, _I64, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP]),
    "isogs_blend_bwd_geom": (C.c_int, [C.POINTER(Camera), _I32, _I64, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP]),
    "isogs_preprocess_bwd": (C.c_int, [C.POINTER(Camera), C.POINTER(Pose), _VP, _I64, _VP, _VP, _I32, _VP, _VP, _I32,
                                       _I32, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP]),
    "isogs_loss_l1_fused": (C.c_int, [C.POINTER(Camera), _I32, _VP, _VP, _VP, _VP, _F, _F, _F, _I32, _F, _VP, _VP, _VP,
                                    _VP]),
    "isogs_depth_outlier_threshold": (C.c_int, [C.POINTER(Camera), _VP, _VP, _F, _VP, _VP, _SZ, _VP, _VP]),
    "isogs_ssim_loss_fused": (C.c_int, [C.POINTER(Camera), _VP, _VP, _F, _VP, _VP, _VP, _VP, _VP]),
    "isogs_adam_fused": (C.c_int, [C.POINTER(AdamGroup), _I32, _F, _F, _F, _I32, _VP, _VP, _VP]),
    "isogs_step_begin": (C.c_int, [_VP, _I32, _VP, _F, _F, _VP]),
    "isogs_pose_keep_best": (C.c_int, [_VP, _VP, _VP, _VP, _VP]),
    "isogs_depth_outlier_threshold_noalloc": (C.c_int, [C.POINTER(Camera), _VP, _VP, _F, _VP, _VP]),
    "isogs_map_step": (C.c_int, [C.POINTER(Camera), _VP, C.POINTER(Workspace), C.POINTER(Gaussians),
                                 C.POINTER(StepParams), C.POINTER(_VP), _VP]),
    "isogs_map_step_ba": (C.c_int, [C.POINTER(Camera), _VP, C.POINTER(Workspace), C.POINTER(Gaussians), _VP, _VP, _VP,
                                    _VP, C.POINTER(StepParams), C.POINTER(_VP), _VP]),
    "isogs_track_step": (C.c_int, [C.POINTER(Camera), _VP, C.POINTER(Workspace), C.POINTER(Gaussians), _VP, _VP, _VP,
                                   _VP, _VP, C.POINTER(StepParams), C.POINTER(_VP), _VP]),
    "isogs_unpack_rgbd": (C.c_int, [C.POINTER(Camera), _VP, _VP, _F, _VP, _VP, _VP]),
    "isogs_graph_begin": (C.c_int, [_VP]),
    "isogs_graph_end": (C.c_int, [_VP, C.POINTER(_VP)]),
    "isogs_graph_launch": (C.c_int, [_VP, _VP]),
    "isogs_graph_destroy": (C.c_int, [_VP]),
    "isogs_graph_info": (C.c_int, [_VP, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]),
    "isogs_densify_stats": (C.c_int, [_I64, _VP, _VP, _I32, _F, _F, _VP, _VP, _VP, _VP]),
    "isogs_densify_select": (C.c_int, [_I64, _VP, _VP, _VP, _I32, _F, _F, _I32, _VP, _VP, _SZ, _VP, _VP]),
    "isogs_densify_emit": (C.c_int, [_I64, _I64, _VP, _VP, _VP, _VP, _VP, _VP, _I32, _I32, _VP, C.c_uint64, _VP, _VP]),
    "isogs_mask_and": (C.c_int, [_I64, _VP, _VP, _VP]),
    "isogs_keyframe_overlap": (C.c_int, [C.POINTER(Camera), _VP, C.POINTER(C.c_float), _VP, _I32, C.c_uint64, _VP, _I32,
                                         _I32, _VP, _VP, _VP, _SZ, _VP, _VP]),
    "isogs_knn_workspace": (C.c_int, [_I64, _I32, C.POINTER(C.c_size_t), C.POINTER(C.c_size_t)]),
    "isogs_knn16": (C.c_int, [_I64, _VP, _I32, _I32, _VP, C.c_uint64, _VP, _VP, _VP, _VP, _SZ, _VP, _VP]),
    "isogs_iso_flat_loss": (C.c_int, [_I64, _VP, _VP, _VP, _VP, _I32, _VP, _VP, _F, _F, _F, _VP, _VP, _VP, _VP, _VP,
                                      _VP, _VP]),
    "isogs_peer_alloc": (C.c_int, [_SZ, C.POINTER(_VP), _VP]),
    "isogs_peer_open": (C.c_int, [_VP, C.POINTER(_VP)]),
    "isogs_peer_close": (C.c_int, [_VP]),
    "isogs_peer_free": (C.c_int, [_VP]),
    "isogs_allreduce_avg_p2p": (C.c_int, [_I32, _I32, C.POINTER(_VP), C.POINTER(_VP), _I64, C.c_uint32, _I32, _VP, _VP]),
    "isogs_peer_wait": (C.c_int, [_I32, _VP, C.c_uint32, _VP, _VP]),
    "isogs_reduce_adam_p2p": (C.c_int, [_I32, _I32, C.POINTER(_VP), C.POINTER(_VP), _I64, C.POINTER(PeerAdamGroup), _I32,
                                        _F, _F, _F, _VP, C.c_uint32, _I32, _VP, _VP]),
    "isogs_reduce_adam_nvls": (C.c_int, [_I32, _I32, C.POINTER(_VP), C.POINTER(_VP), _I64, C.POINTER(PeerAdamGroup), _I32,
                                         _F, _F, _F, _VP, C.c_uint32, _I32, _VP, _VP, C.POINTER(_VP), _VP]),
    "isogs_prune_mask": (C.c_int, [_I64, _VP, _VP, _I32, _F, _F, _VP, _VP]),
    "isogs_compact": (C.c_int, [_I64, _VP, C.POINTER(Rows), _I32, _VP, _VP, _SZ, _VP, _VP]),
    "isogs_backproject_append": (C.c_int, [C.POINTER(Camera), C.POINTER(Pose), _VP, _VP, _VP, _VP, _F, _F, _VP, _VP,
                                           _VP, _VP, _VP, _I32, _F, C.c_uint64, _I64, _VP, _VP, _SZ, _VP, _VP]),
}

EXPORTED_SYMBOLS = tuple(_SIGS.keys())


def lib():
    """Load libisogs_rast.so; raises if it is not built (no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `make -C {_HERE}` (or __graft_entry__.build()). "
                "There is no CPU fallback.")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in _SIGS.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


class IsogsError(RuntimeError):
    pass


def check(rc):
    if rc != 0:
        msg = lib().isogs_last_error()
        raise IsogsError(f"isogs error {rc}: {msg.decode() if msg else ''}")


def make_camera(width, height, fx, fy, cx, cy, viewmat=None, near=0.01, far=100.0, eps2d=0.3, bg=(0.0, 0.0, 0.0)):
    """viewmat: row-major 4x4 (or 3x4) world->camera, any array-like on the host."""
    cam = Camera()
    cam.width, cam.height = int(width), int(height)
    cam.fx, cam.fy, cam.cx, cam.cy = float(fx), float(fy), float(cx), float(cy)
    if viewmat is None:
        R = [1, 0, 0, 0, 1, 0, 0, 0, 1]
        t = [0, 0, 0]
    else:
        R = [float(viewmat[i][j]) for i in range(3) for j in range(3)]
        t = [float(viewmat[i][3]) for i in range(3)]
    cam.R[:] = R
    cam.t[:] = t
    cam.near_plane, cam.far_plane, cam.eps2d = float(near), float(far), float(eps2d)
    cam.bg[:] = [float(b) for b in bg]
    return cam


def make_pose(q, t):
    p = Pose()
    p.q[:] = [float(v) for v in q]
    p.t[:] = [float(v) for v in t]
    return p


def temp_bytes(N, cap):
    a, b = _SZ(0), _SZ(0)
    check(lib().isogs_temp_bytes(int(N), int(cap), C.byref(a), C.byref(b)))
    return int(a.value), int(b.value)
