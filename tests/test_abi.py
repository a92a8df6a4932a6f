"""The C-ABI library loads and exports every symbol include/isogs_rast.h declares (no compute)."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    txt = open(os.path.join(ROOT, "include", "isogs_rast.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(isogs_[a-z0-9_]+)\s*\(", txt)))


def test_header_symbols_exported(pkg):
    pkg.build()
    L = ctypes.CDLL(pkg._lib.LIB_PATH)
    names = _declared()
    assert len(names) >= 12
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/isogs_rast.h but not exported"


def test_binding_covers_header(pkg):
    assert set(_declared()) == set(pkg._lib.EXPORTED_SYMBOLS)


def test_version_and_error_string(pkg):
    L = pkg.lib()
    assert L.isogs_version() >= 100
    # bad argument -> error code + message, never an exception/exit across the ABI
    rc = L.isogs_blend_fwd(None, 3, None, None, None, None, None, None, None)
    assert rc == pkg._lib.E_ARG
    assert b"isogs_blend_fwd" in L.isogs_last_error()


def test_blend_rejects_misaligned_records(pkg):
    # rec / v_rec are read and accumulated as 16-byte rows (float4 loads, red.global.add.v4.f32): a misaligned
    # pointer is an argument error, reported before any CUDA call
    L = pkg.lib()
    cam = pkg._lib.Camera()
    cam.width, cam.height = 64, 48
    rc = L.isogs_blend_fwd(ctypes.byref(cam), 5, 0x1004, None, None, None, None, None, None)
    assert rc == pkg._lib.E_ARG and b"16-byte" in L.isogs_last_error()
    for fn in (L.isogs_blend_bwd, L.isogs_blend_bwd_geom):
        rc = fn(ctypes.byref(cam), 5, 0, 0x1000, None, None, None, None, None, None, 0x2008, None)
        assert rc == pkg._lib.E_ARG and b"16-byte" in L.isogs_last_error()


def test_struct_layout_matches_header(pkg):
    # isogs_camera: 2 i32 + 4 f + 9 f + 3 f + 3 f + 3 f = 96 bytes; isogs_pose 28 bytes
    assert ctypes.sizeof(pkg._lib.Camera) == 96
    assert ctypes.sizeof(pkg._lib.Pose) == 28


def test_step_struct_layouts_match_the_compiler(pkg, tmp_path):
    """The ctypes mirrors of the fused-step structs against sizeof / offsetof as gcc sees the header."""
    import subprocess
    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "isogs_rast.h"\nint main(void){printf("%zu %zu %zu %zu %zu %zu %zu %zu\\n",'
                   'sizeof(isogs_adam_state), sizeof(isogs_frame), offsetof(isogs_frame, pose), sizeof(isogs_gaussians),'
                   'sizeof(isogs_workspace), offsetof(isogs_workspace, adam_state), sizeof(isogs_step_params),'
                   'offsetof(isogs_step_params, lr));return 0;}\n')
    exe = tmp_path / "sz"
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    got = [int(v) for v in subprocess.check_output([str(exe)], text=True).split()]
    L = pkg._lib
    want = [ctypes.sizeof(L.AdamState), ctypes.sizeof(L.Frame), L.Frame.pose.offset, ctypes.sizeof(L.Gaussians),
            ctypes.sizeof(L.Workspace), L.Workspace.adam_state.offset, ctypes.sizeof(L.StepParams), L.StepParams.lr.offset]
    assert got == want, (got, want)
    assert ctypes.sizeof(L.AdamState) == 32 and ctypes.sizeof(L.Frame) == 48


def test_no_cpu_fallback(pkg):
    import torch
    import pytest
    cam = pkg.make_camera(32, 32, 30, 30, 15.5, 15.5)
    z = torch.zeros(4, 3)
    with pytest.raises(RuntimeError):
        pkg.raster.forward(cam, z, torch.ones(4, 1), None, torch.ones(4), z)


def test_dropin_names():
    import diff_gaussian_rasterization as d
    assert d.GaussianRasterizationSettings._fields[:11] == (
        "image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier", "viewmatrix", "projmatrix",
        "sh_degree", "campos", "prefiltered")
    assert callable(d.GaussianRasterizer)
