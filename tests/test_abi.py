"""-m "not gpu": the C-ABI library loads, exports every symbol include/rtb200.h declares, struct layouts match the
header, and — there being no GPU here — the product path fails loudly instead of falling back to a CPU path."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_functions():
    src = open(os.path.join(ROOT, "include", "rtb200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(rt_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_expected_entry_points():
    fns = header_functions()
    for f in ("rt_create", "rt_destroy", "rt_scene_begin", "rt_scene_spheres", "rt_scene_triangles", "rt_scene_mesh",
              "rt_scene_lights", "rt_scene_commit", "rt_render", "rt_last_error"):
        assert f in fns


def test_library_exports_every_declared_symbol(rtb):
    L = rtb.load_library()
    for f in header_functions():
        assert hasattr(L, f), f"librtb200.so does not export {f}"
    assert sorted(rtb.host.ABI_SYMBOLS) == header_functions()
    assert L.rt_abi_version() == 2


def test_struct_layouts_match_header(rtb, scenes):
    assert C.sizeof(rtb.host.RenderParams) == 88          # rt_render_params
    assert scenes.MAT_DTYPE.itemsize == 56                # rt_material
    assert C.sizeof(rtb.host.Stats) == 8 * 3 + 4 * 4 + 8 * 2 + 4 * 4 + 4 * 8 + 4 * 8 + 8


def test_no_cpu_fallback_without_a_device(rtb):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    with pytest.raises(rtb.host.RtError) as ei:
        rtb.RayTracer((0,))
    assert ei.value.code == -2 and "no CPU fallback" in str(ei.value)       # RT_E_CUDA


def test_null_arguments_are_rejected(rtb):
    L = rtb.load_library()
    assert L.rt_create(None, None, 1) == -1
    assert L.rt_scene_begin(None) == -1
    assert L.rt_render(None, None, None, None, None, None, None) == -1
    L.rt_destroy(None)


def test_product_package_does_not_reference_the_oracle():
    """Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may touch oracle/."""
    pkg = os.path.join(ROOT, "unity-raytracer_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                txt = open(os.path.join(dp, f), errors="ignore").read()
                assert "librt_oracle" not in txt and "import oracle" not in txt and "from oracle" not in txt, f
                assert "librt_emu" not in txt, f


def build_abi_smoke(out_dir):
    """tests/abi_smoke.c as a strict C99 translation unit against include/rtb200.h, linked to the product library"""
    import subprocess
    exe = os.path.join(out_dir, "abi_smoke")
    lib_dir = os.path.join(ROOT, "unity-raytracer_b200")
    subprocess.check_call(["gcc", "-std=c99", "-pedantic", "-Wall", "-Wextra", "-Werror", "-I", os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "tests", "abi_smoke.c"), "-o", exe, "-L", lib_dir, "-lrtb200", "-lm",
                           "-Wl,-rpath," + lib_dir])
    return exe


def test_header_is_plain_c_and_the_c_caller_fails_loudly_without_a_device(rtb, tmp_path):
    """the boundary is bound from C (a managed [DllImport] sees exactly this view). Without a GPU the C caller
    must get RT_E_CUDA from rt_create, never a CPU path; with one, the GPU suite runs the full smoke (test_gpu_parity.py)."""
    import subprocess
    import torch
    rtb.load_library()
    exe = build_abi_smoke(str(tmp_path))
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present: the full C smoke test runs in the gpu suite")
    r = subprocess.run([exe, "--no-gpu"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
