"""CPU checks of the drop-in boundary: the C-ABI library loads, exports every symbol include/hf_b200.h declares,
and fails loudly (no CPU fallback) when no CUDA device is usable. No compute calls here."""
import ctypes as C
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "hf_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(hf_[a-z0-9_]+)\s*\(", src)))


@pytest.fixture(scope="module")
def lib():
    import hfb200
    if not os.path.exists(hfb200.lib_path()):
        import __graft_entry__
        __graft_entry__.build()
    return C.CDLL(hfb200.lib_path())


def test_every_declared_symbol_is_exported(lib):
    names = _declared()
    assert len(names) >= 30 and "hf_ray_intersect_preliminary" in names and "hf_backward" in names
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, f"declared in include/hf_b200.h but not exported: {missing}"
    assert lib.hf_abi_version() == 4


def test_python_mirror_binds_every_compute_symbol(lib):
    import hfb200
    L = hfb200._load()
    for n in _declared():
        assert hasattr(L, n)


def test_no_cpu_fallback(lib):
    """Without a CUDA device the constructor must fail with HF_ERR_CUDA and a message -- never compute on the CPU."""
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    import hfb200
    d = hfb200._Desc()
    d.width = d.height = 4
    d.max_height = 1.0
    d.to_world[:] = [1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0]
    ctx = C.c_void_p()
    rc = lib.hf_create(C.byref(ctx), C.byref(d))
    lib.hf_last_error.restype = C.c_char_p
    assert rc == 2 and not ctx.value
    assert b"no CPU path" in lib.hf_last_error() or b"CUDA" in lib.hf_last_error()
    with pytest.raises(hfb200.HeightfieldError, match="no CPU fallback"):
        hfb200.load_dict({"type": "heightfield", "heights": torch.zeros(4, 4)})


def test_invalid_arguments_are_rejected_before_any_device_work(lib):
    import hfb200
    lib.hf_last_error.restype = C.c_char_p
    assert lib.hf_create(None, None) == 1
    d = hfb200._Desc()
    d.width, d.height = 1, 5
    ctx = C.c_void_p()
    assert lib.hf_create(C.byref(ctx), C.byref(d)) == 1 and b"2x2" in lib.hf_last_error()
    assert lib.hf_destroy(None) == 0
    assert lib.hf_zero_grad(None, None) == 1


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "mitsuba3-differentiable-heightfield-rendering_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                s = open(os.path.join(dirpath, f)).read()
                bad = re.findall(r"#include[^\n]*oracle|^\s*(?:from|import)\s+oracle|libhf_oracle|oracle/_build|oracle/_ref|hfo_[a-z_]+\s*\(", s, flags=re.M)
                assert not bad, (f, bad)


# ---- the C++ host class (include/hf_heightfield.hpp) ---------------------------------------------------------------------------
def test_cpp_host_class_compiles_standalone_and_mirrors_the_shape_virtuals(tmp_path):
    """The header needs nothing but hf_b200.h (no CUDA headers) and declares the Shape methods the plugin provides, in the
    reference's order (include/mitsuba/render/shape.h:137-197, 253, 432, 533-550)."""
    import subprocess
    src = tmp_path / "tu.cpp"
    src.write_text('#include "hf_heightfield.hpp"\nint main() { return sizeof(hf::Heightfield) > 0 ? 0 : 1; }\n')
    subprocess.check_call(["g++", "-std=c++17", "-fsyntax-only", "-Wall", "-Wextra", "-Werror", "-I" + os.path.join(ROOT, "include"), str(src)])
    hpp = open(os.path.join(ROOT, "include", "hf_heightfield.hpp")).read()
    body = hpp[hpp.index("class Heightfield {"):]
    order = ["ray_intersect_preliminary(", "ray_test(", "compute_surface_interaction(", "ray_intersect(", "bbox()", "primitive_count()",
             "void traverse(", "void parameters_changed(", "parameters_grad_enabled()"]
    pos = [body.index(" " + m if not m.startswith("void") else m) for m in order]
    assert pos == sorted(pos), "methods are not in the order of the reference's Shape interface"


def test_cpp_smoke_binary_is_built(lib):
    import __graft_entry__
    exe = os.path.join(ROOT, "tests", "cpp", "abi_smoke")
    if not os.path.exists(exe):
        __graft_entry__.build_cpp_smoke()
    assert os.access(exe, os.X_OK)


@pytest.mark.gpu
def test_cpp_host_class_on_the_gpu(lib):
    """tests/cpp/abi_smoke.cpp: hf::Heightfield from plain C++ -- the reference's known answers (test_mesh.py test10 / test16) on the
    flat 2x2 field, bbox / traverse / parameters_changed, exceptions with the reference's messages."""
    import subprocess
    import __graft_entry__
    exe = os.path.join(ROOT, "tests", "cpp", "abi_smoke")
    if not os.path.exists(exe):
        __graft_entry__.build_cpp_smoke()
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "abi_smoke ok" in r.stdout, r.stdout + r.stderr
