"""CPU: the C-ABI shared library loads and exports every symbol include/*.h declares; the CLI keeps the
reference's surface for bad scene ids; nothing routes through the oracle."""
import ctypes as C
import os
import re
import subprocess

import pytest

import common

ROOT = common.ROOT


def _declared_functions():
    names = set()
    for hdr in ("rt_api.h", "rt_scene.h"):
        src = open(os.path.join(ROOT, "include", hdr)).read()
        src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
        for m in re.finditer(r"^\s*(?:const\s+)?[A-Za-z_][\w\s\*]*?\b(rt_[a-z0-9_]+)\s*\(", src, flags=re.M):
            names.add(m.group(1))
    return names


def test_exports_every_declared_symbol(product_lib):
    L = product_lib.lib()
    declared = _declared_functions()
    assert len(declared) >= 40
    missing = [n for n in sorted(declared) if not hasattr(L, n)]
    assert not missing, missing
    assert set(product_lib.EXPORTS) <= declared


def test_struct_layouts_match_header(product_lib):
    # sizes the C side asserts too (rt_world.c _Static_asserts); these guard the ctypes mirrors
    assert C.sizeof(product_lib.api.RenderParams) == 144
    assert C.sizeof(product_lib.api.ProbeHit) == 72
    assert C.sizeof(product_lib.api.RenderSettings) == 80       # == the reference's RenderSettings
    assert C.sizeof(product_lib.api.SceneDesc) == 56


def test_no_gpu_fails_loudly(product_lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(product_lib.RtError) as e:
        product_lib.Context(0)
    assert "no CPU fallback" in str(e.value)


def test_group_without_gpu_fails_loudly(product_lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(product_lib.RtError) as e:
        product_lib.Group(2)
    assert "no CPU fallback" in str(e.value)


def test_product_does_not_link_oracle(product_lib):
    out = subprocess.run(["ldd", product_lib.LIB_PATH], capture_output=True, text=True).stdout
    assert "oracle" not in out and "libref_host" not in out
    syms = subprocess.run(["nm", "-D", "--defined-only", product_lib.LIB_PATH], capture_output=True, text=True).stdout
    assert " orc_" not in syms and " ref_" not in syms
    for dirpath, _, files in os.walk(os.path.join(ROOT, "cuda_raytracing_b200")):
        for f in files:
            if f.endswith((".c", ".cu", ".cpp", ".h", ".cuh", ".py")):
                text = open(os.path.join(dirpath, f)).read()
                assert "rt_oracle.h" not in text and "liboracle" not in text and "oracleabi" not in text, f


def test_cli_surface():
    exe = os.path.join(ROOT, "bin", "raytracing")
    if not os.path.exists(exe):
        subprocess.check_call(["make", "-s", "-C", ROOT, "bin/raytracing"])
    r = subprocess.run([exe, "42"], capture_output=True, text=True)
    assert r.returncode == 1 and r.stdout == "Unknown scene\n"       # main.cu:113-116
    r = subprocess.run([exe, "-3"], capture_output=True, text=True)
    assert r.returncode == 1
