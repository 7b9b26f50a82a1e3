"""CPU checks of the drop-in boundary (no GPU, no compute calls): the C-ABI library loads, exports every symbol
include/spk_cuda.h declares, the ctypes mirror has the header's struct sizes, and the product fails loudly — never falls
back — when there is no CUDA device. Also: the product never links or imports the oracle."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

from sparkle_b200 import _abi as A

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "spk_cuda.h")


def declared_functions(path):
    src = open(path).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(spk_\w+)\s*\(", src)))


def test_header_symbols_are_exported():
    lib = A.cuda_lib()
    names = declared_functions(HEADER)
    assert len(names) >= 18
    for n in names:
        assert hasattr(lib, n), f"libsparkle_cuda.so does not export {n}"
        assert n in A.CUDA_SYMBOLS or n in ("spk_cuda_last_error", "spk_cuda_abi_version", "spk_cuda_create",
                                            "spk_cuda_destroy"), f"{n} has no ctypes prototype in _abi.py"
    assert lib.spk_cuda_abi_version() == A.ABI_VERSION


def test_host_library_symbols():
    lib = A.host_lib()
    for hdr in ("bvh_builder.h", "cuda_renderer.h", "gltf_loader.h"):
        path = os.path.join(ROOT, "sparkle_b200", "host", hdr)
        if not os.path.exists(path):
            continue
        for n in declared_functions(path):
            assert hasattr(lib, n), f"libsparkle_host.so does not export {n}"


def test_struct_sizes_match_header(tmp_path):
    """compile a tiny C program against the header and compare sizeof() with the ctypes mirror"""
    names = {"SpkBvhNode": A.BvhNode, "SpkBvhDesc": A.BvhDesc, "SpkMeshDesc": A.MeshDesc,
             "SpkPrimitiveDesc": A.PrimitiveDesc, "SpkMaterialDesc": A.MaterialDesc, "SpkTextureDesc": A.TextureDesc,
             "SpkSkyDesc": A.SkyDesc, "SpkSceneDesc": A.SceneDesc, "SpkCamera": A.Camera,
             "SpkRenderParams": A.RenderParams, "SpkRay": A.Ray, "SpkHit": A.Hit, "SpkStats": A.Stats,
             "SpkCreateInfo": A.CreateInfo}
    src = tmp_path / "sz.c"
    body = "\n".join(f'  printf("{n} %zu\\n", sizeof({n}));' for n in names)
    src.write_text(f'#include <stdio.h>\n#include "{HEADER}"\nint main(void) {{\n{body}\n  return 0; }}\n')
    exe = tmp_path / "sz"
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Werror", "-o", str(exe), str(src)])  # the header is plain C
    out = dict(line.split() for line in subprocess.check_output([str(exe)], text=True).splitlines())
    for n, cls in names.items():
        assert int(out[n]) == C.sizeof(cls), f"{n}: header {out[n]} B, ctypes {C.sizeof(cls)} B"
    assert C.sizeof(A.MaterialDesc) == 64 and C.sizeof(A.Ray) == 32 and C.sizeof(A.Hit) == 20


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.mark.skipif(_has_gpu(), reason="this test asserts the no-device behaviour")
def test_no_device_fails_loudly():
    """no CPU fallback: without a CUDA device the context cannot be created and says why"""
    lib = A.cuda_lib()
    info = A.CreateInfo(A.ABI_VERSION, 0, 64, 64)
    handle = A.CtxP()
    rc = lib.spk_cuda_create(C.byref(info), C.byref(handle))
    assert rc == A.ERR_NO_DEVICE and not handle
    assert b"device" in lib.spk_cuda_last_error().lower()
    from sparkle_b200.device import CudaContext
    with pytest.raises(A.SpkError) as e:
        CudaContext(64, 64)
    assert e.value.code == A.ERR_NO_DEVICE


def test_bad_arguments_are_rejected_without_a_device():
    lib = A.cuda_lib()
    handle = A.CtxP()
    assert lib.spk_cuda_create(None, C.byref(handle)) == A.ERR_INVALID_ARGUMENT
    bad = A.CreateInfo(A.ABI_VERSION + 7, 0, 64, 64)
    assert lib.spk_cuda_create(C.byref(bad), C.byref(handle)) == A.ERR_INVALID_ARGUMENT
    zero = A.CreateInfo(A.ABI_VERSION, 0, 0, 64)
    assert lib.spk_cuda_create(C.byref(zero), C.byref(handle)) == A.ERR_INVALID_ARGUMENT
    assert lib.spk_cuda_render(None, None) == A.ERR_INVALID_ARGUMENT
    assert lib.spk_cuda_destroy(None) in (A.OK, A.ERR_INVALID_ARGUMENT)


def test_product_does_not_touch_the_oracle():
    """only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline/reference legs may use oracle/"""
    pkg = os.path.join(ROOT, "sparkle_b200")
    for dirpath, _, files in os.walk(pkg):
        if "build" in dirpath.split(os.sep):
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp", ".c", "Makefile")):
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "oracle_lib" not in text and "liboracle" not in text and "libsparkle_ref" not in text, \
                    f"{os.path.join(dirpath, f)} references the oracle"
                assert not re.search(r'#include\s+"[^"]*oracle', text), f"{f} includes an oracle header"
    for so in ("libsparkle_cuda.so", "libsparkle_host.so"):
        needed = subprocess.check_output(["readelf", "-d", os.path.join(pkg, so)], text=True)
        assert "oracle" not in needed and "sparkle_ref" not in needed
