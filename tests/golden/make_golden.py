"""Generates tests/golden/reference_vectors.npz from the REFERENCE's own CPU path tracer (oracle/_ref/libsparkle_ref.so:
the reference translation units compiled in place from /root/reference by oracle/Makefile). Run in the container that
has /root/reference mounted:   python tests/golden/make_golden.py
The vectors pin the C++ port (oracle/liboracle.so) and, through it, the CUDA path on machines without the reference.
TEST INFRASTRUCTURE."""
import ctypes as C
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import golden_cases as G  # noqa: E402
import oracle_lib as O  # noqa: E402


def main():
    if not O.available("reference"):
        raise SystemExit("oracle/_ref/libsparkle_ref.so missing: run `make -C oracle` where /root/reference is mounted")
    orc = O.Oracle("reference")
    assert orc.kind == "reference"
    out = G.evaluate(orc)
    # the reference's scene-graph float math (Transform, glTF node, orbit camera): pins the product's host/ref_math.cpp
    out.update(G.evaluate_host_math(orc.kat_transform, orc.kat_gltf_node, orc.kat_camera))
    path = os.path.join(HERE, "reference_vectors.npz")
    np.savez_compressed(path, **out)
    print(f"wrote {path}: {len(out)} arrays, {os.path.getsize(path) / 1024:.0f} KiB")


if __name__ == "__main__":
    main()
