"""Headless driver over the C++ host (sparkle_b200/host/cuda_renderer.{h,cpp}): RenderConfig with the reference's cvar
names, the procedural TestScene, CudaRenderer and the RenderFramework-style frame loop (Tick -> ready? -> Render,
libraries/source/application/RenderFramework.cpp:146-162). Usage:
    python -m sparkle_b200.app --pipeline cuda --width 1280 --height 720 --max_spp 64 --spp 8 --output out.png
There is no fallback: without a CUDA device creation fails with the library's error text."""
import ctypes as C
import sys

import numpy as np

from . import _abi as A

PIPELINES = {"cpu": 0, "gpu": 1, "forward": 2, "deferred": 3, "cuda": 4}


def parse_pipeline(name):
    """Str2Enum<RenderConfig::Pipeline> (core/Enum.h:20-32): case-insensitive, -1 when unknown"""
    return A.host_lib().spk_host_parse_pipeline(name.encode())


class HostApp:
    def __init__(self, args, device=0):
        self.lib = A.host_lib()
        self.args = args
        self.handle = self.lib.spk_host_app_create(args.encode(), device)
        if not self.handle:
            raise RuntimeError(f"sparkle_cuda host: cannot start with '{args}': {self.lib.spk_host_last_error().decode()}")
        opts = dict(zip(args.split()[0::2], args.split()[1::2]))
        self.width, self.height = int(opts.get("--width", 1280)), int(opts.get("--height", 720))

    def close(self):
        if self.handle:
            self.lib.spk_host_app_destroy(self.handle)
            self.handle = None

    def frame(self):
        return self.lib.spk_host_app_frame(self.handle)

    def render_frame(self):
        """the reference's frame (Tick, Render, ready flag): keeps rendering past max_spp"""
        return self.lib.spk_host_app_render_frame(self.handle)

    def run(self, max_frames=1 << 20):
        return self.lib.spk_host_app_run(self.handle, max_frames)

    @property
    def scene(self):
        return self.lib.spk_host_app_scene(self.handle).contents

    @property
    def camera(self):
        return self.lib.spk_host_app_camera(self.handle).contents

    @property
    def cumulated_samples(self):
        return self.lib.spk_host_app_cumulated_samples(self.handle)

    def read_linear(self):
        out = np.empty((self.height, self.width, 4), np.float32)
        if self.lib.spk_host_app_read_linear(self.handle, out.ctypes.data_as(A.fp), out.size) != 0:
            raise RuntimeError(self.lib.spk_host_last_error().decode())
        return out

    def save_screenshot(self, path):
        if self.lib.spk_host_app_save_screenshot(self.handle, path.encode()) != 0:
            raise RuntimeError(self.lib.spk_host_last_error().decode())

    def move_primitive(self, primitive, translation):
        if self.lib.spk_host_app_move_primitive(self.handle, primitive, *[float(x) for x in translation]) != 0:
            raise IndexError(primitive)


    def set_orbit(self, center=(0, 0, 0), radius=25.0, pitch=10.0, yaw=-20.0):
        """OrbitCameraComponent::Setup; the accumulation restarts on the next frame"""
        self.lib.spk_host_app_set_orbit(self.handle, *[float(x) for x in center], float(radius), float(pitch), float(yaw))


def gltf_probe(path, instancing=False):
    """counters of GLTFLoader-style ingestion of `path` (sparkle_b200/host/gltf_loader.h)"""
    out = (C.c_uint64 * 7)()
    if A.host_lib().spk_host_gltf_probe(str(path).encode(), int(instancing), out) != 0:
        raise RuntimeError(A.host_lib().spk_host_last_error().decode())
    keys = ("nodes", "primitives", "unique_meshes", "materials", "textures", "instanced_triangles", "unique_triangles")
    return dict(zip(keys, [int(v) for v in out]))


def main(argv):
    output = "screenshot.png"
    args = []
    it = iter(argv)
    for a in it:
        if a == "--output":
            output = next(it)
        else:
            args.append(a)
    if "--pipeline" not in args:
        args += ["--pipeline", "cuda"]
    app = HostApp(" ".join(args))
    frames = app.run()
    app.save_screenshot(output)
    print(f"{frames} frames, {app.cumulated_samples} spp -> {output}")
    app.close()


if __name__ == "__main__":
    main(sys.argv[1:])
