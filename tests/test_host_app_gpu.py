"""GPU tests of the C++ host side (sparkle_b200/host/cuda_renderer.cpp) through its C shims: the RenderFramework-style
frame loop (Tick -> IsReadyForAutoScreenshot -> Render) on the `cuda` pipeline, checked against the oracle on the scene
and camera the C++ side built. Mirrors the reference's functional tests: cpu_render_static (accumulate to max_spp, take
the screenshot) and tests/scene/PrimitiveMoveTest.cpp (moving a primitive restarts accumulation and changes the image)."""
import os

import numpy as np
import pytest

from sparkle_b200 import _abi as A
from sparkle_b200 import app, scenes
from sparkle_b200.scene import orbit_camera

pytestmark = pytest.mark.gpu


def rel_err(gpu, ref):
    g, r = gpu[..., :3].astype(np.float64), ref[..., :3].astype(np.float64)
    return float((np.abs(g - r).sum(-1) / np.maximum(r.sum(-1), 1e-3)).mean())


def test_frame_loop_matches_reference_accumulation(oracle, tmp_path):
    w, h, spp = 96, 64, 6
    a = app.HostApp(f"--pipeline CUDA --width {w} --height {h} --max_spp {spp} --spp 1 --bounce 8 --spatial_denoise false")
    frames = a.run(100)
    # one sample per frame; the camera's cumulated count advances in the Update of the NEXT frame (CameraRenderProxy.cpp:66-72)
    assert frames == spp and a.cumulated_samples == spp
    gpu = a.read_linear()
    ref, cum = oracle.render(a.scene, a.camera, w, h, 0, spp, max_bounce=8, accum_mode=A.ACCUM_RUNNING_AVERAGE)
    assert cum == spp and rel_err(gpu, ref) <= 1e-3
    # the C++ TestScene equals the Python recipe: same sphere placement from the fixed-seed-42 stream, same camera
    sb, camkw = scenes.testscene(models=False)
    pdesc = sb.finalize()
    assert a.scene.num_primitives == pdesc.num_primitives == 14
    for k in range(14):
        cp, pp = a.scene.primitives[k], pdesc.primitives[k]
        assert cp.kind == pp.kind and np.allclose(list(cp.fwd), list(pp.fwd), rtol=1e-6, atol=1e-6)
        assert np.float32(cp.radius2) == np.float32(pp.radius2)
    pcam = orbit_camera(w, h, **camkw)
    for f in ("position", "front", "max_u", "max_v", "lower_left"):
        assert np.allclose(list(getattr(a.camera, f)), list(getattr(pcam, f)), rtol=1e-5, atol=1e-5), f
    # screenshot: PNG signature, IHDR dimensions
    png = tmp_path / "shot.png"
    a.save_screenshot(str(png))
    data = png.read_bytes()
    assert data[:8] == b"\x89PNG\r\n\x1a\n" and int.from_bytes(data[16:20], "big") == w and int.from_bytes(data[20:24], "big") == h
    a.close()


def test_multiple_samples_per_frame_and_cap(oracle):
    w, h = 64, 48
    a = app.HostApp(f"--pipeline cuda --width {w} --height {h} --max_spp 10 --spp 4 --spatial_denoise false")
    frames = a.run(100)
    assert frames == 3 and a.cumulated_samples == 10  # 4 + 4 + 2: never past max_spp
    ref, _ = oracle.render(a.scene, a.camera, w, h, 0, 10, accum_mode=A.ACCUM_RUNNING_AVERAGE)
    assert rel_err(a.read_linear(), ref) <= 1e-3
    a.close()


def test_primitive_move_restarts_accumulation(oracle):
    """PrimitiveMoveTest: move a sphere after convergence -> accumulation restarts, the image changes and matches the
    oracle on the updated scene (TLAS rebuilt host-side, instances refreshed through spk_cuda_update_instances)"""
    w, h, spp = 96, 64, 4
    a = app.HostApp(f"--pipeline cuda --width {w} --height {h} --max_spp {spp} --spp 1 --spatial_denoise false")
    a.run(100)
    before = a.read_linear()
    a.move_primitive(3, (2.0, -1.0, 3.5))  # the gold sphere
    assert a.frame() == 0 and a.cumulated_samples == 0  # restarted, first new sample dispatched
    a.run(100)
    assert a.cumulated_samples == spp
    after = a.read_linear()
    assert np.abs(after[..., :3] - before[..., :3]).max() > 0.05
    ref, _ = oracle.render(a.scene, a.camera, w, h, 0, spp, accum_mode=A.ACCUM_RUNNING_AVERAGE)
    assert rel_err(after, ref) <= 1e-3
    a.close()


def test_primitive_move_with_device_side_tlas_rebuild(oracle):
    """8f-4: --device_tlas true -> PrimitiveChange::Update rebuilds the TLAS on the device (LBVH over the instances' world
    boxes); the closest hit does not depend on the topology, so the image still matches the oracle on the host's scene"""
    w, h, spp = 96, 64, 4
    a = app.HostApp(f"--pipeline cuda --width {w} --height {h} --max_spp {spp} --spp 1 --spatial_denoise false "
                    "--device_tlas true")
    a.run(100)
    before = a.read_linear()
    a.move_primitive(3, (2.0, -1.0, 3.5))
    a.run(100)
    after = a.read_linear()
    assert np.abs(after[..., :3] - before[..., :3]).max() > 0.05
    ref, _ = oracle.render(a.scene, a.camera, w, h, 0, spp, accum_mode=A.ACCUM_RUNNING_AVERAGE)
    assert rel_err(after, ref) <= 1e-3
    a.close()


def test_device_built_trees_in_the_host_app(oracle):
    """8f-4: --device_bvh true -> every BLAS and the TLAS are built on the device at upload (and rebuilt there on a
    move); closest hits do not depend on the topology (exact-t ties apart), so the image matches the oracle walking the
    host's trees"""
    w, h, spp = 96, 64, 4
    a = app.HostApp(f"--pipeline cuda --width {w} --height {h} --max_spp {spp} --spp 2 --spatial_denoise false "
                    "--device_bvh true")
    a.run(100)
    ref, _ = oracle.render(a.scene, a.camera, w, h, 0, spp, accum_mode=A.ACCUM_RUNNING_AVERAGE)
    assert rel_err(a.read_linear(), ref) <= 1e-3
    a.move_primitive(3, (2.0, -1.0, 3.5))
    a.run(100)
    ref, _ = oracle.render(a.scene, a.camera, w, h, 0, spp, accum_mode=A.ACCUM_RUNNING_AVERAGE)
    assert rel_err(a.read_linear(), ref) <= 1e-3
    a.close()


def test_gltf_scene_through_the_host_app(tmp_path, oracle):
    """8f-1: --scene file.gltf -> GLTFLoader semantics -> CudaRenderer; the image equals the oracle's on the app's own
    scene description"""
    from sparkle_b200 import gltf_export, scenes
    sb, cam = scenes.city(num_meshes=4, tris_per_mesh=800, num_instances=24)
    path = gltf_export.write_gltf(sb, tmp_path / "city.gltf")
    w, h, spp = 96, 64, 4
    a = app.HostApp(f"--pipeline cuda --width {w} --height {h} --max_spp {spp} --spp 1 --spatial_denoise false "
                    f"--scene {path} --gltf_instancing true")
    a.set_orbit(center=(0, 0, 1.0), radius=30.0, pitch=25.0, yaw=-30.0)
    a.run()
    assert a.cumulated_samples == spp
    gpu = a.read_linear()
    ref, _ = oracle.render(a.scene, a.camera, w, h, 0, spp, accum_mode=A.ACCUM_RUNNING_AVERAGE)
    rel = np.abs(gpu[..., :3] - ref[..., :3]).sum(-1) / np.maximum(ref[..., :3].sum(-1), 1e-3)
    assert rel.mean() <= 1e-3, rel.mean()
    assert (gpu[..., :3].std() > 0.01)  # an actual picture, not a constant
    a.close()


def test_gpus_cvar_splits_samples_and_equals_one_gpu(oracle):
    """--gpus 2: one context per GPU in ONE process (local SpkGroup), samples interleaved, frames combined by the fused
    reduce + resolve kernel. The frame equals the 1-GPU frame (float association only) and the CPU path tracer's. On a
    1-GPU box both contexts sit on device 0 (--gpu_devices 0,0): the local group orders its kernels with stream events."""
    import torch
    w, h, spp = 96, 64, 10
    ndev = torch.cuda.device_count()
    devs = "0,1" if ndev >= 2 else "0,0"
    common = f"--pipeline cuda --width {w} --height {h} --max_spp {spp} --spp 4 --spatial_denoise false"
    one = app.HostApp(common)
    assert one.run(100) == 3
    img1 = one.read_linear()
    two = app.HostApp(common + f" --gpus 2 --gpu_devices {devs}")
    assert two.run(100) == 3 and two.cumulated_samples == spp
    img2 = two.read_linear()
    assert np.allclose(img1, img2, rtol=2e-6, atol=2e-6), float(np.abs(img1 - img2).max())
    ref, _ = oracle.render(two.scene, two.camera, w, h, 0, spp, accum_mode=A.ACCUM_RUNNING_AVERAGE)
    assert rel_err(img2, ref) <= 1e-3
    # frames past max_spp add nothing in the multi-GPU form: the frame stays
    two.render_frame()
    two.render_frame()
    assert np.array_equal(two.read_linear(), img2)
    one.close()
    two.close()


def test_usd_scene_through_the_host_app(tmp_path, oracle):
    """8f-2: --scene file.usda -> USDLoader semantics (host/usd_loader.cpp) -> CudaRenderer. The file is TestScene-
    procedural written in the reference exporter's layout (spheres baked to meshes, textures as PNG files, camera and
    dome light in the file); the image equals the oracle's on the app's own scene description and camera."""
    from sparkle_b200 import usd_export
    sb, camkw = scenes.testscene()
    path = usd_export.write_usda(sb, tmp_path / "testscene.usda", camera=camkw)
    w, h, spp = 96, 64, 4
    a = app.HostApp(f"--pipeline cuda --width {w} --height {h} --max_spp {spp} --spp 2 --spatial_denoise false "
                    f"--scene {path}")
    a.run()
    assert a.cumulated_samples == spp and a.scene.num_primitives == len(sb.prims)
    gpu = a.read_linear()
    ref, _ = oracle.render(a.scene, a.camera, w, h, 0, spp, accum_mode=A.ACCUM_RUNNING_AVERAGE)
    assert rel_err(gpu, ref) <= 1e-3
    # the camera in the file is the orbit camera of the Python recipe
    pcam = orbit_camera(w, h, **camkw)
    for f in ("position", "front", "max_u", "max_v", "lower_left"):
        assert np.allclose(list(getattr(a.camera, f)), list(getattr(pcam, f)), rtol=1e-5, atol=2e-5), f
    a.close()


def test_usd_scene_with_a_sky_map_through_the_host_app(tmp_path, oracle):
    """DomeLight with inputs:texture:file: the .hdr is cooked into the cube on the host (host/sky_cook.cpp, bit-equal to the
    reference's SkyLightCookJob in tests/test_sky_cook_cpu.py) and the path tracer lights the scene with it; the image
    equals the CPU path tracer's on the app's own scene description."""
    import hdr_files as H
    from sparkle_b200 import usd_export
    sb, camkw = scenes.testscene(models=False)
    data = H.write_hdr(H.float_to_rgbe(H.make_sky()))
    path = usd_export.write_usda(sb, tmp_path / "sky.usda", camera=camkw, sky_hdr=("sky.hdr", data))
    w, h, spp = 96, 64, 4
    a = app.HostApp(f"--pipeline cuda --width {w} --height {h} --max_spp {spp} --spp 2 --spatial_denoise false "
                    f"--scene {path}")
    a.run()
    assert a.cumulated_samples == spp and a.scene.sky.has_map == 1 and a.scene.sky.faces[0].width == 1024
    gpu = a.read_linear()
    ref, _ = oracle.render(a.scene, a.camera, w, h, 0, spp, accum_mode=A.ACCUM_RUNNING_AVERAGE)
    assert rel_err(gpu, ref) <= 1e-3
    # the map lights the scene: the same file without it (gradient sky of the dome light's colour) is another image
    plain = usd_export.write_usda(sb, tmp_path / "plain.usda", camera=camkw)
    b = app.HostApp(f"--pipeline cuda --width {w} --height {h} --max_spp {spp} --spp 2 --spatial_denoise false "
                    f"--scene {plain}")
    b.run()
    assert rel_err(b.read_linear(), ref) > 0.05
    # --sky_map: the default sky of a scene without one (SceneManager::AddDefaultSky) from a user-supplied .hdr
    c = app.HostApp(f"--pipeline cuda --width {w} --height {h} --max_spp {spp} --spp 2 --spatial_denoise false "
                    f"--scene {plain} --sky_map {tmp_path / 'textures' / 'sky.hdr'}")
    c.run()
    assert c.scene.sky.has_map == 1 and np.array_equal(c.read_linear(), gpu)
    a.close()
    b.close()
    c.close()
