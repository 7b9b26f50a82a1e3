"""Procedural scenes for the `cuda` pipeline's tests and benchmarks (all assets of the reference
are absent from the mount, so every scene is generated from fixed seeds).

  testscene()  TestScene-procedural: tests/usd/UsdRoundTripTest.cpp:25-105 minus the two glTF models (replaced by two
               textured procedural meshes at the same transforms) and the HDR sky (gradient sky, SkyRenderProxy.cpp:58-63)
  soup()       C3: 1M-triangle soup, one mesh, identity transform
  city()       C4: 16 base meshes x 160 instances (10M instanced triangles), diffuse + glossy, floor, gradient sky
  room()       C5: closed room, emissive ceiling panel, textured materials, boxes + spheres"""
import math

import numpy as np

from . import _abi as A
from .geometry import Xoshiro128Plus, f32, grid_terrain, triangle_soup, unit_cube, uv_sphere
from .scene import SceneBuilder


def checker_texture(size, c0, c1, cells=8, seed=0):
    """RGBA8 checker with seeded per-texel jitter (so bilinear taps differ)"""
    rng = np.random.default_rng(seed)
    y, x = np.mgrid[0:size, 0:size]
    m = (((x * cells) // size + (y * cells) // size) % 2).astype(bool)
    img = np.zeros((size, size, 4), np.uint8)
    img[..., :3] = np.where(m[..., None], np.array(c1, np.uint8), np.array(c0, np.uint8))
    img[..., :3] = np.clip(img[..., :3].astype(np.int32) + rng.integers(-12, 13, (size, size, 3)), 0, 255)
    img[..., 3] = 255
    return img


def mr_texture(size, seed=0):
    """metallic-roughness (roughness = .y, metallic = .z), smooth seeded pattern"""
    y, x = np.mgrid[0:size, 0:size] / float(size)
    rng = np.random.default_rng(seed)
    a, b = rng.uniform(1, 4, 2)
    rough = 0.25 + 0.6 * (0.5 + 0.5 * np.sin(2 * math.pi * a * x) * np.cos(2 * math.pi * b * y))
    metal = (np.sin(2 * math.pi * (x + y) * 2) > 0.3).astype(np.float64)
    img = np.zeros((size, size, 4), np.uint8)
    img[..., 1] = np.clip(rough * 255, 0, 255)
    img[..., 2] = np.clip(metal * 255, 0, 255)
    img[..., 3] = 255
    return img


def normal_texture(size, seed=0, strength=0.25):
    y, x = np.mgrid[0:size, 0:size] / float(size)
    rng = np.random.default_rng(seed)
    a, b = rng.integers(3, 9, 2)
    nx = strength * np.cos(2 * math.pi * a * x)
    ny = strength * np.cos(2 * math.pi * b * y)
    nz = np.sqrt(np.maximum(0.0, 1 - nx * nx - ny * ny))
    img = np.zeros((size, size, 4), np.uint8)
    img[..., 0] = np.clip((nx * 0.5 + 0.5) * 255, 0, 255)
    img[..., 1] = np.clip((ny * 0.5 + 0.5) * 255, 0, 255)
    img[..., 2] = np.clip((nz * 0.5 + 0.5) * 255, 0, 255)
    img[..., 3] = 255
    return img


def _textured_material(sb, seed, size=256, c0=(200, 60, 40), c1=(230, 220, 200)):
    bc = sb.add_texture(checker_texture(size, c0, c1, seed=seed), A.FMT_RGBA8_SRGB)
    mr = sb.add_texture(mr_texture(size, seed=seed), A.FMT_RGBA8_UNORM)
    nm = sb.add_texture(normal_texture(size, seed=seed), A.FMT_RGBA8_UNORM)
    return sb.add_material(base_color=(1, 1, 1), metallic=1.0, roughness=1.0, base_color_tex=bc,
                           metallic_roughness_tex=mr, normal_tex=nm)


METALS = [(1.0, 0.7, 0.29), (0.56, 0.57, 0.58), (0.95, 0.64, 0.54), (0.92, 0.92, 0.92), (0.95, 0.93, 0.88)]


def _collides(lo, hi, boxes):
    """Scene::BoxCollides -> AABB::Intersect (AABB.h:50-55): strict overlap on every axis"""
    for blo, bhi in boxes:
        c0, h0 = (lo + hi) * f32(0.5), (hi - lo) * f32(0.5)
        c1, h1 = (blo + bhi) * f32(0.5), (bhi - blo) * f32(0.5)
        if np.all(np.abs(c0 - c1) < h0 + h1):
            return True
    return False


def testscene(models=True, max_leaf=4):
    """Returns (SceneBuilder, camera kwargs). Primitive order follows BuildTestScene: floor, glass, lambert, gold,
    [bottle stand-in, boombox stand-in], 10 random spheres from the fixed-seed-42 stream."""
    sb = SceneBuilder(max_leaf=max_leaf)
    white = sb.lambertian((1, 1, 1))
    glass = sb.dielectric(1.6)
    gold = sb.metal(METALS[0])
    cube = sb.add_mesh(unit_cube())
    sb.add_mesh_instance(cube, white, (0, 0, -0.001), (0, 0, 0), (50.0, 50.0, 0.001))
    sb.add_sphere((0, 0, 2), 2.0, glass)
    sb.add_sphere((-4, 4, 2), 2.0, white)
    sb.add_sphere((4, -4, 2), 2.0, gold)
    if models:
        ball = sb.add_mesh(uv_sphere(24, 0.5, "bottle_standin"))
        m0 = _textured_material(sb, 11)
        sb.add_mesh_instance(ball, m0, (-4, -4, 2.7), (0, 0, math.radians(-30.0)), (2, 2, 2))
        m1 = _textured_material(sb, 12, c0=(40, 90, 200), c1=(240, 240, 120))
        sb.add_mesh_instance(cube, m1, (5, 4, 3), (0, 0, math.radians(30.0)), (1.5, 1.2, 0.9))
    # GenerateRandomSpheres (SceneManager.cpp:51-105), one sphere per call, fixed-seed stream
    rng = Xoshiro128Plus(42)
    boxes = [(p["lo"], p["hi"]) for p in sb.prims]
    for _ in range(10):
        radius = f32(f32(0.4) + f32(f32(1.0) - f32(0.4)) * rng.unit())
        kind = rng.unit()
        if kind < f32(0.7):
            color = (rng.unit(), rng.unit(), rng.unit())
            mat = sb.lambertian(color)
        elif kind < f32(0.7) + f32(0.1):
            eta = f32(rng.unit() * f32(2.0) + f32(1.3))
            mat = sb.dielectric(float(eta))
        else:
            mat = sb.metal(METALS[int(f32(rng.unit() * f32(4)))])
        while True:
            x, y = rng.unit_disk()
            pos = np.array([x * f32(13.0), y * f32(13.0), radius], f32)
            lo, hi = pos - radius, pos + radius
            if not _collides(lo, hi, boxes):
                break
        sb.add_sphere(pos, float(radius), mat)
        boxes.append((lo, hi))
    sb.set_sky((0.5, 0.7, 1.0))
    cam = dict(center=(0, 0, 0), radius=25.0, pitch=10.0, yaw=-20.0)
    return sb, cam


def soup(num_tris=1_000_000, seed=1, max_leaf=4, size=0.01):
    sb = SceneBuilder(max_leaf=max_leaf)
    mat = sb.lambertian((0.8, 0.8, 0.8))
    m = sb.add_mesh(triangle_soup(num_tris, seed, size=size))
    sb.add_mesh_instance(m, mat)
    sb.set_sky((0.5, 0.7, 1.0))
    return sb, dict(center=(0, 0, 0), radius=4.0, pitch=15.0, yaw=30.0)


def terrain(nx=1000, ny=500, max_leaf=4):
    sb = SceneBuilder(max_leaf=max_leaf)
    mat = sb.lambertian((0.7, 0.75, 0.6))
    m = sb.add_mesh(grid_terrain(nx, ny))
    sb.add_mesh_instance(m, mat)
    sb.set_sky((0.5, 0.7, 1.0))
    return sb, dict(center=(0, 0, 0), radius=3.0, pitch=30.0, yaw=20.0)


def city(num_meshes=16, tris_per_mesh=62_500, num_instances=160, seed=5, max_leaf=4):
    """C4 'city-10M': base meshes are bumpy towers (grid terrain wrapped as a block), instanced with seeded TRS
    (uniform scale). 50% diffuse / 50% glossy materials, floor, gradient sky."""
    rng = np.random.default_rng(seed)
    sb = SceneBuilder(max_leaf=max_leaf)
    n = int(round(math.sqrt(tris_per_mesh / 2)))
    nx, ny = n, max(1, tris_per_mesh // (2 * n))
    mesh_ids = []
    for k in range(num_meshes):
        mesh_ids.append(sb.add_mesh(grid_terrain(nx, ny, size=2.0, height=0.5 + 0.1 * k, seed=100 + k,
                                                 name=f"block{k}")))
    mats = []
    for k in range(32):
        col = rng.uniform(0.2, 0.95, 3)
        if k % 2 == 0:
            mats.append(sb.lambertian(tuple(col)))
        else:
            mats.append(sb.metal(tuple(col), roughness=float(rng.uniform(0.2, 0.6))))
    floor_mat = sb.lambertian((0.8, 0.8, 0.8))
    cube = sb.add_mesh(unit_cube())
    sb.add_mesh_instance(cube, floor_mat, (0, 0, -0.01), (0, 0, 0), (60.0, 60.0, 0.01))
    side = int(math.ceil(math.sqrt(num_instances)))
    for i in range(num_instances):
        gx, gy = i % side, i // side
        s = float(rng.uniform(1.2, 2.2))
        pos = ((gx - side / 2 + 0.5) * 5.0 + rng.uniform(-0.5, 0.5), (gy - side / 2 + 0.5) * 5.0 + rng.uniform(-0.5, 0.5),
               0.6 * s + float(rng.uniform(0, 2.0)))
        euler = (float(rng.uniform(-0.3, 0.3)), float(rng.uniform(-0.3, 0.3)), float(rng.uniform(0, 2 * math.pi)))
        sb.add_mesh_instance(mesh_ids[i % num_meshes], mats[int(rng.integers(0, len(mats)))], pos, euler, (s, s, s))
    sb.set_sky((0.5, 0.7, 1.0))
    return sb, dict(center=(0, 0, 1.0), radius=55.0, pitch=25.0, yaw=-30.0)


def room(num_objects=50, seed=9, tex=256, max_leaf=4):
    """C5 closed room 10x10x5 (inside of 6 thin slabs), emissive ceiling panel, textured boxes and spheres"""
    rng = np.random.default_rng(seed)
    sb = SceneBuilder(max_leaf=max_leaf)
    cube = sb.add_mesh(unit_cube())
    wall_mats = [_textured_material(sb, 20 + k, tex, c0=tuple(rng.integers(60, 200, 3)), c1=(235, 235, 225))
                 for k in range(3)]
    t = 0.05
    walls = [((0, 0, -t), (5 + t, 5 + t, t)), ((0, 0, 5 + t), (5 + t, 5 + t, t)), ((-5 - t, 0, 2.5), (t, 5 + t, 2.5 + t)),
             ((5 + t, 0, 2.5), (t, 5 + t, 2.5 + t)), ((0, -5 - t, 2.5), (5 + t, t, 2.5 + t)),
             ((0, 5 + t, 2.5), (5 + t, t, 2.5 + t))]
    for k, (pos, scl) in enumerate(walls):
        sb.add_mesh_instance(cube, wall_mats[k % 3], pos, (0, 0, 0), scl)
    light = sb.add_material(base_color=(1, 1, 1), emissive=(10, 10, 10), metallic=0.0, roughness=1.0)
    sb.add_mesh_instance(cube, light, (0, 0, 4.98), (0, 0, 0), (1.5, 1.5, 0.01))
    ball = sb.add_mesh(uv_sphere(16, 1.0, "ball"))
    obj_mats = [_textured_material(sb, 40 + k, tex, c0=tuple(rng.integers(30, 255, 3)), c1=tuple(rng.integers(30, 255, 3)))
                for k in range(6)]
    glass = sb.dielectric(1.5)
    for i in range(num_objects):
        s = float(rng.uniform(0.25, 0.6))
        pos = (float(rng.uniform(-4, 4)), float(rng.uniform(-4, 4)), s + float(rng.uniform(0, 2.5)))
        kind = i % 3
        if kind == 0:
            sb.add_mesh_instance(cube, obj_mats[i % 6], pos, tuple(rng.uniform(0, 1.0, 3)), (s, s, s))
        elif kind == 1:
            sb.add_mesh_instance(ball, obj_mats[(i + 1) % 6], pos, tuple(rng.uniform(0, 1.0, 3)), (s, s, s))
        else:
            sb.add_sphere(pos, s, glass if i % 2 else obj_mats[i % 6])
    # closed: no sky light at all
    return sb, dict(center=(0, 0, 2.2), radius=4.5, pitch=5.0, yaw=35.0, focal_length=0.02)
