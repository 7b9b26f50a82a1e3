"""SceneBuilder: collects meshes, primitives, materials, textures and a sky, builds the BLAS/TLAS with the host BVH
builder and produces the SpkSceneDesc the C ABI consumes (include/spk_cuda.h). It plays the role of
SceneRenderProxy::Update/UpdateBVH (libraries/source/renderer/proxy/SceneRenderProxy.cpp:195-297) for the headless
driver: primitives keep their registration order (== SpkHit.prim), the TLAS is built over world AABBs
(AABB::TransformTo, core/math/AABB.h:57-78) and the BLAS over per-triangle boxes (MeshRenderProxy.cpp:31-65)."""
import ctypes as C

import numpy as np

from . import _abi as A
from .geometry import MeshData, f32, trs


class HostBvh:
    """bvh::v2-layout tree built by libsparkle_host.so (sparkle_b200/host/bvh_builder.cpp)"""

    def __init__(self, handle):
        lib = A.host_lib()
        if not handle:
            raise RuntimeError("BVH build failed (invalid arguments or depth budget)")
        n, p = lib.spk_host_bvh_num_nodes(handle), lib.spk_host_bvh_num_prims(handle)
        self.depth = lib.spk_host_bvh_depth(handle)
        self.sah = lib.spk_host_bvh_sah_cost(handle)
        nodes = np.ctypeslib.as_array(C.cast(lib.spk_host_bvh_nodes(handle), C.POINTER(C.c_uint32)), (n, 8))
        ids = np.ctypeslib.as_array(C.cast(lib.spk_host_bvh_prim_ids(handle), C.POINTER(C.c_uint32)), (p,))
        self.nodes = np.array(nodes, copy=True)  # (n, 8) uint32 view of {6 float bounds, index, pad}
        self.prim_ids = np.array(ids, copy=True)
        lib.spk_host_bvh_free(handle)

    @staticmethod
    def from_boxes(boxes, max_leaf=4, max_depth=32):
        boxes = np.ascontiguousarray(boxes, f32).reshape(-1, 6)
        return HostBvh(A.host_lib().spk_host_bvh_build(boxes.ctypes.data_as(A.fp), None, len(boxes), max_leaf, max_depth))

    @staticmethod
    def from_mesh(mesh: MeshData, max_leaf=4, max_depth=32):
        return HostBvh(A.host_lib().spk_host_bvh_build_triangles(
            mesh.vertices.ctypes.data_as(A.fp), mesh.indices.ctypes.data_as(A.u32p), mesh.num_faces, max_leaf, max_depth))

    def desc(self):
        d = A.BvhDesc()
        d.nodes = self.nodes.ctypes.data
        d.num_nodes = len(self.nodes)
        d.node_stride = 32
        d.prim_ids = self.prim_ids.ctypes.data
        d.num_prims = len(self.prim_ids)
        d.prim_id_bytes = 4
        return d


class DeviceBvh:
    """a tree read back from the device (spk_cuda_download_bvh) in the bvh::v2 layout: what the kernels traverse"""

    def __init__(self, ctx, mesh):
        lib = ctx.lib
        n, p = A.u32(), A.u32()
        A.check(lib.spk_cuda_download_bvh(ctx.handle, mesh, None, 0, C.byref(n), None, 0, C.byref(p)))
        self.nodes = np.zeros((n.value, 8), np.uint32)
        self.prim_ids = np.zeros(max(1, p.value), np.uint32)[:p.value]
        A.check(lib.spk_cuda_download_bvh(ctx.handle, mesh, C.cast(self.nodes.ctypes.data, C.POINTER(A.BvhNode)), n.value,
                                          C.byref(n), C.cast(self.prim_ids.ctypes.data, A.u32p), p.value, C.byref(p)))
        info = A.BuildInfo()
        A.check(lib.spk_cuda_get_build_info(ctx.handle, mesh, C.byref(info)))
        self.info = {k: getattr(info, k) for k, _ in A.BuildInfo._fields_}
        self.depth, self.sah = info.depth, info.sah_cost

    desc = HostBvh.desc


class SceneBuilder:
    def __init__(self, max_leaf=4, device_build=False):
        """device_build: meshes and the TLAS come WITHOUT trees (SpkBvhDesc.nodes == NULL): spk_cuda_upload_scene builds
        them on the device (csrc/spk_build.cu); adopt_device_trees() reads them back for CPU consumers"""
        self.max_leaf = max_leaf
        self.device_build = device_build
        self.meshes = []      # MeshData
        self.blas = []        # HostBvh
        self.materials = []   # dict
        self.textures = []    # (pixels ndarray, format)
        self.prims = []       # dict(kind, mesh, material, radius2, fwd, inv, lo, hi)
        self.sky = dict(has_sky=0, has_map=0, color=(0.5, 0.7, 1.0), faces=None)
        self.tlas = None
        self._keep = []
        self.desc = None

    # ---- resources ----
    def add_texture(self, pixels, fmt):
        px = np.ascontiguousarray(pixels)
        assert px.ndim == 3 and px.shape[2] == 4
        self.textures.append((px, fmt))
        return len(self.textures) - 1

    def add_material(self, base_color=(1, 1, 1), emissive=(0, 0, 0), metallic=1.0, roughness=1.0, eta=0.0,
                     kind=A.MAT_PBR, base_color_tex=A.INVALID_ID, emissive_tex=A.INVALID_ID,
                     metallic_roughness_tex=A.INVALID_ID, normal_tex=A.INVALID_ID):
        """io/Material.h:7-22 defaults: base_color 1, emissive 0, metallic 1, roughness 1, eta 0"""
        self.materials.append(dict(base_color=base_color, emissive=emissive, metallic=metallic, roughness=roughness,
                                   eta=eta, type=kind, base_color_tex=base_color_tex, emissive_tex=emissive_tex,
                                   metallic_roughness_tex=metallic_roughness_tex, normal_tex=normal_tex))
        return len(self.materials) - 1

    def lambertian(self, color=(1, 1, 1), **kw):  # LambertianMaterial.h: metallic 0, roughness 1
        return self.add_material(base_color=color, metallic=0.0, roughness=1.0, **kw)

    def metal(self, color, roughness=0.2, **kw):  # MetalMaterial.h: metallic 1, roughness 0.2
        return self.add_material(base_color=color, metallic=1.0, roughness=roughness, **kw)

    def dielectric(self, eta, color=(1, 1, 1)):
        return self.add_material(base_color=color, eta=eta, kind=A.MAT_DIELECTRIC)

    def add_mesh(self, mesh: MeshData):
        self.meshes.append(mesh)
        self.blas.append(None if self.device_build else HostBvh.from_mesh(mesh, self.max_leaf))
        return len(self.meshes) - 1

    # ---- primitives (registration order == SpkHit.prim) ----
    def add_mesh_instance(self, mesh_id, material, translate=(0, 0, 0), euler=(0, 0, 0), scale=(1, 1, 1)):
        fwd, inv = trs(translate, euler, scale)
        lo, hi = self.meshes[mesh_id].bounds()
        # AABB::TransformTo: transform the 8 corners, take min/max
        corners = np.array([[x, y, z] for x in (lo[0], hi[0]) for y in (lo[1], hi[1]) for z in (lo[2], hi[2])], f32)
        wc = corners @ fwd[:, :3].T + fwd[:, 3]
        self.prims.append(dict(kind=A.PRIM_MESH, mesh=mesh_id, material=material, radius2=0.0, fwd=fwd, inv=inv,
                               lo=wc.min(axis=0), hi=wc.max(axis=0), trs=(translate, euler, scale)))
        return len(self.prims) - 1

    def add_sphere(self, center, radius, material):
        """SpherePrimitive: translation = centre, uniform scale = radius (SpherePrimitive.cpp:22-38)"""
        fwd, inv = trs(center, (0, 0, 0), (radius,) * 3)
        c = np.asarray(center, f32)
        r = f32(radius)
        self.prims.append(dict(kind=A.PRIM_SPHERE, mesh=0, material=material, radius2=float(r * r), fwd=fwd, inv=inv,
                               lo=c - r, hi=c + r))
        return len(self.prims) - 1

    def set_sky(self, color=(0.5, 0.7, 1.0), faces=None, fmt=A.FMT_RGBA32F):
        self.sky = dict(has_sky=1, has_map=1 if faces is not None else 0, color=color,
                        faces=None if faces is None else [np.ascontiguousarray(f) for f in faces], fmt=fmt)

    # ---- finalize ----
    def world_bounds(self):
        lo = np.min([p["lo"] for p in self.prims], axis=0)
        hi = np.max([p["hi"] for p in self.prims], axis=0)
        return lo, hi

    def primitive_array(self):
        arr = (A.PrimitiveDesc * max(1, len(self.prims)))()
        for i, p in enumerate(self.prims):
            d = arr[i]
            d.kind, d.mesh, d.material, d.radius2 = p["kind"], p["mesh"], p["material"], p["radius2"]
            d.fwd[:] = p["fwd"].reshape(-1).tolist()
            d.inv[:] = p["inv"].reshape(-1).tolist()
        return arr

    def build_tlas(self):
        boxes = np.array([np.concatenate([p["lo"], p["hi"]]) for p in self.prims], f32)
        self.tlas = HostBvh.from_boxes(boxes, max_leaf=1 if len(self.prims) > 1 else 1)
        return self.tlas

    def finalize(self):
        keep = self._keep = []
        d = A.SceneDesc()
        d.abi_version = A.ABI_VERSION
        marr = (A.MeshDesc * max(1, len(self.meshes)))()
        for i, (m, b) in enumerate(zip(self.meshes, self.blas)):
            md = marr[i]
            md.vertices = m.vertices.ctypes.data_as(A.fp)
            md.normals = m.normals.ctypes.data_as(A.fp)
            md.tangents = m.tangents.ctypes.data_as(A.fp)
            md.uvs = m.uvs.ctypes.data_as(A.fp) if m.uvs is not None else None
            md.indices = m.indices.ctypes.data_as(A.u32p)
            md.num_vertices = len(m.vertices)
            md.num_faces = m.num_faces
            md.blas = b.desc() if b is not None else A.BvhDesc()
        d.num_meshes = len(self.meshes)
        d.meshes = marr
        parr = self.primitive_array()
        d.num_primitives = len(self.prims)
        d.primitives = parr
        matarr = (A.MaterialDesc * max(1, len(self.materials)))()
        for i, m in enumerate(self.materials):
            md = matarr[i]
            md.base_color_tex, md.emissive_tex = m["base_color_tex"], m["emissive_tex"]
            md.metallic_roughness_tex, md.normal_tex = m["metallic_roughness_tex"], m["normal_tex"]
            md.metallic, md.roughness, md.eta, md.type = m["metallic"], m["roughness"], m["eta"], m["type"]
            md.base_color[:] = [float(x) for x in m["base_color"]]
            md.emissive[:] = [float(x) for x in m["emissive"]]
        d.num_materials = len(self.materials)
        d.materials = matarr
        tarr = (A.TextureDesc * max(1, len(self.textures)))()
        for i, (px, fmt) in enumerate(self.textures):
            tarr[i].pixels = px.ctypes.data
            tarr[i].height, tarr[i].width = px.shape[0], px.shape[1]
            tarr[i].format = fmt
        d.num_textures = len(self.textures)
        d.textures = tarr
        if self.prims:
            if not self.device_build:
                self.build_tlas()
            if self.tlas is not None:
                d.tlas = self.tlas.desc()
            lo, hi = self.world_bounds()
            d.world_extent = float(max(np.abs(lo).max(), np.abs(hi).max()))
        d.sky.has_sky = self.sky["has_sky"]
        d.sky.has_map = self.sky["has_map"]
        d.sky.color[:] = [float(x) for x in self.sky["color"]]
        if self.sky["has_map"]:
            for f in range(6):
                px = self.sky["faces"][f]
                d.sky.faces[f].pixels = px.ctypes.data
                d.sky.faces[f].height, d.sky.faces[f].width = px.shape[0], px.shape[1]
                d.sky.faces[f].format = self.sky["fmt"]
        keep += [marr, parr, matarr, tarr]
        self.desc = d
        return d

    def adopt_device_trees(self, ctx):
        """after ctx.upload_scene(self.finalize()) of a device_build scene: read every tree back and return a description
        that carries them, so the oracle traverses the topology the device built"""
        self.blas = [DeviceBvh(ctx, m) for m in range(len(self.meshes))]
        self.tlas = DeviceBvh(ctx, A.BVH_TLAS) if self.prims else None
        self.device_build = False
        try:
            return self.finalize_keep_trees()
        finally:
            self.device_build = True

    def finalize_keep_trees(self):
        tlas = self.tlas
        build = self.build_tlas
        self.build_tlas = lambda: tlas  # keep the downloaded TLAS instead of building one on the host
        try:
            return self.finalize()
        finally:
            self.build_tlas = build

    def stats(self):
        return dict(meshes=len(self.meshes), primitives=len(self.prims),
                    unique_triangles=int(sum(m.num_faces for m in self.meshes)),
                    instanced_triangles=int(sum(self.meshes[p["mesh"]].num_faces for p in self.prims
                                                if p["kind"] == A.PRIM_MESH)),
                    blas_nodes=int(sum(len(b.nodes) for b in self.blas)),
                    blas_depth=int(max([b.depth for b in self.blas], default=0)),
                    tlas_nodes=0 if self.tlas is None else len(self.tlas.nodes))


def orbit_camera(width, height, center=(0, 0, 0), radius=25.0, pitch=10.0, yaw=-20.0, focal_length=0.035,
                 sensor_height=0.024, aperture=22.0, exposure=1.0):
    """OrbitCameraComponent::Setup + CameraRenderProxy::{Update,OnTransformDirty}
    (OrbitCameraComponent.cpp:16-30, CameraComponent.cpp:120-128, CameraRenderProxy.cpp:21-33,113-136) -> SpkCamera, by
    the host library in the reference's float operation order (sparkle_b200/host/ref_math.cpp)."""
    cam = A.Camera()
    c = np.ascontiguousarray(center, f32)
    A.host_lib().spk_host_orbit_camera(c.ctypes.data_as(A.fp), radius, pitch, yaw, focal_length, sensor_height, aperture,
                                       exposure, width, height, C.byref(cam))
    return cam
