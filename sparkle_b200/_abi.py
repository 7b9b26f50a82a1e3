"""ctypes mirror of include/spk_cuda.h (the C ABI of the `cuda` pipeline) and loaders for the in-tree libraries.

Product code: loads sparkle_b200/libsparkle_cuda.so and libsparkle_host.so and fails loudly when they are missing.
Nothing here touches oracle/ (test infrastructure)."""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
ABI_VERSION = 1
INVALID_ID = 0xFFFFFFFF

OK, ERR_INVALID_ARGUMENT, ERR_NO_DEVICE, ERR_CUDA, ERR_OUT_OF_MEMORY, ERR_NO_SCENE, ERR_UNSUPPORTED = 0, -1, -2, -3, -4, -5, -6
PRIM_MESH, PRIM_SPHERE = 0, 1
MAT_PBR, MAT_DIELECTRIC = 0, 1
FMT_RGBA8_UNORM, FMT_RGBA8_SRGB, FMT_RGBA16F, FMT_RGBA32F = 0, 1, 2, 3
ACCUM_SUM, ACCUM_RUNNING_AVERAGE = 0, 1
NOALLOC_MAX_RAYS = 262144  # SPK_NOALLOC_MAX_RAYS
FLAG_EXACT_TRAVERSAL, FLAG_COUNTERS, FLAG_WAVEFRONT, FLAG_THREAD_PER_RAY, FLAG_AUTO_TRACER = 1, 2, 4, 8, 16
FLAG_SPATIAL_DENOISE = 32
RESOLVE_LINEAR_F32, RESOLVE_TONEMAP_F16, RESOLVE_TONEMAP_SRGB8 = 0, 1, 2
(OPT_MEGAKERNEL_FORM, OPT_WAVEFRONT_SLOTS, OPT_WAVEFRONT_CHUNK, OPT_TRACE_VOTE, OPT_SMEM_CARVEOUT,
 OPT_VERBOSE, OPT_BUILD_MAX_LEAF, OPT_TRACE_SMEM_TOP, OPT_COST_ORDERED_STRIPS, OPT_WAVEFRONT_SORT) = range(10)
BVH_TLAS = 0xFFFFFFFF

u32, i32, f32, u64 = C.c_uint32, C.c_int32, C.c_float, C.c_uint64
fp = C.POINTER(f32)
u32p = C.POINTER(u32)


class BvhNode(C.Structure):
    _fields_ = [("bounds", f32 * 6), ("index", u32), ("pad", u32)]


class BvhDesc(C.Structure):
    _fields_ = [("nodes", C.c_void_p), ("num_nodes", u32), ("node_stride", u32), ("prim_ids", C.c_void_p),
                ("num_prims", u32), ("prim_id_bytes", u32)]


class MeshDesc(C.Structure):
    _fields_ = [("vertices", fp), ("normals", fp), ("tangents", fp), ("uvs", fp), ("indices", u32p),
                ("num_vertices", u32), ("num_faces", u32), ("blas", BvhDesc)]


class PrimitiveDesc(C.Structure):
    _fields_ = [("kind", u32), ("mesh", u32), ("material", u32), ("radius2", f32), ("fwd", f32 * 12), ("inv", f32 * 12)]


class MaterialDesc(C.Structure):
    _fields_ = [("base_color_tex", u32), ("emissive_tex", u32), ("metallic_roughness_tex", u32), ("normal_tex", u32),
                ("metallic", f32), ("roughness", f32), ("eta", f32), ("type", u32), ("base_color", f32 * 3),
                ("pad0", f32), ("emissive", f32 * 3), ("pad1", f32)]


class TextureDesc(C.Structure):
    _fields_ = [("pixels", C.c_void_p), ("width", u32), ("height", u32), ("format", u32), ("pad", u32)]


class SkyDesc(C.Structure):
    _fields_ = [("has_sky", u32), ("has_map", u32), ("color", f32 * 3), ("pad", f32), ("faces", TextureDesc * 6)]


class SceneDesc(C.Structure):
    _fields_ = [("abi_version", u32), ("num_meshes", u32), ("meshes", C.POINTER(MeshDesc)), ("num_primitives", u32),
                ("num_materials", u32), ("primitives", C.POINTER(PrimitiveDesc)),
                ("materials", C.POINTER(MaterialDesc)), ("num_textures", u32), ("pad", u32),
                ("textures", C.POINTER(TextureDesc)), ("tlas", BvhDesc), ("sky", SkyDesc), ("world_extent", f32),
                ("pad2", f32)]


class Camera(C.Structure):
    _fields_ = [("position", f32 * 3), ("up", f32 * 3), ("front", f32 * 3), ("right", f32 * 3), ("max_u", f32 * 3),
                ("max_v", f32 * 3), ("lower_left", f32 * 3), ("aperture_radius", f32), ("exposure", f32)]


class RenderParams(C.Structure):
    _fields_ = [("first_sample", u32), ("sample_count", u32), ("sample_stride", u32), ("max_bounce", u32),
                ("output_limit", f32), ("accum_mode", u32), ("flags", u32), ("cumulated_cap", u32)]


class Ray(C.Structure):
    _fields_ = [("origin", f32 * 3), ("pad0", f32), ("direction", f32 * 3), ("pad1", f32)]


class Hit(C.Structure):
    _fields_ = [("t", f32), ("u", f32), ("v", f32), ("prim", u32), ("face", u32)]


class Stats(C.Structure):
    _fields_ = [("paths", u64), ("rays", u64), ("node_visits", u64), ("tri_tests", u64), ("inst_visits", u64),
                ("sphere_tests", u64), ("last_ms", C.c_double), ("kernel_launches", u64)]


class CreateInfo(C.Structure):
    _fields_ = [("abi_version", u32), ("device", i32), ("width", u32), ("height", u32)]


class BuildInfo(C.Structure):
    _fields_ = [("device_built", u32), ("num_nodes", u32), ("num_prims", u32), ("depth", u32), ("sah_cost", C.c_double),
                ("build_ms", C.c_double)]


CtxP = C.c_void_p
GroupP = C.c_void_p
PEER_HANDLE_BYTES = 192


class PeerHandle(C.Structure):
    _fields_ = [("bytes", C.c_uint8 * PEER_HANDLE_BYTES)]


# every symbol include/spk_cuda.h declares: name -> (restype, argtypes)
CUDA_SYMBOLS = {
    "spk_cuda_last_error": (C.c_char_p, []),
    "spk_cuda_abi_version": (u32, []),
    "spk_cuda_create": (C.c_int, [C.POINTER(CreateInfo), C.POINTER(CtxP)]),
    "spk_cuda_destroy": (C.c_int, [CtxP]),
    "spk_cuda_host_alloc": (C.c_int, [C.POINTER(C.c_void_p), C.c_size_t]),
    "spk_cuda_host_free": (C.c_int, [C.c_void_p]),
    "spk_cuda_set_stream": (C.c_int, [CtxP, C.c_void_p]),
    "spk_cuda_set_option": (C.c_int, [CtxP, u32, C.c_int64]),
    "spk_cuda_upload_scene": (C.c_int, [CtxP, C.POINTER(SceneDesc)]),
    "spk_cuda_update_instances": (C.c_int, [CtxP, C.POINTER(PrimitiveDesc), u32, C.POINTER(BvhDesc)]),
    "spk_cuda_download_bvh": (C.c_int, [CtxP, u32, C.POINTER(BvhNode), u32, u32p, u32p, u32, u32p]),
    "spk_cuda_get_build_info": (C.c_int, [CtxP, u32, C.POINTER(BuildInfo)]),
    "spk_cuda_set_camera": (C.c_int, [CtxP, C.POINTER(Camera)]),
    "spk_cuda_reset_accumulation": (C.c_int, [CtxP]),
    "spk_cuda_render": (C.c_int, [CtxP, C.POINTER(RenderParams)]),
    "spk_cuda_resolve": (C.c_int, [CtxP, u32, u32, C.c_void_p, C.c_size_t]),
    "spk_cuda_tracer_choice": (C.c_int, [CtxP, C.POINTER(C.c_int32), C.POINTER(u32)]),
    "spk_cuda_choose_tracer": (C.c_int, [CtxP, u32, C.POINTER(u32), C.POINTER(C.c_float)]),
    "spk_cuda_accum_device_ptr": (C.c_int, [CtxP, C.POINTER(C.c_void_p), C.POINTER(C.c_size_t)]),
    "spk_cuda_accum_sample_count": (C.c_int, [CtxP, u32p]),
    "spk_cuda_probe_libm": (C.c_int, [CtxP, u32, C.c_void_p, C.c_void_p, u64, C.c_void_p]),
    "spk_cuda_trace_closest": (C.c_int, [CtxP, C.c_void_p, u64, C.c_void_p, u32]),
    "spk_cuda_trace_closest_device": (C.c_int, [CtxP, C.c_void_p, u64, C.c_void_p, u32]),
    "spk_cuda_trace_closest_noalloc": (C.c_int, [CtxP, C.c_void_p, u64, C.c_void_p, u32]),
    "spk_cuda_group_export": (C.c_int, [CtxP, C.POINTER(PeerHandle)]),
    "spk_cuda_group_create": (C.c_int, [CtxP, u32, u32, C.POINTER(PeerHandle), C.POINTER(GroupP)]),
    "spk_cuda_group_create_local": (C.c_int, [C.POINTER(CtxP), u32, C.POINTER(GroupP)]),
    "spk_cuda_group_resolve": (C.c_int, [GroupP, u32, u32, C.c_void_p, C.c_size_t]),
    "spk_cuda_group_set_async": (C.c_int, [GroupP, u32]),
    "spk_cuda_group_destroy": (C.c_int, [GroupP]),
    "spk_cuda_get_stats": (C.c_int, [CtxP, C.POINTER(Stats)]),
    "spk_cuda_reset_stats": (C.c_int, [CtxP]),
    "spk_cuda_synchronize": (C.c_int, [CtxP]),
}

HOST_SYMBOLS = {
    "spk_host_bvh_build": (C.c_void_p, [fp, fp, u32, u32, u32]),
    "spk_host_bvh_build_triangles": (C.c_void_p, [fp, u32p, u32, u32, u32]),
    "spk_host_bvh_free": (None, [C.c_void_p]),
    "spk_host_bvh_num_nodes": (u32, [C.c_void_p]),
    "spk_host_bvh_num_prims": (u32, [C.c_void_p]),
    "spk_host_bvh_depth": (u32, [C.c_void_p]),
    "spk_host_bvh_nodes": (C.c_void_p, [C.c_void_p]),
    "spk_host_bvh_prim_ids": (C.c_void_p, [C.c_void_p]),
    "spk_host_bvh_sah_cost": (C.c_double, [C.c_void_p]),
    # host/cuda_renderer.h: RenderConfig / CudaRenderer / headless frame loop
    "spk_host_app_create": (C.c_void_p, [C.c_char_p, C.c_int]),
    "spk_host_app_destroy": (None, [C.c_void_p]),
    "spk_host_last_error": (C.c_char_p, []),
    "spk_host_app_frame": (C.c_int, [C.c_void_p]),
    "spk_host_app_run": (C.c_int, [C.c_void_p, u32]),
    "spk_host_app_render_frame": (C.c_int, [C.c_void_p]),
    "spk_host_app_scene": (C.POINTER(SceneDesc), [C.c_void_p]),
    "spk_host_app_camera": (C.POINTER(Camera), [C.c_void_p]),
    "spk_host_app_cumulated_samples": (u32, [C.c_void_p]),
    "spk_host_app_read_linear": (C.c_int, [C.c_void_p, fp, C.c_size_t]),
    "spk_host_app_save_screenshot": (C.c_int, [C.c_void_p, C.c_char_p]),
    "spk_host_app_move_primitive": (C.c_int, [C.c_void_p, u32, f32, f32, f32]),
    "spk_host_parse_pipeline": (C.c_int, [C.c_char_p]),
    "spk_host_app_set_orbit": (None, [C.c_void_p, f32, f32, f32, f32, f32, f32]),
    "spk_host_scene_load_gltf": (C.c_void_p, [C.c_char_p, C.c_int]),
    "spk_host_scene_desc": (C.POINTER(SceneDesc), [C.c_void_p]),
    "spk_host_scene_free": (None, [C.c_void_p]),
    "spk_host_usd_probe": (C.c_int, [C.c_char_p, C.POINTER(C.c_uint64)]),
    "spk_host_decode_hdr": (C.c_int, [C.c_char_p, C.c_size_t, u32p, u32p, C.POINTER(C.c_void_p)]),
    "spk_host_free_buffer": (None, [C.c_void_p]),
    "spk_host_cook_sky_cube": (C.c_int, [C.c_void_p, u32, u32, u32, C.c_void_p]),
    "spk_host_scene_load_usd": (C.c_void_p, [C.c_char_p, u32, u32]),
    "spk_host_scene_camera": (C.POINTER(Camera), [C.c_void_p]),
    "spk_host_gltf_probe": (C.c_int, [C.c_char_p, C.c_int, C.POINTER(C.c_uint64)]),
    # host/ref_math.h: Transform / orbit camera in the reference's float operation order
    "spk_host_trs": (None, [fp, fp, fp, fp, fp]),
    "spk_host_gltf_node": (None, [fp, fp, fp, fp]),
    "spk_host_orbit_camera": (None, [fp, f32, f32, f32, f32, f32, f32, f32, u32, u32, C.POINTER(Camera)]),
}

_libs = {}


def _load(name, symbols):
    if name in _libs:
        return _libs[name]
    path = os.path.join(HERE, name)
    if not os.path.exists(path):
        raise RuntimeError(f"{path} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(make -C sparkle_b200). The cuda pipeline has no fallback.")
    lib = C.CDLL(path)  # RTLD_LOCAL: the static cudart inside must not interpose on torch's
    for sym, (res, args) in symbols.items():
        fn = getattr(lib, sym)  # AttributeError if the library does not export what the header declares
        fn.restype = res
        fn.argtypes = args
    _libs[name] = lib
    return lib


def cuda_lib():
    # SPK_CUDA_LIB: an A/B build of the same library next to the default one (kernel experiments only)
    return _load(os.environ.get("SPK_CUDA_LIB", "libsparkle_cuda.so"), CUDA_SYMBOLS)


def host_lib():
    return _load("libsparkle_host.so", HOST_SYMBOLS)


class SpkError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"spk_cuda error {code}: {msg}")
        self.code = code


def check(rc):
    if rc != OK:
        raise SpkError(rc, cuda_lib().spk_cuda_last_error().decode())
