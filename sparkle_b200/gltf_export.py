"""Writes a SceneBuilder scene of mesh instances as glTF 2.0 (.gltf + .bin + two 2x2 PNG textures): the "instanced glTF
scene" of BASELINE.json configs[3], one node per instance referencing shared meshes, consumed by the C++ loader
(sparkle_b200/host/gltf_loader.cpp) with the reference's GLTFLoader semantics:
  * the loader pre-multiplies every node rotation by +90 degrees about X (Y-up -> Z-up), so nodes are written with
    rotation = Rx(-90) * scene rotation and come back where the scene has them;
  * the loader bakes node.scale into the vertices, so the unscaled mesh is written once and the scale per node;
  * the loader reads material factors only next to a texture, so every material references a white base-colour texture
    (sRGB 255 decodes to exactly 1.0) and a white metallic-roughness texture, which leave the factors unchanged.
Analytic spheres have no glTF form and are rejected."""
import json
import math
import os
import struct
import zlib

import numpy as np

from . import _abi as A


def write_png(path, rgba):
    """8-bit RGBA PNG (zlib-compressed, filter 0)"""
    h, w, _ = rgba.shape
    raw = b"".join(b"\x00" + rgba[y].astype(np.uint8).tobytes() for y in range(h))

    def chunk(tag, body):
        return struct.pack(">I", len(body)) + tag + body + struct.pack(">I", zlib.crc32(tag + body) & 0xFFFFFFFF)

    with open(path, "wb") as f:
        f.write(b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 6, 0, 0, 0)) +
                chunk(b"IDAT", zlib.compress(raw, 6)) + chunk(b"IEND", b""))


def _quat_from_matrix(R):
    """unit quaternion (x, y, z, w) of a rotation matrix"""
    t = R[0, 0] + R[1, 1] + R[2, 2]
    if t > 0:
        s = math.sqrt(t + 1.0) * 2
        q = ((R[2, 1] - R[1, 2]) / s, (R[0, 2] - R[2, 0]) / s, (R[1, 0] - R[0, 1]) / s, 0.25 * s)
    elif R[0, 0] > R[1, 1] and R[0, 0] > R[2, 2]:
        s = math.sqrt(1.0 + R[0, 0] - R[1, 1] - R[2, 2]) * 2
        q = (0.25 * s, (R[0, 1] + R[1, 0]) / s, (R[0, 2] + R[2, 0]) / s, (R[2, 1] - R[1, 2]) / s)
    elif R[1, 1] > R[2, 2]:
        s = math.sqrt(1.0 + R[1, 1] - R[0, 0] - R[2, 2]) * 2
        q = ((R[0, 1] + R[1, 0]) / s, 0.25 * s, (R[1, 2] + R[2, 1]) / s, (R[0, 2] - R[2, 0]) / s)
    else:
        s = math.sqrt(1.0 + R[2, 2] - R[0, 0] - R[1, 1]) * 2
        q = ((R[0, 2] + R[2, 0]) / s, (R[1, 2] + R[2, 1]) / s, 0.25 * s, (R[1, 0] - R[0, 1]) / s)
    return np.array(q, np.float64)


def _euler_matrix(euler):
    rx, ry, rz = [float(a) for a in euler]
    cx, sx, cy, sy, cz, sz = math.cos(rx), math.sin(rx), math.cos(ry), math.sin(ry), math.cos(rz), math.sin(rz)
    Rx = np.array([[1, 0, 0], [0, cx, -sx], [0, sx, cx]])
    Ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
    Rz = np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]])
    return Rx @ Ry @ Rz  # Transform.cpp:62-63 order, as geometry.trs


def write_gltf(sb, path):
    """sb: SceneBuilder whose primitives are all mesh instances added with add_mesh_instance. Returns the .gltf path."""
    path = str(path)
    base = os.path.splitext(os.path.basename(path))[0]
    out_dir = os.path.dirname(path)
    blob = bytearray()
    views, accessors, meshes = [], [], []

    def add(arr, kind, target=None):
        arr = np.ascontiguousarray(arr)
        while len(blob) % 4:
            blob.append(0)
        view = {"buffer": 0, "byteOffset": len(blob), "byteLength": arr.nbytes}
        if target:
            view["target"] = target
        blob.extend(arr.tobytes())
        views.append(view)
        acc = {"bufferView": len(views) - 1, "componentType": 5125 if arr.dtype == np.uint32 else 5126,
               "count": int(arr.shape[0]), "type": kind}
        if kind == "VEC3" and arr.dtype == np.float32:
            acc["min"] = [float(x) for x in arr.min(axis=0)]
            acc["max"] = [float(x) for x in arr.max(axis=0)]
        accessors.append(acc)
        return len(accessors) - 1

    for m in sb.meshes:
        attrs = {"POSITION": add(m.vertices, "VEC3", 34962), "NORMAL": add(m.normals, "VEC3", 34962),
                 "TANGENT": add(m.tangents, "VEC4", 34962)}
        if m.uvs is not None:
            attrs["TEXCOORD_0"] = add(m.uvs, "VEC2", 34962)
        meshes.append({"name": m.name, "primitives": [{"attributes": attrs, "indices": add(m.indices, "SCALAR", 34963),
                                                       "mode": 4}]})
    white = np.full((2, 2, 4), 255, np.uint8)
    write_png(os.path.join(out_dir, base + "_white.png"), white)
    images = [{"uri": base + "_white.png"}]
    textures = [{"source": 0}]
    materials = []
    for mat in sb.materials:
        if mat["type"] != A.MAT_PBR or any(mat[k] != A.INVALID_ID for k in
                                           ("base_color_tex", "emissive_tex", "metallic_roughness_tex", "normal_tex")):
            raise ValueError("write_gltf: factor-only PBR materials only")
        materials.append({"pbrMetallicRoughness": {
            "baseColorFactor": [float(c) for c in mat["base_color"]] + [1.0], "baseColorTexture": {"index": 0},
            "metallicFactor": float(mat["metallic"]), "roughnessFactor": float(mat["roughness"]),
            "metallicRoughnessTexture": {"index": 0}}})
    half = math.sqrt(0.5)
    unrot = np.array([-half, 0.0, 0.0, half])  # Rx(-90 degrees)

    def qmul(a, b):
        return np.array([a[3] * b[0] + a[0] * b[3] + a[1] * b[2] - a[2] * b[1], a[3] * b[1] - a[0] * b[2] + a[1] * b[3] + a[2] * b[0],
                         a[3] * b[2] + a[0] * b[1] - a[1] * b[0] + a[2] * b[3], a[3] * b[3] - a[0] * b[0] - a[1] * b[1] - a[2] * b[2]])

    nodes = []
    for p in sb.prims:
        if p["kind"] != A.PRIM_MESH or "trs" not in p:
            raise ValueError("write_gltf: mesh instances only (analytic spheres have no glTF form)")
        t, e, s = p["trs"]
        q = qmul(unrot, _quat_from_matrix(_euler_matrix(e)))
        nodes.append({"mesh": int(p["mesh"]), "translation": [float(x) for x in t], "rotation": [float(x) for x in q],
                      "scale": [float(x) for x in s]})
        # the material belongs to the glTF primitive, i.e. to the mesh: instances of one mesh with different materials
        # get their own mesh entry pointing at the same accessors
        prim = dict(meshes[p["mesh"]]["primitives"][0])
        if "material" in prim and prim["material"] != p["material"]:
            meshes.append({"name": meshes[p["mesh"]]["name"], "primitives": [dict(prim, material=int(p["material"]))],
                           "_shares": int(p["mesh"])})
            nodes[-1]["mesh"] = len(meshes) - 1
        else:
            meshes[p["mesh"]]["primitives"][0]["material"] = int(p["material"])
    for m in meshes:
        m.pop("_shares", None)
    doc = {"asset": {"version": "2.0", "generator": "sparkle_b200.gltf_export"}, "scene": 0,
           "scenes": [{"nodes": list(range(len(nodes)))}], "nodes": nodes, "meshes": meshes, "materials": materials,
           "textures": textures, "images": images, "accessors": accessors, "bufferViews": views,
           "buffers": [{"uri": base + ".bin", "byteLength": len(blob)}]}
    with open(os.path.join(out_dir, base + ".bin"), "wb") as f:
        f.write(bytes(blob))
    with open(path, "w") as f:
        json.dump(doc, f)
    return path
