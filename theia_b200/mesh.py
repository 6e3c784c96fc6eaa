"""
Triangle meshes: loading (binary/ASCII STL, PLY), vertex welding, vertex normals and a few
procedural shapes for synthetic scenes.

The reference loads meshes through ``trimesh.load_mesh`` (src/theia/scene.py:434-449), which
is not available here. What the tracer needs from it is reproduced: welded vertices, outward
vertex normals, ``float32[N,6]`` (position, normal) + ``uint32[M,3]``. The normals are only
used for the sign of the geometric normal (sh/scene.intersect.glsl:51-58).
"""

from __future__ import annotations

import struct
from pathlib import Path

import numpy as np
from numpy.typing import NDArray

__all__ = ["Mesh", "SceneFile", "loadMesh", "loadSceneFile", "icosphere", "box", "torus"]


class Mesh:
    """Host side triangle mesh (stand-in for ``hephaistos.Mesh``)."""

    def __init__(self, vertices: NDArray | None = None, indices: NDArray | None = None) -> None:
        self.vertices = np.zeros((0, 6), np.float32) if vertices is None else vertices
        self.indices = np.zeros((0, 3), np.uint32) if indices is None else indices

    @property
    def vertices(self) -> NDArray[np.float32]:
        """float32[N, 6]: position, normal"""
        return self._vertices

    @vertices.setter
    def vertices(self, value) -> None:
        v = np.ascontiguousarray(value, dtype=np.float32)
        if v.ndim != 2 or v.shape[1] != 6:
            raise ValueError("vertices must have shape (N, 6)")
        self._vertices = v

    @property
    def indices(self) -> NDArray[np.uint32]:
        """uint32[M, 3]"""
        return self._indices

    @indices.setter
    def indices(self, value) -> None:
        i = np.ascontiguousarray(value, dtype=np.uint32)
        if i.ndim != 2 or i.shape[1] != 3:
            raise ValueError("indices must have shape (M, 3)")
        self._indices = i


def vertexNormals(positions: NDArray, faces: NDArray) -> NDArray:
    """Area weighted vertex normals."""
    p = positions.astype(np.float64)
    fn = np.cross(p[faces[:, 1]] - p[faces[:, 0]], p[faces[:, 2]] - p[faces[:, 0]])
    n = np.zeros_like(p)
    for k in range(3):
        np.add.at(n, faces[:, k], fn)
    length = np.linalg.norm(n, axis=1, keepdims=True)
    length[length == 0.0] = 1.0
    return n / length


def weld(positions: NDArray, faces: NDArray) -> tuple[NDArray, NDArray]:
    """Merges bit-identical vertex positions, keeping first-occurrence order."""
    key = np.ascontiguousarray(positions, dtype=np.float32).view([("", np.float32)] * 3).reshape(-1)
    _, first, inverse = np.unique(key, return_index=True, return_inverse=True)
    order = np.argsort(first)
    rank = np.empty_like(order)
    rank[order] = np.arange(len(order))
    return positions[first[order]], rank[inverse.reshape(-1)][faces]


def fromTriangles(positions: NDArray, faces: NDArray, normals: NDArray | None = None) -> Mesh:
    faces = np.asarray(faces)
    # drop degenerate faces
    ok = (faces[:, 0] != faces[:, 1]) & (faces[:, 1] != faces[:, 2]) & (faces[:, 0] != faces[:, 2])
    faces = faces[ok]
    if normals is None:
        normals = vertexNormals(positions, faces)
    return Mesh(np.concatenate([positions, normals], axis=-1).astype(np.float32), faces.astype(np.uint32))


def _loadSTL(data: bytes) -> Mesh:
    n = struct.unpack("<I", data[80:84])[0] if len(data) >= 84 else 0
    if len(data) == 84 + 50 * n and n > 0:
        rec = np.frombuffer(data, dtype=np.dtype([("n", "<f4", 3), ("v", "<f4", (3, 3)), ("a", "<u2")]), count=n, offset=84)
        tri = rec["v"].reshape(-1, 3)
    else:  # ASCII
        tri = np.array([[float(x) for x in line.split()[1:4]] for line in data.decode(errors="ignore").splitlines()
                        if line.strip().startswith("vertex")], dtype=np.float32)
    faces = np.arange(len(tri)).reshape(-1, 3)
    pos, faces = weld(tri, faces)
    return fromTriangles(pos, faces)


def _loadPLY(data: bytes) -> Mesh:
    end = data.index(b"end_header\n") + len(b"end_header\n")
    header = data[:end].decode().splitlines()
    fmt = [l.split()[1] for l in header if l.startswith("format")][0]
    elements, cur = [], None
    for l in header:
        t = l.split()
        if t and t[0] == "element":
            cur = [t[1], int(t[2]), []]
            elements.append(cur)
        elif t and t[0] == "property" and cur is not None:
            cur[2].append(t[1:])
    types = {"float": "f4", "float32": "f4", "double": "f8", "float64": "f8", "uchar": "u1", "uint8": "u1",
             "char": "i1", "int8": "i1", "short": "i2", "int16": "i2", "ushort": "u2", "uint16": "u2", "int": "i4",
             "int32": "i4", "uint": "u4", "uint32": "u4"}
    verts, faces = None, None
    if fmt == "ascii":
        tokens = data[end:].split()
        pos = 0
        for name, count, props in elements:
            if name == "vertex":
                k = len(props)
                arr = np.array(tokens[pos : pos + k * count], dtype=np.float64).reshape(count, k)
                pos += k * count
                verts = {p[-1]: arr[:, j] for j, p in enumerate(props)}
            elif name == "face":
                fl = []
                for _ in range(count):
                    n = int(tokens[pos])
                    idx = [int(x) for x in tokens[pos + 1 : pos + 1 + n]]
                    pos += 1 + n
                    fl += [[idx[0], idx[j], idx[j + 1]] for j in range(1, n - 1)]
                faces = np.array(fl)
    else:
        e = "<" if "little" in fmt else ">"
        off = end
        for name, count, props in elements:
            if name == "vertex":
                dt = np.dtype([(p[-1], e + types[p[0]]) for p in props])
                arr = np.frombuffer(data, dtype=dt, count=count, offset=off)
                off += dt.itemsize * count
                verts = {n: arr[n].astype(np.float64) for n in arr.dtype.names}
            elif name == "face":
                ct, it = e + types[props[0][1]], e + types[props[0][2]]
                fl = []
                for _ in range(count):
                    n = int(np.frombuffer(data, dtype=ct, count=1, offset=off)[0])
                    off += np.dtype(ct).itemsize
                    idx = np.frombuffer(data, dtype=it, count=n, offset=off)
                    off += np.dtype(it).itemsize * n
                    fl += [[idx[0], idx[j], idx[j + 1]] for j in range(1, n - 1)]
                faces = np.array(fl)
    pos = np.stack([verts["x"], verts["y"], verts["z"]], -1).astype(np.float32)
    normals = None
    if "nx" in verts:
        normals = np.stack([verts["nx"], verts["ny"], verts["nz"]], -1)
    else:
        pos, faces = weld(pos, faces)
    return fromTriangles(pos, faces, normals)


def loadMesh(filepath: str | Path) -> Mesh:
    """Loads an STL or PLY file (reference: ``loadMesh``, src/theia/scene.py:443-449)."""
    path = Path(filepath)
    data = path.read_bytes()
    suffix = path.suffix.lower()
    if suffix == ".stl":
        return _loadSTL(data)
    if suffix == ".ply":
        return _loadPLY(data)
    raise ValueError(f"unsupported mesh format '{suffix}' (STL and PLY are available)")


# ------------------------------------------------------------------ procedural shapes
def icosphere(subdivisions: int = 3, radius: float = 1.0) -> Mesh:
    """Icosphere with 20 * 4^subdivisions triangles (3 -> 1280, the shape of assets/sphere.stl)."""
    t = (1.0 + np.sqrt(5.0)) / 2.0
    v = np.array([[-1, t, 0], [1, t, 0], [-1, -t, 0], [1, -t, 0], [0, -1, t], [0, 1, t], [0, -1, -t], [0, 1, -t],
                  [t, 0, -1], [t, 0, 1], [-t, 0, -1], [-t, 0, 1]], dtype=np.float64)
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    f = np.array([[0, 11, 5], [0, 5, 1], [0, 1, 7], [0, 7, 10], [0, 10, 11], [1, 5, 9], [5, 11, 4], [11, 10, 2],
                  [10, 7, 6], [7, 1, 8], [3, 9, 4], [3, 4, 2], [3, 2, 6], [3, 6, 8], [3, 8, 9], [4, 9, 5], [2, 4, 11],
                  [6, 2, 10], [8, 6, 7], [9, 8, 1]])
    for _ in range(subdivisions):
        verts = list(v)
        cache: dict[tuple[int, int], int] = {}

        def mid(a: int, b: int) -> int:
            key = (min(a, b), max(a, b))
            if key not in cache:
                m = verts[a] + verts[b]
                verts.append(m / np.linalg.norm(m))
                cache[key] = len(verts) - 1
            return cache[key]

        nf = []
        for a, b, c in f:
            ab, bc, ca = mid(a, b), mid(b, c), mid(c, a)
            nf += [[a, ab, ca], [b, bc, ab], [c, ca, bc], [ab, bc, ca]]
        v, f = np.array(verts), np.array(nf)
    return fromTriangles((v * radius).astype(np.float32), f, v)


def box(halfExtent: float = 1.0) -> Mesh:
    """Axis aligned cube with per-face normals (24 vertices, 12 triangles, like assets/cube.ply)."""
    pos, nrm, faces = [], [], []
    for axis in range(3):
        for sign in (1.0, -1.0):
            n = np.zeros(3)
            n[axis] = sign
            a, b = np.zeros(3), np.zeros(3)
            a[(axis + 1) % 3] = 1.0
            b[(axis + 2) % 3] = 1.0
            if sign < 0:
                a, b = b, a
            base = len(pos)
            for s, t in ((-1, -1), (1, -1), (1, 1), (-1, 1)):
                pos.append((n + s * a + t * b) * halfExtent)
                nrm.append(n)
            faces += [[base, base + 1, base + 2], [base, base + 2, base + 3]]
    return fromTriangles(np.array(pos, np.float32), np.array(faces), np.array(nrm))


def torus(major: float = 1.0, minor: float = 0.25, nu: int = 48, nv: int = 16) -> Mesh:
    """Torus around the z axis with 2 * nu * nv triangles (48 x 16 -> 1536)."""
    uu, vv = np.meshgrid(np.arange(nu) * 2 * np.pi / nu, np.arange(nv) * 2 * np.pi / nv, indexing="ij")
    c = np.stack([np.cos(uu), np.sin(uu), np.zeros_like(uu)], -1)
    n = c * np.cos(vv)[..., None] + np.array([0.0, 0.0, 1.0]) * np.sin(vv)[..., None]
    p = c * major + n * minor
    idx = np.arange(nu * nv).reshape(nu, nv)
    i0, i1 = idx, np.roll(idx, -1, 0)
    j0 = (i0, np.roll(i0, -1, 1))
    j1 = (i1, np.roll(i1, -1, 1))
    f = np.concatenate([np.stack([j0[0], j1[0], j1[1]], -1).reshape(-1, 3),
                        np.stack([j0[0], j1[1], j0[1]], -1).reshape(-1, 3)])
    return fromTriangles(p.reshape(-1, 3).astype(np.float32), f, n.reshape(-1, 3))


# ------------------------------------------------------------------ scene files (SceneTemplate)
class SceneFile:
    """
    What ``trimesh.load(file, force="scene")`` provides to the reference's ``SceneTemplate``
    (src/theia/scene.py:757-806): the meshes of a file by name, the material name of each mesh, and the
    instances of the scene graph (name, mesh name, 3x4 transformation to the base frame).
    """

    def __init__(self) -> None:
        self.meshes: dict[str, Mesh] = {}
        self.materials: dict[str, str] = {}   # mesh name -> material name ("" = none assigned)
        self.instances: list[tuple[str, str, NDArray]] = []  # (instance name, mesh name, float64[3, 4])


_GLTF_TYPES = {5120: np.int8, 5121: np.uint8, 5122: np.int16, 5123: np.uint16, 5125: np.uint32, 5126: np.float32}
_GLTF_WIDTH = {"SCALAR": 1, "VEC2": 2, "VEC3": 3, "VEC4": 4, "MAT4": 16}


def _quatMatrix(q) -> NDArray:
    x, y, z, w = (float(v) for v in q)
    return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                     [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                     [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])


def _loadGLTF(path: Path) -> SceneFile:
    """glTF 2.0 (.gltf with external / base64 buffers, or binary .glb): triangle primitives, node hierarchy."""
    import base64
    import json
    raw = path.read_bytes()
    binChunk = None
    if raw[:4] == b"glTF":
        length = struct.unpack("<I", raw[8:12])[0]
        at, doc = 12, None
        while at < length:
            n, kind = struct.unpack("<II", raw[at:at + 8])
            chunk = raw[at + 8:at + 8 + n]
            if kind == 0x4E4F534A:
                doc = json.loads(chunk.decode())
            elif kind == 0x004E4942 and binChunk is None:
                binChunk = chunk
            at += 8 + n + (-n % 4)
        if doc is None:
            raise ValueError(f"{path}: GLB file without JSON chunk")
    else:
        doc = json.loads(raw.decode())
    buffers = []
    for b in doc.get("buffers", []):
        uri = b.get("uri")
        if uri is None:
            buffers.append(binChunk)
        elif uri.startswith("data:"):
            buffers.append(base64.b64decode(uri.split(",", 1)[1]))
        else:
            buffers.append((path.parent / uri).read_bytes())

    def accessor(i: int) -> NDArray:
        a = doc["accessors"][i]
        view = doc["bufferViews"][a["bufferView"]]
        dt, w = np.dtype(_GLTF_TYPES[a["componentType"]]), _GLTF_WIDTH[a["type"]]
        start = view.get("byteOffset", 0) + a.get("byteOffset", 0)
        stride = view.get("byteStride", 0) or dt.itemsize * w
        data = np.frombuffer(buffers[view["buffer"]], dtype=np.uint8, count=(a["count"] - 1) * stride + dt.itemsize * w, offset=start)
        rows = np.lib.stride_tricks.as_strided(data, shape=(a["count"], dt.itemsize * w), strides=(stride, 1))
        return np.ascontiguousarray(rows).view(dt).reshape(a["count"], w)

    out = SceneFile()
    matNames = [m.get("name", "") for m in doc.get("materials", [])]
    meshParts: list[list[str]] = []
    for k, m in enumerate(doc.get("meshes", [])):
        base = m.get("name") or f"mesh_{k}"
        parts = []
        for p, prim in enumerate(m["primitives"]):
            if prim.get("mode", 4) != 4:
                continue  # triangles only
            pos = accessor(prim["attributes"]["POSITION"]).astype(np.float32)
            idx = accessor(prim["indices"]).reshape(-1, 3) if "indices" in prim else np.arange(len(pos)).reshape(-1, 3)
            nrm = accessor(prim["attributes"]["NORMAL"]).astype(np.float32) if "NORMAL" in prim["attributes"] else None
            if nrm is None:
                pos, idx = weld(pos, idx.astype(np.int64))
            name = base if len(m["primitives"]) == 1 else f"{base}_{p}"
            while name in out.meshes:
                name += "_"
            out.meshes[name] = fromTriangles(pos, idx, nrm)
            out.materials[name] = matNames[prim["material"]] if "material" in prim else ""
            parts.append(name)
        meshParts.append(parts)

    nodes = doc.get("nodes", [])

    def local(n) -> NDArray:
        if "matrix" in n:
            return np.array(n["matrix"], dtype=np.float64).reshape(4, 4).T  # column major in the file
        m = np.eye(4)
        m[:3, :3] = _quatMatrix(n.get("rotation", (0, 0, 0, 1))) * np.array(n.get("scale", (1, 1, 1)), dtype=np.float64)[None, :]
        m[:3, 3] = n.get("translation", (0, 0, 0))
        return m

    def walk(i: int, parent: NDArray) -> None:
        n = nodes[i]
        world = parent @ local(n)
        if "mesh" in n:
            for part in meshParts[n["mesh"]]:
                name = n.get("name") or f"node_{i}"
                if len(meshParts[n["mesh"]]) > 1:
                    name = f"{name}/{part}"
                while any(name == inst[0] for inst in out.instances):
                    name += "_"
                out.instances.append((name, part, world[:3, :].copy()))
        for c in n.get("children", []):
            walk(c, world)

    scenes = doc.get("scenes") or [{"nodes": list(range(len(nodes)))}]
    for root in scenes[doc.get("scene", 0)]["nodes"]:
        walk(root, np.eye(4))
    return out


def _loadOBJ(path: Path) -> SceneFile:
    """Wavefront OBJ: every object / group is one mesh and one instance (identity transformation), ``usemtl`` names its material."""
    verts: list[list[float]] = []
    groups: dict[str, list[list[int]]] = {}
    mats: dict[str, str] = {}
    cur, curMat = "default", ""
    for line in path.read_text(errors="ignore").splitlines():
        t = line.split()
        if not t or t[0].startswith("#"):
            continue
        if t[0] == "v":
            verts.append([float(x) for x in t[1:4]])
        elif t[0] in ("o", "g"):
            cur = " ".join(t[1:]) or "default"
        elif t[0] == "usemtl":
            curMat = " ".join(t[1:])
        elif t[0] == "f":
            idx = [int(x.split("/")[0]) for x in t[1:]]
            idx = [i - 1 if i > 0 else len(verts) + i for i in idx]
            faces = groups.setdefault(cur, [])
            mats.setdefault(cur, curMat)
            for k in range(1, len(idx) - 1):  # fan triangulation
                faces.append([idx[0], idx[k], idx[k + 1]])
    out = SceneFile()
    allVerts = np.array(verts, dtype=np.float32).reshape(-1, 3)
    for name, faces in groups.items():
        f = np.array(faces, dtype=np.int64)
        used, inverse = np.unique(f.reshape(-1), return_inverse=True)
        pos, f = weld(allVerts[used], inverse.reshape(-1, 3))
        out.meshes[name] = fromTriangles(pos, f)
        out.materials[name] = mats.get(name, "")
        out.instances.append((name, name, np.eye(4)[:3, :].copy()))
    return out


def loadSceneFile(filepath: str | Path) -> SceneFile:
    """Loads a scene file (glTF 2.0 ``.gltf`` / ``.glb``, Wavefront ``.obj``); a plain mesh file becomes a scene of one instance."""
    path = Path(filepath)
    ext = path.suffix.lower()
    if ext in (".gltf", ".glb"):
        return _loadGLTF(path)
    if ext == ".obj":
        return _loadOBJ(path)
    out = SceneFile()
    out.meshes[path.stem] = loadMesh(path)
    out.materials[path.stem] = ""
    out.instances.append((path.stem, path.stem, np.eye(4)[:3, :].copy()))
    return out
