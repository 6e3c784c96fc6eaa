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

__all__ = ["Mesh", "loadMesh", "icosphere", "box", "torus"]


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
