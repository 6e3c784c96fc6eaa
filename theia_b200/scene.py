"""
Scenes: transforms, bounding boxes, mesh store and the acceleration structures (API of the
reference's src/theia/scene.py).

The reference hands BLAS/TLAS construction to the Vulkan driver (``hp.GeometryStore``
scene.py:550, ``hp.AccelerationStructure`` scene.py:668). Here they are built by hand-written
CUDA kernels (LBVH -> compressed 8-wide BVH, csrc/bvh_build.cu) through
``theia_blas_build`` / ``theia_tlas_build``.
"""

from __future__ import annotations

import ctypes
import itertools
from collections.abc import Iterable, Mapping
from ctypes import Structure, byref, c_float, c_uint64
from pathlib import Path

import numpy as np
import torch
from numpy.typing import ArrayLike, NDArray

import theia_b200.units as u
from theia_b200 import _abi, _device, _lib
from theia_b200.glsl import vec3
from theia_b200.material import MaterialStore
from theia_b200.mesh import Mesh, loadMesh

__all__ = ["loadMesh", "Mesh", "MeshInstance", "MeshStore", "RectBBox", "Scene", "SceneTemplate", "SphereBBox", "Transform"]


class Transform:
    """Affine transformation stored as 4x4 float64, exported as 3x4 float32 (scene.py:42-296)."""

    def __init__(self, matrix: NDArray | None = None) -> None:
        self._arr = np.identity(4)
        if matrix is not None:
            matrix = np.asarray_chkfinite(matrix)
            if matrix.shape != (3, 4):
                raise ValueError("matrix must be of shape (3,4)!")
            self._arr[:3, :] = matrix

    def apply(self, points: NDArray) -> NDArray:
        """Transforms points of shape (N,3)"""
        return points @ self._arr[:3, :3].T + self._arr[:3, 3]

    def applyVec(self, vector: NDArray) -> NDArray:
        """Transforms vectors of shape (N,3), ignoring the translation"""
        return vector @ self._arr[:3, :3].T

    def copy(self) -> "Transform":
        return Transform(self._arr[:3, :].copy())

    def freeze(self) -> "Transform":
        self._arr.setflags(write=False)
        return self

    def inverse(self) -> "Transform":
        inv = Transform()
        inv._arr = np.linalg.inv(self._arr)
        return inv

    def numpy(self) -> NDArray:
        """3x4 row major float32 matrix as used by mesh instances"""
        return np.ascontiguousarray(self._arr[:3, :], dtype=np.float32)

    @property
    def innerMatrix(self) -> NDArray:
        return self.numpy()[:3, :3]

    @property
    def offset(self) -> NDArray:
        return self.numpy()[:3, 3]

    @staticmethod
    def Rotation(dx: float, dy: float, dz: float, angle: float) -> "Transform":
        """Rotation about the axis (dx,dy,dz) by ``angle`` DEGREES (Rodrigues' formula)."""
        axis = np.array([dx, dy, dz], dtype=np.float64)
        axis /= np.sqrt(np.square(axis).sum())
        x, y, z = axis
        K = np.array([[0.0, -z, y], [z, 0.0, -x], [-y, x, 0.0]])
        a = np.deg2rad(angle)
        res = Transform()
        res._arr[:3, :3] += np.sin(a) * K + (1.0 - np.cos(a)) * (K @ K)
        return res

    def rotate(self, dx: float, dy: float, dz: float, angle: float) -> "Transform":
        return Transform.Rotation(dx, dy, dz, angle) @ self

    @staticmethod
    def Scale(x: float, y: float | None = None, z: float | None = None) -> "Transform":
        res = Transform()
        res._arr[0, 0] = x
        res._arr[1, 1] = x if y is None else y
        res._arr[2, 2] = x if z is None else z
        return res

    def scale(self, x: float, y: float | None = None, z: float | None = None) -> "Transform":
        return Transform.Scale(x, y, z) @ self

    @staticmethod
    def Translation(x: float, y: float, z: float) -> "Transform":
        res = Transform()
        res._arr[:3, -1] = (x, y, z)
        return res

    def translate(self, x: float, y: float, z: float) -> "Transform":
        return Transform.Translation(x, y, z) @ self

    @staticmethod
    def TRS(*, scale=1.0, rotate=None, translate=(0.0, 0.0, 0.0)) -> "Transform":
        """Scale, then rotate (axis + angle as in ``Rotation``), then translate."""
        result = Transform()
        result = result.scale(*scale) if type(scale) is tuple else result.scale(scale)
        if rotate is not None:
            result = result.rotate(*rotate)
        return result.translate(*translate)

    @staticmethod
    def View(*, position=(0.0, 0.0, 0.0), direction=(0.0, 0.0, 1.0), up=(0.0, 1.0, 0.0)) -> "Transform":
        """Frame at ``position`` whose z axis is ``direction`` and whose y axis is close to ``up``."""
        position, direction, up = np.array(position), np.array(direction), np.array(up)
        unit = lambda v: v / np.sqrt(np.square(v).sum(-1))
        z = unit(direction)
        x = unit(np.cross(up, direction))
        y = unit(np.cross(z, x))
        return Transform(np.stack([x, y, z, position], -1))

    @staticmethod
    def LookAt(*, position=(0.0, 0.0, 0.0), target=(0.0, 0.0, 1.0), up=(0.0, 1.0, 0.0)) -> "Transform":
        position, target = np.array(position), np.array(target)
        return Transform.View(position=position, direction=(position - target), up=np.array(up))

    def __matmul__(self, other: "Transform") -> "Transform":
        if type(other) is not Transform:
            raise TypeError(other)
        result = Transform()
        result._arr = self._arr @ other._arr
        return result

    def __imatmul__(self, other: "Transform") -> "Transform":
        if type(other) is not Transform:
            raise TypeError(other)
        if not self._arr.flags.writeable:
            raise RuntimeError("Tried to change frozen Transform!")
        self._arr = other._arr @ self._arr
        return self

    def __repr__(self) -> str:
        return "Transform(" + np.array2string(self.numpy(), separator=",", prefix="Transform(", suffix=")") + ")"

    def __str__(self) -> str:
        return np.array_str(self.numpy(), precision=5, suppress_small=True)


class RectBBox:
    """Axis aligned box given by two opposite corners (scene.py:299-378)."""

    class GLSL(Structure):
        _fields_ = [("lowerCorner", vec3), ("upperCorner", vec3)]

    def __init__(self, lowerCorner, upperCorner) -> None:
        self._glsl = self.GLSL()
        self.lowerCorner = lowerCorner
        self.upperCorner = upperCorner

    @property
    def glsl(self) -> "RectBBox.GLSL":
        return self._glsl

    @property
    def diagonal(self) -> float:
        d = np.subtract(self.upperCorner, self.lowerCorner)
        return np.sqrt(np.square(d).sum())

    @property
    def lowerCorner(self) -> tuple[float, float, float]:
        return tuple(self._glsl.lowerCorner)

    @lowerCorner.setter
    def lowerCorner(self, value) -> None:
        self._glsl.lowerCorner.assign(value)

    @property
    def upperCorner(self) -> tuple[float, float, float]:
        return tuple(self._glsl.upperCorner)

    @upperCorner.setter
    def upperCorner(self, value) -> None:
        self._glsl.upperCorner.assign(value)

    def transform(self, trafo: Transform) -> "RectBBox":
        """Box around this box after applying ``trafo``"""
        lo, hi = self.lowerCorner, self.upperCorner
        corners = np.array(list(itertools.product(*[(hi[k], lo[k]) for k in range(3)])))
        corners = trafo.apply(corners)
        return RectBBox(tuple(corners.min(0)), tuple(corners.max(0)))


class SphereBBox:
    """Sphere given by centre and radius (scene.py:381-431)."""

    class GLSL(Structure):
        _fields_ = [("center", vec3), ("radius", c_float)]

    def __init__(self, center, radius: float) -> None:
        self._glsl = self.GLSL()
        self.center = center
        self.radius = radius

    @property
    def glsl(self) -> "SphereBBox.GLSL":
        return self._glsl

    @property
    def center(self) -> tuple[float, float, float]:
        return tuple(self._glsl.center)

    @center.setter
    def center(self, value) -> None:
        self._glsl.center.assign(value)

    @property
    def radius(self) -> float:
        return self._glsl.radius

    @radius.setter
    def radius(self, value: float) -> None:
        self._glsl.radius = value


class _GeometryInstance:
    """Stand-in for ``hp.GeometryInstance``: transform + custom index + BLAS reference."""

    def __init__(self, blas: int, meshIndex: int) -> None:
        self.blas = blas
        self.meshIndex = meshIndex
        self.transform = np.ascontiguousarray(np.identity(4)[:3], dtype=np.float32)
        self.customIndex = 0
        self.mask = 0xFF


class MeshInstance:
    """One placed copy of a mesh with a material name (scene.py:452-526)."""

    def __init__(self, instance: _GeometryInstance, vertices: int, indices: int, triangleCount: int, bbox: RectBBox,
                 material: str, vertexCount: int = 0) -> None:
        self._instance = instance
        self._vertices = vertices
        self._indices = indices
        self._triangleCount = triangleCount
        self._vertexCount = vertexCount
        self._localBbox = bbox
        self._bbox = self.localBBox.transform(self.transform)
        self.material = material

    @property
    def bbox(self) -> RectBBox:
        return self._bbox

    @property
    def localBBox(self) -> RectBBox:
        return self._localBbox

    @property
    def instance(self) -> _GeometryInstance:
        return self._instance

    @property
    def vertices(self) -> int:
        """Device address of the vertex data"""
        return self._vertices

    @property
    def indices(self) -> int:
        """Device address of the index data"""
        return self._indices

    @property
    def triangleCount(self) -> int:
        return self._triangleCount

    @property
    def transform(self) -> Transform:
        return Transform(self.instance.transform)

    @transform.setter
    def transform(self, value: ArrayLike) -> None:
        self.instance.transform = np.ascontiguousarray(value, dtype=np.float32).reshape(3, 4)
        self._bbox = self.localBBox.transform(self.transform)


class MeshStore:
    """
    Owns the meshes and their bottom level acceleration structures (scene.py:529-605).

    Parameters
    ----------
    meshes: mapping name -> ``Mesh`` or file path (STL / PLY)
    """

    def __init__(self, meshes: Mapping[str, Mesh | str]) -> None:
        self._device = _device.currentDevice()
        self._keys = list(meshes.keys())
        values = [loadMesh(v) if isinstance(v, (str, Path)) else v for v in meshes.values()]
        self._meshes = values
        self._triangleCounts = [len(m.indices) for m in values]
        self._bbox = [RectBBox(tuple(m.vertices[:, :3].min(0)), tuple(m.vertices[:, :3].max(0))) for m in values]
        # upload geometry, build one BLAS per mesh
        self._vertexTensors, self._indexTensors, self._blasTensors = [], [], []
        for m in values:
            vt = torch.from_numpy(m.vertices.copy()).to(self._device)
            it = torch.from_numpy(m.indices.astype(np.int32)).to(self._device)
            self._vertexTensors.append(vt)
            self._indexTensors.append(it)
            self._blasTensors.append(self._buildBlas(vt, it) if self._device == "cuda" else None)

    def _buildBlas(self, vt: torch.Tensor, it: torch.Tensor) -> torch.Tensor:
        lib = _lib.load()
        sizes = _abi.ASSizes()
        nTris, nVerts = it.shape[0], vt.shape[0]
        _lib.check(lib.theia_blas_sizes(nTris, byref(sizes)))
        result = _device.alloc(sizes.result_bytes, "cuda")
        scratch = _device.alloc(sizes.scratch_bytes, "cuda")
        _lib.check(lib.theia_blas_build(vt.data_ptr(), nVerts, it.data_ptr(), nTris, result.data_ptr(),
                                        sizes.result_bytes, scratch.data_ptr(), sizes.scratch_bytes,
                                        _device.currentStream("cuda")))
        torch.cuda.current_stream().synchronize()
        return result

    @property
    def keys(self) -> list[str]:
        return list(self._keys)

    def mesh(self, key: str) -> Mesh:
        return self._meshes[self._keys.index(key)]

    def createInstance(self, key: str, material: str, transform: Transform | None = None, *, detectorId: int = 0,
                       scale: float | None = None) -> MeshInstance:
        """
        Creates an instance of the mesh ``key``. ``scale`` is the length unit of the vertex
        positions (default 1 m) and is applied before ``transform``.
        """
        idx = self._keys.index(key)
        blas = self._blasTensors[idx]
        inst = MeshInstance(
            _GeometryInstance(0 if blas is None else blas.data_ptr(), idx),
            self._vertexTensors[idx].data_ptr(), self._indexTensors[idx].data_ptr(), self._triangleCounts[idx],
            self._bbox[idx], material, self._vertexTensors[idx].shape[0])
        inst.instance.customIndex = detectorId
        if scale is None:
            scale = 1.0 * u.m
        if scale != 1.0 or transform is not None:
            trafo = Transform.Scale(scale, scale, scale)
            if transform is not None:
                trafo = transform @ trafo
            inst.transform = trafo.numpy()
        inst._store = self  # keeps the geometry alive
        return inst


class Scene:
    """
    Instances + materials + top level acceleration structure (scene.py:608-710).

    Parameters
    ----------
    instances: the mesh instances; their order defines the instance ids
    materials: ``MaterialStore`` or mapping material name -> device address
    medium: device address of the medium surrounding the scene (0 = vacuum)
    bbox: rays leaving this box are stopped (default +-1 km)
    """

    class GLSLGeometry(Structure):
        _fields_ = [("vertices", c_uint64), ("indices", c_uint64), ("material", c_uint64)]

    def __init__(self, instances: Iterable[MeshInstance], materials: MaterialStore | Mapping[str, int], *,
                 medium: int = 0, bbox: RectBBox | None = None) -> None:
        instances = list(instances)
        if len(instances) == 0:
            raise ValueError("No instances given. Scene cannot be empty!")
        self._device = _device.currentDevice()
        self._materialStore = materials if isinstance(materials, MaterialStore) else None
        if isinstance(materials, MaterialStore):
            materials = materials.material
        self._instances = instances
        n = len(instances)
        self._descs = (_abi.InstanceDesc * n)()
        self._vertexCounts = np.zeros(n, dtype=np.uint32)
        self._triangleCounts = np.zeros(n, dtype=np.uint32)
        for k, inst in enumerate(instances):
            d = self._descs[k]
            o2w = np.identity(4)
            o2w[:3] = inst.instance.transform.astype(np.float64)
            w2o = np.linalg.inv(o2w)
            d.obj_to_world[:] = inst.instance.transform.astype(np.float32).reshape(-1).tolist()
            d.world_to_obj[:] = w2o[:3].astype(np.float32).reshape(-1).tolist()
            d.blas = inst.instance.blas
            d.geometry.vertices = inst.vertices
            d.geometry.indices = inst.indices
            if inst.material is not None:
                if inst.material not in materials:
                    raise ValueError(f'Unknown material "{inst.material}"')
                d.geometry.material = materials.get(inst.material)
            else:
                d.geometry.material = 0
            d.custom_index = inst.instance.customIndex
            self._vertexCounts[k] = inst._vertexCount
            self._triangleCounts[k] = inst.triangleCount
        self._tlas = self._buildTlas() if self._device == "cuda" else None
        self.medium = medium
        self.bbox = RectBBox((-1.0 * u.km,) * 3, (1.0 * u.km,) * 3) if bbox is None else bbox

    def _buildTlas(self) -> torch.Tensor:
        lib = _lib.load()
        n = len(self._instances)
        sizes = _abi.ASSizes()
        _lib.check(lib.theia_tlas_sizes(n, byref(sizes)))
        result = _device.alloc(sizes.result_bytes, "cuda")
        scratch = _device.alloc(sizes.scratch_bytes, "cuda")
        _lib.check(lib.theia_tlas_build(self._descs, n, result.data_ptr(), sizes.result_bytes, scratch.data_ptr(),
                                        sizes.scratch_bytes, _device.currentStream("cuda")))
        torch.cuda.current_stream().synchronize()
        return result

    @property
    def bbox(self) -> RectBBox:
        return self._bbox

    @bbox.setter
    def bbox(self, value: RectBBox) -> None:
        self._bbox = value

    @property
    def instances(self) -> list[MeshInstance]:
        return self._instances

    @property
    def medium(self) -> int:
        return self._medium

    @medium.setter
    def medium(self, value: int) -> None:
        self._medium = value

    @property
    def tlas(self) -> torch.Tensor | None:
        """Buffer holding the top level acceleration structure"""
        return self._tlas


class SceneTemplate:
    """
    Template for creating scenes: loads meshes and a partial scene (instances with their transformations and material
    names) from a file and creates complete scenes from copies of it (src/theia/scene.py:713-935). The reference reads the
    file with ``trimesh``; here ``theia_b200.mesh.loadSceneFile`` reads glTF 2.0 (``.gltf`` / ``.glb``) and Wavefront
    ``.obj`` itself.

    Parameters
    ----------
    file: path of the scene file
    materials: MaterialStore | Mapping[str, int] | None
        material map used when creating scenes; if None it must be passed to ``createScene``
    sceneMedium: device address of the medium the template is embedded within
    templateTransform: optional transformation applied to the loaded template
    detectorIdMap: optional map instance name -> detectorId (others get 0); if None every instance gets a unique id from 1
    detectorMaterial: if given (and no detectorIdMap), only instances with one of these materials get a unique id
    """

    from dataclasses import dataclass as _dataclass

    @_dataclass(frozen=True)
    class InstanceInfo:
        name: str            #: instance name
        meshName: str        #: name of the referenced mesh
        transform: Transform  #: base transformation from mesh to template
        material: str        #: name of the referenced material
        detectorId: int      #: detector id assigned to the instance when creating a scene

    def __init__(self, file: str | Path, *, materials: MaterialStore | Mapping[str, int] | None = None, sceneMedium: int = 0,
                 templateTransform: Transform | None = None, detectorIdMap: Mapping[str, int] | None = None,
                 detectorMaterial: set[str] | None = None) -> None:
        from types import MappingProxyType
        from theia_b200.mesh import loadSceneFile
        sceneFile = loadSceneFile(file)
        for meshName, mat in sceneFile.materials.items():
            if mat == "" or mat == "material_0":  # second: trimesh's default name, kept for files written by it
                raise ValueError(f'Mesh "{meshName}" has no material assigned!')
        self._store = MeshStore(dict(sceneFile.meshes))
        instances: dict[str, SceneTemplate.InstanceInfo] = {}
        nextId = 1
        for instanceName, meshName, matrix in sceneFile.instances:
            trafo = Transform(np.asarray(matrix, dtype=np.float64)[:3, :])
            if templateTransform is not None:
                trafo @= templateTransform
            trafo.freeze()
            mat = sceneFile.materials[meshName]
            detId = 0
            if detectorIdMap is not None:
                detId = detectorIdMap.get(instanceName, 0)
            elif detectorMaterial is not None:
                if mat in detectorMaterial:
                    detId = nextId
                    nextId += 1
            else:
                detId = nextId
                nextId += 1
            instances[instanceName] = self.InstanceInfo(instanceName, meshName, trafo, mat, detId)
        self._instances = MappingProxyType(instances)
        self.materials = materials
        self._sceneMedium = sceneMedium
        self._idStride = nextId - 1

    @property
    def materials(self) -> Mapping[str, int] | None:
        """Material map used when creating scenes from this template (None: pass it to ``createScene``)"""
        return self._materials

    @materials.setter
    def materials(self, value: MaterialStore | Mapping[str, int] | None) -> None:
        self._materialStore = value if isinstance(value, MaterialStore) else None
        if isinstance(value, MaterialStore):
            value = value.material
        self._materials = value

    @property
    def meshStore(self) -> MeshStore:
        """MeshStore containing the meshes used by this template"""
        return self._store

    @property
    def instances(self):
        return self._instances

    @property
    def sceneMedium(self) -> int:
        """Device address of the medium the template is embedded within"""
        return self._sceneMedium

    @sceneMedium.setter
    def sceneMedium(self, value: int) -> None:
        self._sceneMedium = value

    def createScene(self, tempInstance: Iterable[Transform] | None = None, *,
                    materials: MaterialStore | Mapping[str, int] | None = None, sceneMedium: int = 0,
                    sceneTransformation: Transform | None = None, sceneBBox: RectBBox | None = None,
                    detectorIdStride: int | None = None) -> tuple[Scene, dict[tuple[str, int], int]]:
        """
        Creates a scene of one copy of the template per transformation in ``tempInstance`` (None: a single untransformed
        copy). Returns the scene and the map (instance name, copy number) -> detectorId; copy ``i`` offsets the detector
        ids of the template by ``i * detectorIdStride`` (default: the smallest stride that keeps them unique).
        """
        if materials is None:
            materials = self._materialStore if self._materialStore is not None else self.materials
        if materials is None:
            raise ValueError("No material mapping was provided!")
        if tempInstance is None:
            tempInstance = [None]
        if detectorIdStride is None:
            detectorIdStride = self._idStride
        sceneInst = []
        detIdMap: dict[tuple[str, int], int] = {}
        for i, trafo in enumerate(tempInstance):
            offset = i * detectorIdStride
            for tempInst in self.instances.values():
                detId = tempInst.detectorId
                if detId != 0:
                    detId += offset
                    detIdMap[(tempInst.name, i)] = detId
                t = tempInst.transform.copy()
                if trafo is not None:
                    t @= trafo
                if sceneTransformation is not None:
                    t @= sceneTransformation
                sceneInst.append(self.meshStore.createInstance(tempInst.meshName, tempInst.material, t, detectorId=detId))
        kw = {} if sceneBBox is None else {"bbox": sceneBBox}
        return Scene(sceneInst, materials, medium=sceneMedium, **kw), detIdMap
