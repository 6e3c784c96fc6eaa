"""
SceneTemplate (src/theia/scene.py:713-935 of the reference) on this package's own scene file readers (glTF 2.0, OBJ):
instances, material names, transformations through a node hierarchy, detector ids and their stride over template
copies - and a scene created from it traced on the oracle.
"""

import base64
import json
import struct

import numpy as np
import pytest

import harness as H
from theia_b200 import _device, material, mesh, scene
from theia_b200.testing import WaterTestModel


def writeGltf(path, *, binary: bool):
    """two meshes (a box with material 'glass', a box with material 'pmt'), one parent node with two children + one root"""
    box = mesh.box(0.5)
    pos = box.vertices[:, :3].astype(np.float32)
    idx = box.indices.astype(np.uint16).reshape(-1)
    blob = pos.tobytes() + idx.tobytes() + b"\0" * (-len(idx.tobytes()) % 4)
    views = [{"buffer": 0, "byteOffset": 0, "byteLength": pos.nbytes}, {"buffer": 0, "byteOffset": pos.nbytes, "byteLength": idx.nbytes}]
    accessors = [{"bufferView": 0, "componentType": 5126, "count": len(pos), "type": "VEC3"},
                 {"bufferView": 1, "componentType": 5123, "count": len(idx), "type": "SCALAR"}]
    prim = lambda m: {"attributes": {"POSITION": 0}, "indices": 1, "material": m}  # noqa: E731
    doc = {
        "asset": {"version": "2.0"}, "scene": 0, "scenes": [{"nodes": [0, 3]}],
        "nodes": [{"name": "string", "translation": [10.0, 0.0, 0.0], "children": [1, 2]},
                  {"name": "module_a", "mesh": 0, "translation": [0.0, 0.0, 2.0], "scale": [2.0, 2.0, 2.0]},
                  {"name": "module_b", "mesh": 1, "translation": [0.0, 0.0, -2.0],
                   "rotation": [0.0, 0.0, float(np.sin(np.pi / 4)), float(np.cos(np.pi / 4))]},
                  {"name": "housing", "mesh": 0, "matrix": [1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, -5, 1, 0, 1]}],
        "meshes": [{"name": "shell", "primitives": [prim(0)]}, {"name": "sensor", "primitives": [prim(1)]}],
        "materials": [{"name": "glass"}, {"name": "pmt"}],
        "accessors": accessors, "bufferViews": views,
    }
    if binary:
        doc["buffers"] = [{"byteLength": len(blob)}]
        js = json.dumps(doc).encode()
        js += b" " * (-len(js) % 4)
        body = struct.pack("<II", len(js), 0x4E4F534A) + js + struct.pack("<II", len(blob), 0x004E4942) + blob
        path.write_bytes(b"glTF" + struct.pack("<II", 2, 12 + len(body)) + body)
    else:
        doc["buffers"] = [{"byteLength": len(blob), "uri": "data:application/octet-stream;base64," + base64.b64encode(blob).decode()}]
        path.write_text(json.dumps(doc))


@pytest.mark.parametrize("binary", [False, True], ids=["gltf", "glb"])
def test_template_from_gltf(tmp_path, binary):
    file = tmp_path / ("scene.glb" if binary else "scene.gltf")
    writeGltf(file, binary=binary)
    with _device.useDevice("cpu"):
        water = WaterTestModel().createMedium()
        glass = material.BK7Model().createMedium()
        store = material.MaterialStore([material.Material("glass", glass, water, flags="TR"),
                                        material.Material("pmt", None, glass, flags="DB")], media=[water, glass])
        temp = scene.SceneTemplate(file, materials=store, detectorMaterial={"pmt"})
        assert set(temp.instances) == {"module_a", "module_b", "housing"}
        a, b, h = temp.instances["module_a"], temp.instances["module_b"], temp.instances["housing"]
        assert (a.meshName, a.material, a.detectorId) == ("shell", "glass", 0)
        assert (b.meshName, b.material, b.detectorId) == ("sensor", "pmt", 1)
        assert (h.meshName, h.material, h.detectorId) == ("shell", "glass", 0)
        # hierarchy: parent translation (10, 0, 0) on top of the children's own transformations
        assert np.allclose(a.transform.apply(np.array([[0.5, 0.0, 0.0]])), [[11.0, 0.0, 2.0]])
        assert np.allclose(b.transform.apply(np.array([[0.5, 0.0, 0.0]])), [[10.0, 0.5, -2.0]])  # 90 degrees about z
        assert np.allclose(h.transform.apply(np.array([[0.0, 0.0, 0.0]])), [[-5.0, 1.0, 0.0]])
        with pytest.raises(RuntimeError):
            t = a.transform
            t @= scene.Transform.Translation(1.0, 0.0, 0.0)  # frozen
        copies = [scene.Transform.Translation(0.0, 0.0, 0.0), scene.Transform.Translation(0.0, 30.0, 0.0)]
        sc, ids = temp.createScene(copies, sceneMedium=store.media["water"])
        assert len(sc.instances) == 6
        assert ids == {("module_b", 0): 1, ("module_b", 1): 2}
        sc2, ids2 = temp.createScene(copies, detectorIdStride=10)
        assert ids2 == {("module_b", 0): 1, ("module_b", 1): 11}
        # default: every instance its own id
        assert [i.detectorId for i in scene.SceneTemplate(file, materials=store).instances.values()] == [1, 2, 3]
        assert scene.SceneTemplate(file, materials=store, detectorIdMap={"housing": 7}).instances["housing"].detectorId == 7
        with pytest.raises(ValueError):
            scene.SceneTemplate(file).createScene()  # no material map


def test_template_from_obj_and_missing_material(tmp_path):
    obj = tmp_path / "two.obj"
    obj.write_text("o left\nusemtl glass\nv 0 0 0\nv 1 0 0\nv 0 1 0\nv 0 0 1\nf 1 2 3\nf 1 2 4\nf 1 3 4\nf 2 3 4\n"
                   "o right\nusemtl pmt\nv 5 0 0\nv 6 0 0\nv 5 1 0\nv 5 0 1\nf 5 6 7 \nf -4 -3 -1\nf 5 7 8\nf 6 7 8\n")
    with _device.useDevice("cpu"):
        temp = scene.SceneTemplate(obj, materials={"glass": 1, "pmt": 2})
        assert [(i.name, i.material, i.detectorId) for i in temp.instances.values()] == [("left", "glass", 1), ("right", "pmt", 2)]
        assert len(temp.meshStore._meshes[1].indices) == 4
    bad = tmp_path / "bad.obj"
    bad.write_text("o thing\nv 0 0 0\nv 1 0 0\nv 0 1 0\nf 1 2 3\n")
    with _device.useDevice("cpu"), pytest.raises(ValueError, match="no material"):
        scene.SceneTemplate(bad)


def test_scene_from_template_traces_on_the_oracle(tmp_path, orc):
    from theia_b200 import light, random, response, trace
    file = tmp_path / "scene.glb"
    writeGltf(file, binary=True)
    with _device.useDevice("cpu"):
        water = WaterTestModel().createMedium()
        glass = material.BK7Model().createMedium()
        store = material.MaterialStore([material.Material("glass", glass, water, flags="TR"),
                                        material.Material("pmt", None, water, flags="DB")], media=[water, glass])
        temp = scene.SceneTemplate(file, materials=store, detectorMaterial={"pmt"})
        sc, ids = temp.createScene(sceneMedium=store.media["water"])
        rec = response.HitRecorder()
        rng = random.PhiloxRNG(key=3)
        src = light.ConeLightSource(position=(10.0, 0.0, 6.0), direction=(0.0, 0.0, -1.0), cosOpeningAngle=0.9, budget=1.0)
        tr = trace.SceneForwardTracer(4096, src, light.ConstWavelengthSource(450.0), rec, rng, scene=sc, maxPathLength=6,
                                      targetId=ids[("module_b", 0)])
        H.runOracle({"tracer": tr, "response": rec, "callback": None, "rng": rng})
        v = rec.queue.view(0)
    n = min(v.count, rec.capacity)
    assert n > 50 and np.all(v["objectId"][:n] == 1)  # light through the glass shell (transmission) onto the sensor below it
