"""The C++ drop-in host API (pathtracing_b200/host/PathTracing.hpp): the reference's class names and
constructor signatures over the C-ABI.  CPU tier: the example program compiles against the bundled glm
stand-in (and against the reference's real glm where that tree exists) and fails loudly without a
device.  GPU tier: it renders the Cornell box and the estimators agree."""
import json
import os
import subprocess
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
HOST = ROOT / "pathtracing_b200" / "host"
LIBDIR = ROOT / "pathtracing_b200"
REF_NAMES = ["Scene", "Material", "Primitive", "Light", "LightSampler", "Sampler", "Filter", "Film", "Camera",
             "Integrators", "Model", "Mesh", "Texture", "Medium", "PhaseFunction", "BVH", "Ray", "Interaction", "AABB",
             "Shape", "ResourceManager", "Util", "Random"]


def _compile(out, glm_dir, src=ROOT / "examples" / "cornell_box.cpp", strict=True):
    cmd = ["/usr/bin/g++", "-std=c++20", "-O1"] + (["-Wall", "-Wextra", "-Werror"] if strict else []) + [
           f"-I{ROOT / 'include'}", f"-I{HOST}", f"-I{glm_dir}",
           str(src), f"-L{LIBDIR}", "-lptb200", f"-Wl,-rpath,{LIBDIR}", "-o", str(out)]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]


def test_forwarding_headers_cover_the_reference_include_names():
    for n in REF_NAMES:
        assert (HOST / f"{n}.hpp").read_text().count('#include "PathTracing.hpp"') == 1


def test_example_compiles_with_bundled_glm(tmp_path):
    _compile(tmp_path / "cornell_box", HOST / "glm_min")


@pytest.mark.skipif(not Path("/root/reference/glm/glm.hpp").exists(), reason="reference tree (real glm) not present")
def test_example_compiles_with_the_reference_glm(tmp_path):
    _compile(tmp_path / "cornell_box", "/root/reference")


@pytest.mark.parametrize("glm", ["bundled", "reference"])
def test_shapes_example_compiles(tmp_path, glm):
    if glm == "reference" and not Path("/root/reference/glm/glm.hpp").exists():
        pytest.skip("reference tree (real glm) not present")
    _compile(tmp_path / "shapes", HOST / "glm_min" if glm == "bundled" else "/root/reference", ROOT / "examples" / "shapes.cpp")


@pytest.mark.skipif(not Path("/root/reference/examples/example_1.cpp").exists(), reason="reference tree not present")
@pytest.mark.parametrize("glm", ["bundled", "reference"])
def test_the_reference_example_compiles_unmodified(tmp_path, glm):
    """examples/example_1.cpp of the reference, byte for byte, with the host API's headers standing where the
    reference's own are (it includes "../Model.hpp" ... "../Integrators.hpp"): spheres, quads, an area light,
    a medium boundary, ResourceManager, three integrators, Film::WriteJPG."""
    import shutil
    tree = tmp_path / "tree"
    (tree / "examples").mkdir(parents=True)
    for h in HOST.glob("*.hpp"):
        shutil.copy(h, tree / h.name)
    shutil.copy("/root/reference/examples/example_1.cpp", tree / "examples" / "example_1.cpp")
    _compile(tmp_path / "example_1", HOST / "glm_min" if glm == "bundled" else "/root/reference", tree / "examples" / "example_1.cpp", strict=False)
    import torch
    if not torch.cuda.is_available():
        r = subprocess.run([str(tmp_path / "example_1")], capture_output=True, text=True, cwd=tmp_path)
        assert r.returncode != 0 and "no CPU fallback" in r.stderr


def test_example_fails_loudly_without_a_device(tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a device is present")
    _compile(tmp_path / "cornell_box", HOST / "glm_min")
    r = subprocess.run([str(tmp_path / "cornell_box"), "8", "8", "1"], capture_output=True, text=True, cwd=tmp_path)
    assert r.returncode != 0 and "no CPU fallback" in r.stderr


@pytest.mark.gpu
def test_example_renders_on_the_gpu(tmp_path):
    _compile(tmp_path / "cornell_box", HOST / "glm_min")
    r = subprocess.run([str(tmp_path / "cornell_box"), "128", "128", "64"], capture_output=True, text=True, cwd=tmp_path, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    out = json.loads(r.stdout.strip().splitlines()[-1])
    assert out["samples"] == 128 * 128 * 64
    # NEE+MIS and BSDF-only sampling estimate the same image
    for a, b in zip(out["path_mean"], out["simple_mean"]):
        assert 0.05 < a < 2.0 and abs(a - b) / a < 0.05, out
    assert out["hit"] == 1 and abs(out["hit_t"] - 2.4) < 1e-4 and out["hit_instance"] == 0
    ppm = (tmp_path / "cornell_path.ppm").read_bytes()
    assert ppm.startswith(b"P6\n128 128\n255\n") and len(ppm) == len(b"P6\n128 128\n255\n") + 3 * 128 * 128


def _read_png(path):
    import struct
    import zlib
    import numpy as np
    b = path.read_bytes()
    assert b[:8] == b"\x89PNG\r\n\x1a\n"
    pos, chunks = 8, {}
    while pos < len(b):
        n, typ = struct.unpack(">I4s", b[pos:pos + 8])
        data = b[pos + 8:pos + 8 + n]
        assert struct.unpack(">I", b[pos + 8 + n:pos + 12 + n])[0] == zlib.crc32(typ + data)
        chunks.setdefault(typ, b"")
        chunks[typ] += data
        pos += 12 + n
    w, h, depth, colour = struct.unpack(">IIBB", chunks[b"IHDR"][:10])
    assert (depth, colour) == (8, 2)
    raw = np.frombuffer(zlib.decompress(chunks[b"IDAT"]), np.uint8).reshape(h, 1 + 3 * w)
    assert not raw[:, 0].any()
    return raw[:, 1:].reshape(h, w, 3)


def test_film_png_and_ppm_writers(tmp_path):
    """Film::{WritePNG, WritePPM, WriteJPG} on a synthetic accumulator (host only): valid PNG chunks / CRCs /
    zlib stream incl. the multi-block case (> 65535 raw bytes), top row first, same bytes as the PPM; a JPEG that decodes."""
    import numpy as np
    src = tmp_path / "film.cpp"
    src.write_text('''
#include "Film.hpp"
#include "Filter.hpp"
int main(){
    const int W = 200, H = 150;
    Film film(glm::ivec2 { W, H }, std::make_shared<BoxFilter>());
    std::vector<double> acc(4 * (size_t)W * H);
    for(int y = 0; y < H; y++)for(int x = 0; x < W; x++){
        double* q = &acc[4 * ((size_t)y * W + x)];
        q[3] = 2.0; q[0] = 2.0 * x / W; q[1] = 2.0 * y / H; q[2] = 0.5;
    }
    film.Merge(acc);
    auto id = [](glm::vec3 c){ return c; };
    film.WritePNG("a", id); film.WritePPM("a", id); film.WriteJPG("b", 90, id);
    return 0;
}
''')
    _compile(tmp_path / "film", HOST / "glm_min", src)
    r = subprocess.run([str(tmp_path / "film")], capture_output=True, text=True, cwd=tmp_path)
    assert r.returncode == 0, r.stderr
    png = _read_png(tmp_path / "a.png")
    hdr = b"P6\n200 150\n255\n"
    ppm = np.frombuffer((tmp_path / "a.ppm").read_bytes()[len(hdr):], np.uint8).reshape(150, 200, 3)
    assert np.array_equal(png, ppm)
    assert png[0, :, 1].min() > 250 and png[-1, :, 1].max() == 0        # film row H-1 (top of the image) comes first
    assert png[:, 0, 0].max() == 0 and png[:, -1, 0].min() > 250
    # WriteJPG writes a baseline JPEG of the same picture (tests/test_imageio.py compares it with the reference's encoder)
    from PIL import Image
    jpg = np.array(Image.open(tmp_path / "b.jpg")).astype(int)
    assert jpg.shape == png.shape and np.abs(jpg - png.astype(int)).mean() < 2.0


@pytest.mark.parametrize("glm", ["bundled", "reference"])
def test_motion_blur_example_compiles(tmp_path, glm):
    """AnimatedPrimitive and TransformedPrimitive over GeometricPrimitives (Primitive.hpp:34-66) in the host API"""
    if glm == "reference" and not Path("/root/reference/glm/glm.hpp").exists():
        pytest.skip("reference tree (real glm) not present")
    _compile(tmp_path / "motion_blur", HOST / "glm_min" if glm == "bundled" else "/root/reference", ROOT / "examples" / "motion_blur.cpp")


@pytest.mark.gpu
def test_motion_blur_example_renders_on_the_gpu(tmp_path):
    """examples/motion_blur.cpp: the moving sphere of main.cpp:187-190, an ellipsoid (TransformedPrimitive over a
    sphere) and a transformed quad light"""
    _compile(tmp_path / "motion_blur", HOST / "glm_min", ROOT / "examples" / "motion_blur.cpp")
    r = subprocess.run([str(tmp_path / "motion_blur"), "160", "90", "64", "8"], capture_output=True, text=True, cwd=tmp_path, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    out = json.loads(r.stdout.strip().splitlines()[-1])
    assert out["samples"] == 160 * 90 * 64 and out["shadow_rays"] > 0
    # the sphere's top: y = 0.6 at shutter open, 0.8 at shutter close, seen from y = 3
    assert out["hit0"] == 1 and abs(out["t0"] - 2.4) < 1e-5 and out["hit1"] == 1 and abs(out["t1"] - 2.2) < 1e-5
    # the ellipsoid's long semi-axis is 0.6: a ray from 3 away along it enters at 2.4
    assert out["hit2"] == 1 and abs(out["t2"] - 2.4) < 1e-4 and out["inst2"] == 2
    img = _read_png(tmp_path / "motion_blur.png")
    assert img.shape == (90, 160, 3) and 10 < img.mean() < 240


@pytest.mark.gpu
def test_shapes_example_renders_on_the_gpu(tmp_path):
    """examples/shapes.cpp = the scene of the reference's example_1.cpp: three integrators over spheres, quads,
    an area light and a medium boundary; PNG output."""
    _compile(tmp_path / "shapes", HOST / "glm_min", ROOT / "examples" / "shapes.cpp")
    r = subprocess.run([str(tmp_path / "shapes"), "160", "90", "64", "32"], capture_output=True, text=True, cwd=tmp_path, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    out = json.loads(r.stdout.strip().splitlines()[-1])
    assert out["samples"] == 160 * 90 * 64
    assert out["hit"] == 1 and abs(out["hit_t"] - 1.4) < 1e-5 and out["hit_instance"] == 1
    assert out["path_shadow_rays"] > 0 and out["volpath_shadow_rays"] > 0
    for k in ("simple_mean", "path_mean", "volpath_mean"):
        assert all(0.05 < v < 3.0 for v in out[k]), out
    # BSDF-only and NEE+MIS estimate nearly the same image (they differ only where the medium boundary
    # blocks PathIntegrator's shadow rays, Integrators.cpp:274)
    for a, b in zip(out["simple_mean"], out["path_mean"]):
        assert abs(a - b) / a < 0.1, out
    for name in ("shapes_simple", "shapes_path", "shapes_volpath"):
        img = _read_png(tmp_path / f"{name}.png")
        assert img.shape == (90, 160, 3) and 20 < img.mean() < 240
    sky, floor = _read_png(tmp_path / "shapes_path.png")[:10].mean(axis=(0, 1)), _read_png(tmp_path / "shapes_path.png")[-10:].mean(axis=(0, 1))
    assert sky[2] > sky[0] and abs(float(floor.mean()) - float(sky.mean())) > 5      # blue sky on top, checker floor below


OBJ = """# a cube of quads without normals (smooth normals are generated), a textured floor, an emissive quad
mtllib scene.mtl
v 0 0 0
v 1 0 0
v 1 1 0
v 0 1 0
v 0 0 1
v 1 0 1
v 1 1 1
v 0 1 1
usemtl cube
f 1 4 3 2
f 5 6 7 8
f 1 5 8 4
f 2 3 7 6
f 1 2 6 5
f -5 -1 -2 -6
v -2 -0.01 -2
v 3 -0.01 -2
v 3 -0.01 3
v -2 -0.01 3
vt 0 0
vt 1 0
vt 1 1
vt 0 1
vn 0 1 0
usemtl floor
f 9/1/1 12/4/1 11/3/1 10/2/1
v 0 2.5 0
v 1 2.5 0
v 1 2.5 1
v 0 2.5 1
usemtl lamp
f 13//1 14//1 15//1
f 13//1 15//1 16//1
l 1 2
"""
MTL = """newmtl cube
Kd 0.05 0.05 0.05
Ks 0.0 0.0 0.0
Ni 1.5
newmtl floor
Kd 0.8 0.8 0.8
map_Kd -s 1 1 1 checker.ppm
newmtl lamp
Kd 0 0 0
Ke 12 11 9
"""


def _write_obj(d):
    (d / "scene.obj").write_text(OBJ)
    (d / "scene.mtl").write_text(MTL)
    px = bytes(((x // 4 + y // 4) % 2) * 200 + 30 for y in range(16) for x in range(16) for _ in range(3))
    (d / "checker.ppm").write_bytes(b"P6\n# generated\n16 16\n255\n" + px)


def test_obj_import_model_from_path(tmp_path):
    """Model(path) / ResourceManager::CacheModel<BLAS4>(name, path) (Model.cpp:14-18, ResourceManager.hpp:45-55)
    over the OBJ / MTL reader: one mesh per material, polygons fan-triangulated, identical corners joined, negative
    indices, generated normals, tangents only where there are texture coordinates, PPM map_Kd, Ke emitters."""
    _write_obj(tmp_path)
    _compile(tmp_path / "obj_model", HOST / "glm_min", ROOT / "examples" / "obj_model.cpp")
    r = subprocess.run([str(tmp_path / "obj_model"), str(tmp_path / "scene.obj")], capture_output=True, text=True, cwd=tmp_path)
    assert r.returncode == 0, r.stderr
    out = json.loads(r.stdout)
    assert out["meshes"] == 3 and out["triangles"] == 12 + 2 + 2
    assert out["vertices"] == 8 + 4 + 4                      # cube corners joined across its six faces
    assert out["emissive_meshes"] == 1 and out["meshes_with_tangents"] == 1
    assert out["lo"] == [-2, -0.01, -2] and out["hi"] == [3, 2.5, 3]
    r = subprocess.run([str(tmp_path / "obj_model"), str(tmp_path / "missing.obj")], capture_output=True, text=True, cwd=tmp_path)
    assert r.returncode != 0 and "cannot open" in r.stderr


def test_blas_from_a_primitive_list_and_nested_wrappers(tmp_path):
    """BLAS4(std::vector<GeometricPrimitive>) with the recipe of Model::BuildBlas (BVH.hpp:743, Model.hpp:47-59) and nested
    wrappers (Primitive.hpp:37) in the host API: runs of the list that share mesh / material / emitter / medium become sub-meshes
    in list order (4 grey, 2 red emitters, 2 red, 4 mirror)."""
    _compile(tmp_path / "blas_list", HOST / "glm_min", ROOT / "examples" / "blas_list.cpp")
    r = subprocess.run([str(tmp_path / "blas_list")], capture_output=True, text=True, cwd=tmp_path)
    assert r.returncode == 0, r.stderr
    assert json.loads(r.stdout.replace('"4": 0, "2": 1, "2": 0, "4": 0', '"split": [4, 2, 2, 4]')) == {"list": 12, "sub_meshes": 4, "split": [4, 2, 2, 4]}, r.stdout


@pytest.mark.gpu
def test_blas_list_and_nested_wrappers_on_the_gpu(tmp_path):
    _compile(tmp_path / "blas_list", HOST / "glm_min", ROOT / "examples" / "blas_list.cpp")
    r = subprocess.run([str(tmp_path / "blas_list"), "64", "48", "32"], capture_output=True, text=True, cwd=tmp_path, timeout=300)
    assert r.returncode == 0, r.stderr
    out = r.stdout
    assert '"ids_are_list_positions": 1' in out and '"samples": %d' % (64 * 48 * 32) in out, out


def _write_gltf(d, glb=False):
    """A two-node glTF 2.0 scene: a quad with normals / uv / u16 indices under a translated + scaled node, and a non-indexed
    triangle without normals under a rotated child node; three materials (textured metallic-roughness with MASK alpha and an
    emitter, a transmissive one with a volume -> dielectric + medium, a thin transmissive one)."""
    import base64
    import struct
    import numpy as np
    from PIL import Image
    Image.fromarray((np.arange(8 * 8 * 3).reshape(8, 8, 3) % 256).astype(np.uint8), "RGB").save(d / "base.png")
    quad_p = np.array([[0, 0, 0], [1, 0, 0], [1, 0, 1], [0, 0, 1]], np.float32)
    quad_n = np.array([[0, 1, 0]] * 4, np.float32)
    quad_uv = np.array([[0, 0], [1, 0], [1, 1], [0, 0.25]], np.float32)
    quad_i = np.array([0, 2, 1, 0, 3, 2], np.uint16)
    tri_p = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0]], np.float32)
    blob = b""
    views, accessors = [], []

    def add(arr, ctype, atype, target=None):
        nonlocal blob
        while len(blob) % 4:
            blob += b"\0"
        views.append({"buffer": 0, "byteOffset": len(blob), "byteLength": arr.nbytes})
        acc = {"bufferView": len(views) - 1, "componentType": ctype, "count": len(arr), "type": atype}
        if atype == "VEC3" and ctype == 5126:
            acc["min"], acc["max"] = arr.min(0).tolist(), arr.max(0).tolist()
        accessors.append(acc)
        blob += arr.tobytes()
        return len(accessors) - 1
    a_qp, a_qn, a_quv, a_qi, a_tp = add(quad_p, 5126, "VEC3"), add(quad_n, 5126, "VEC3"), add(quad_uv, 5126, "VEC2"), add(quad_i, 5123, "SCALAR"), add(tri_p, 5126, "VEC3")
    doc = {
        "asset": {"version": "2.0"}, "scene": 0, "scenes": [{"nodes": [0]}],
        "nodes": [{"mesh": 0, "translation": [10, 0, 0], "scale": [2, 1, 3], "children": [1]},
                  {"mesh": 1, "rotation": [0, 0, 0.70710678, 0.70710678], "translation": [0, 5, 0]}],       # +90 degrees about z
        "meshes": [{"primitives": [{"attributes": {"POSITION": a_qp, "NORMAL": a_qn, "TEXCOORD_0": a_quv}, "indices": a_qi, "material": 0},
                                   {"attributes": {"POSITION": a_tp}, "material": 2}]},
                   {"primitives": [{"attributes": {"POSITION": a_tp}, "material": 1}]}],
        "materials": [
            {"pbrMetallicRoughness": {"baseColorFactor": [0.5, 1.0, 0.25, 1.0], "baseColorTexture": {"index": 0}}, "alphaMode": "MASK", "alphaCutoff": 0.3,
             "emissiveFactor": [1.0, 0.5, 0.0], "extensions": {"KHR_materials_emissive_strength": {"emissiveStrength": 4.0}}},
            {"extensions": {"KHR_materials_transmission": {"transmissionFactor": 1.0}, "KHR_materials_ior": {"ior": 1.33},
                            "KHR_materials_volume": {"thicknessFactor": 0.5, "attenuationColor": [0.5, 0.8, 0.9], "attenuationDistance": 2.0}}},
            {"extensions": {"KHR_materials_transmission": {"transmissionFactor": 1.0}, "KHR_materials_volume": {"thicknessFactor": 0.0}}}],
        "textures": [{"source": 0}], "images": [{"uri": "base.png"}],
        "accessors": accessors, "bufferViews": views,
    }
    if glb:
        doc["buffers"] = [{"byteLength": len(blob)}]
        js = json.dumps(doc).encode()
        js += b" " * (-len(js) % 4)
        bin_ = blob + b"\0" * (-len(blob) % 4)
        out = struct.pack("<III", 0x46546C67, 2, 12 + 8 + len(js) + 8 + len(bin_)) + struct.pack("<II", len(js), 0x4E4F534A) + js + struct.pack("<II", len(bin_), 0x004E4942) + bin_
        (d / "scene.glb").write_bytes(out)
        return d / "scene.glb"
    doc["buffers"] = [{"byteLength": len(blob), "uri": "data:application/octet-stream;base64," + base64.b64encode(blob).decode()}]
    (d / "scene.gltf").write_text(json.dumps(doc))
    return d / "scene.gltf"


@pytest.mark.parametrize("glb", [False, True])
def test_gltf_import_model_from_path(tmp_path, glb):
    """Model(path) over the glTF 2.0 reader (GltfImport.hpp): node transforms baked into the vertices (TRS and nested nodes),
    flipped V, generated normals, indexed and non-indexed primitives, .gltf with a data-URI buffer and .glb, the material
    mapping of Model::SetupMaterial for the glTF material model (Model.cpp:123-190) and the volume medium (:343-351)."""
    path = _write_gltf(tmp_path, glb)
    _compile(tmp_path / "obj_model", HOST / "glm_min", ROOT / "examples" / "obj_model.cpp")
    r = subprocess.run([str(tmp_path / "obj_model"), str(path)], capture_output=True, text=True, cwd=tmp_path)
    assert r.returncode == 0, r.stderr
    out = json.loads(r.stdout)
    assert out["meshes"] == 3 and out["triangles"] == 2 + 1 + 1 and out["vertices"] == 4 + 3 + 3
    quad, tri_root, tri_child = out["per_mesh"]
    # node 0: translate (10, 0, 0) * scale (2, 1, 3)
    assert quad["material"] == "diffuse" and quad["emissive"] == 1 and quad["medium"] == 0 and quad["p0"] == [10, 0, 0] and quad["n0"] == [0, 1, 0]
    assert quad["uv0"] == [0, 1]                                   # V flipped
    assert tri_root["material"] == "thin" and tri_root["p0"] == [10, 0, 0]
    assert tri_root["n0"] == [0, 0, 1]                             # generated: the triangle lies in the xy plane
    # node 1 under node 0: rotate 90 degrees about z, translate (0, 5, 0), then the parent's scale and translation
    assert tri_child["material"] == "dielectric" and tri_child["medium"] == 1 and tri_child["p0"] == [10, 5, 0]
    assert out["lo"][0] == pytest.approx(8.0, abs=1e-5) and out["hi"][1] == pytest.approx(6.0, abs=1e-5)    # (0,1,0) -> rotated (-1,0,0) -> scaled (-2,..) + 10
    assert out["emissive_meshes"] == 1 and out["meshes_with_tangents"] == 1


@pytest.mark.gpu
def test_gltf_import_renders_on_the_gpu(tmp_path):
    path = _write_gltf(tmp_path)
    _compile(tmp_path / "obj_model", HOST / "glm_min", ROOT / "examples" / "obj_model.cpp")
    r = subprocess.run([str(tmp_path / "obj_model"), str(path), "64", "64", "16"], capture_output=True, text=True, cwd=tmp_path, timeout=300)
    assert r.returncode == 0, r.stderr
    out = json.loads(r.stdout)
    assert out["samples"] == 64 * 64 * 16 and all(0.01 < v < 50.0 for v in out["mean"])


@pytest.mark.gpu
def test_obj_import_renders_on_the_gpu(tmp_path):
    _write_obj(tmp_path)
    _compile(tmp_path / "obj_model", HOST / "glm_min", ROOT / "examples" / "obj_model.cpp")
    r = subprocess.run([str(tmp_path / "obj_model"), str(tmp_path / "scene.obj"), "96", "96", "32"], capture_output=True, text=True, cwd=tmp_path, timeout=300)
    assert r.returncode == 0, r.stderr
    out = json.loads(r.stdout)
    assert out["samples"] == 96 * 96 * 32
    assert all(0.05 < v < 5.0 for v in out["mean"])
    assert (tmp_path / "obj_model.ppm").stat().st_size > 96 * 96 * 3
