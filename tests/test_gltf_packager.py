"""the image part of `Packager <file.gltf>` (Gltf.cpp:97-286) - routing of material
images to generate_srgb_texture / generate_normal_map_texture, output names, de-duplication."""
import json
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CLI = os.path.join(ROOT, "rosy_b200", "csrc", "rosy_packager_tex")


def _write_gltf(tmp_path, with_images=True, size=64):
    """4 images; material 0: colour 0, metal-rough 1, normal 2; material 1: colour 0 again (shared),
    normal 2 again, emissive 3 (must be ignored); material 2: colour = image 1 (also a
    metallic image: the later list wins)."""
    from PIL import Image

    from rosy_b200 import synth

    names = ["base color.jpg", "mr.v2.png", "normal.png", "emissive.png"]
    kinds = ["albedo", "metal_rough", "normal", "albedo"]
    if with_images:
        for i, (n, k) in enumerate(zip(names, kinds)):
            im = Image.fromarray(synth.make(k, size, size, i))
            (im.convert("RGB") if n.endswith(".jpg") else im).save(tmp_path / n)
    doc = {
        "asset": {"version": "2.0"},
        "images": [{"uri": n.replace(" ", "%20")} for n in names],
        "samplers": [{}],
        "textures": [{"source": 0, "sampler": 0}, {"source": 1}, {"source": 2}, {"source": 3}, {"sampler": 0}],
        "materials": [
            {"pbrMetallicRoughness": {"baseColorTexture": {"index": 0}, "metallicRoughnessTexture": {"index": 1}},
             "normalTexture": {"index": 2, "scale": 1.0}},
            {"pbrMetallicRoughness": {"baseColorTexture": {"index": 0}}, "normalTexture": {"index": 2},
             "emissiveTexture": {"index": 3}, "occlusionTexture": {"index": 3}},
            {"pbrMetallicRoughness": {"baseColorTexture": {"index": 1}}},
            {"name": "no textures é \"quoted\"", "pbrMetallicRoughness": {"baseColorTexture": {"index": 4}}},
        ],
    }
    p = tmp_path / "scene.gltf"
    p.write_text(json.dumps(doc, indent=1))
    return p, names


def test_gltf_plan_routing_and_names(tmp_path):
    p, names = _write_gltf(tmp_path, with_images=False)
    out = subprocess.run([CLI, "--dry-run", str(p)], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    lines = out.stdout.splitlines()
    assert "[info] num color_images 2" in lines and "[info] num metallic_images 1" in lines
    assert "[info] num normal_map_images 1" in lines
    plan = json.JSONDecoder().raw_decode(out.stdout[out.stdout.index("{"):])[0]
    # rosy_asset::image names: <dir>/<uri up to the first '.'>.dds (Gltf.cpp:102-108)
    assert [os.path.basename(n) for n in plan["images"]] == ["base%20color.dds", "mr.dds", "normal.dds", "emissive.dds"]
    # 6 generate_* calls in the reference (broken dedupe, F14), 3 jobs here; emissive / occlusion never processed
    assert plan["reference_calls"] == 6
    jobs = plan["jobs"]
    assert [(j["image_index"], j["image_type"]) for j in jobs] == [(0, 0), (1, 2), (2, 1)]
    assert [os.path.basename(j["source"]) for j in jobs] == ["base color.jpg", "mr.v2.png", "normal.png"]
    # generate_* writes image_path.replace_extension(".dds") (Packager.cpp:231)
    assert [os.path.basename(j["dds"]) for j in jobs] == ["base color.dds", "mr.v2.dds", "normal.dds"]
    assert [j["uses"] for j in jobs] == [2, 2, 2]


def test_gltf_all_textures_opt_in(tmp_path):
    """--all-textures (N4) also routes emissive (colour) and occlusion (data) images"""
    p, names = _write_gltf(tmp_path, with_images=False)
    out = subprocess.run([CLI, "--dry-run", "--all-textures", str(p)], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    plan = json.JSONDecoder().raw_decode(out.stdout[out.stdout.index("{"):])[0]
    # image 3 is emissive (colour list) and occlusion (data list): the later list wins -> image_type 2
    assert [(j["image_index"], j["image_type"]) for j in plan["jobs"]] == [(0, 0), (1, 2), (3, 2), (2, 1)]
    assert plan["reference_calls"] == 6


def test_gltf_errors(tmp_path):
    bad = tmp_path / "bad.gltf"
    bad.write_text('{"images": [{"bufferView": 0, "mimeType": "image/png"}]}')
    out = subprocess.run([CLI, "--dry-run", str(bad)], capture_output=True, text=True)
    assert out.returncode != 0 and "not a URI datasource" in out.stderr  # Gltf.cpp:213-217
    bad.write_text('{"images": [')
    out = subprocess.run([CLI, "--dry-run", str(bad)], capture_output=True, text=True)
    assert out.returncode != 0 and "failed to load" in out.stderr
    out = subprocess.run([CLI, "--dry-run", str(tmp_path / "missing.gltf")], capture_output=True, text=True)
    assert out.returncode != 0


def test_gltf_reference_sample():
    """the reference tree's own fastgltf sample parses (no images referenced by materials there)"""
    sample = "/root/reference/src/libs/fastgltf/tests/gltf/basic_gltf.gltf"
    if not os.path.exists(sample):
        pytest.skip("reference tree not present")
    out = subprocess.run([CLI, "--dry-run", sample], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    json.JSONDecoder().raw_decode(out.stdout[out.stdout.index("{"):])[0]


@pytest.mark.gpu
@pytest.mark.parametrize("serial", [False, True])
def test_gltf_packager_run(tmp_path, serial, tex, oracle):
    p, names = _write_gltf(tmp_path)
    out = subprocess.run([CLI] + (["--serial"] if serial else []) + [str(p)], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr + out.stdout
    manifest = json.loads((tmp_path / "scene.textures.json").read_text())
    assert len(manifest["jobs"]) == 3
    assert not (tmp_path / "emissive.dds").exists()
    from PIL import Image

    for j in manifest["jobs"]:
        dds = np.fromfile(j["dds"], dtype=np.uint8)
        img = tex.load_image(j["source"])  # PNG and JPEG: the loader is pinned against stb_image separately
        kind = tex.KIND_NORMAL_MAP if j["image_type"] == 1 else (tex.KIND_METALLIC_ROUGHNESS if j["image_type"] == 2 else tex.KIND_COLOR)
        want = tex.encode_rgba8(img, kind, tex.FORMAT_BC7, tex.QUALITY_NORMAL)
        assert dds.size == want.size and (dds == want).all()
        # and the Engine's reader accepts it (dds.hpp:409-574)
        if oracle.have_ref_dds():
            info = oracle.ref_dds_read(dds.tobytes())
            assert info["format"] == 98 and info["numMips"] == 7
