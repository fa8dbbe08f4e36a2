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


def _mosaic(base, times):
    """a (times x times) mosaic of flipped / rolled copies of `base`: a large texture without the minutes the noise
    stack would take at 4096^2 (values are only moved, so a normal map stays a normal map)"""
    rows = []
    for j in range(times):
        row = []
        for i in range(times):
            t = np.roll(base, (37 * (i + times * j) + 1, 53 * (i + times * j) + 3), axis=(0, 1))
            if i & 1:
                t = t[:, ::-1]
            if j & 1:
                t = t[::-1]
            row.append(t)
        rows.append(np.concatenate(row, axis=1))
    return np.ascontiguousarray(np.concatenate(rows, axis=0))


@pytest.mark.gpu
def test_sponza_scale_gltf_one_packager_run_against_the_oracle(tmp_path, tex, oracle, synth):
    """BASELINE.json configs[2] under test: one `rosy_packager_tex scene.gltf` run over a
    Sponza-like material set - 23 materials x (albedo, normal, metallic-roughness) + a shared detail map = 70 images,
    the 1K-4K mix scaled by 1/8 so that the CPU oracle's share stays well under a minute, square and 2:1, PNG and
    JPEG - plus one full 4096^2 image of each kind.  Every .dds the run writes is checked against the ORACLE
    (oracle.generate_texture on the decoded pixels) by the rules of the BASELINE-size parity tests: header equal,
    every level's blocks bit-exact on identical pixels, mip levels within 1 LSB, PSNR within 0.1 dB.  The loop being
    replaced: Gltf.cpp:209-285 (one generate_* call per image, BC7 Quality_Normal for all three kinds)."""
    from PIL import Image

    from test_gpu_parity import _full_size_parity

    sizes = [(128, 128), (256, 256), (512, 512), (256, 128), (512, 256)]  # 1K, 2K, 4K, 2Kx1K, 4Kx2K at 1/8
    images, textures, materials = [], [], []

    def add(name, arr, jpeg=False, level=6):
        im = Image.fromarray(arr)
        if jpeg:
            im.convert("RGB").save(tmp_path / name, quality=92)
        else:
            im.save(tmp_path / name, compress_level=level)
        images.append({"uri": name})
        textures.append({"source": len(images) - 1})
        return len(textures) - 1

    for m in range(23):
        w, h = sizes[m % len(sizes)]
        a = add(f"mat{m:02d}_albedo.{'jpg' if m % 3 == 0 else 'png'}", synth.albedo(w, h, m if m % 4 else 7), jpeg=m % 3 == 0)
        n = add(f"mat{m:02d}_normal.png", synth.normal(w, h, m))
        r = add(f"mat{m:02d}_orm.png", synth.make("metal_rough", w, h, m))
        materials.append({"pbrMetallicRoughness": {"baseColorTexture": {"index": a}, "metallicRoughnessTexture": {"index": r}},
                          "normalTexture": {"index": n}})
    detail = add("shared_detail.png", synth.albedo(256, 256, 3))
    materials.append({"pbrMetallicRoughness": {"baseColorTexture": {"index": detail}}})
    big = {"albedo": _mosaic(synth.albedo(1024, 1024, 5), 4), "normal": _mosaic(synth.normal(1024, 1024, 2), 4),
           "metal_rough": _mosaic(synth.make("metal_rough", 1024, 1024, 1), 4)}
    ba = add("big_albedo.png", big["albedo"], level=0)
    bn = add("big_normal.png", big["normal"], level=0)
    br = add("big_orm.png", big["metal_rough"], level=0)
    materials.append({"pbrMetallicRoughness": {"baseColorTexture": {"index": ba}, "metallicRoughnessTexture": {"index": br}},
                      "normalTexture": {"index": bn}})
    assert len(images) == 73
    p = tmp_path / "sponza_like.gltf"
    p.write_text(json.dumps({"asset": {"version": "2.0"}, "images": images, "textures": textures, "materials": materials}))

    out = subprocess.run([CLI, str(p)], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    manifest = json.loads((tmp_path / "sponza_like.textures.json").read_text())
    assert len(manifest["jobs"]) == 73
    differing = total = 0
    for j in manifest["jobs"]:
        kind = {0: tex.KIND_COLOR, 1: tex.KIND_NORMAL_MAP, 2: tex.KIND_METALLIC_ROUGHNESS}[j["image_type"]]
        name = os.path.basename(j["source"])
        assert ("_normal." in name) == (kind == tex.KIND_NORMAL_MAP) and ("_orm." in name) == (kind == tex.KIND_METALLIC_ROUGHNESS)
        img = tex.load_image(j["source"])  # the decoders are pinned against stb_image in test_abi_host.py
        if name.startswith("big_") and not name.endswith(".jpg"):
            assert (img == big[{"big_albedo.png": "albedo", "big_normal.png": "normal", "big_orm.png": "metal_rough"}[name]]).all()
        dds = np.fromfile(j["dds"], dtype=np.uint8)
        d, t = _full_size_parity(tex, oracle, img, kind, tex.FORMAT_BC7, tex.QUALITY_NORMAL, got=dds,
                                 per_level_min_px=1024 if img.shape[0] >= 1024 else 16384)
        differing += d
        total += t
    print(f"configs[2]-like run: 73 .dds files, {differing} of {total} blocks differ end to end (1-LSB mip differences)")


# ---- .rsy image table (Asset.cpp:39-113, :296-340) and the FBX image routing (Fbx.cpp:421-474) -------------
def _parse_rsy(data):
    """hand-written reader of what asset::write stores for an asset that holds only images"""
    import struct

    magic, version, endian = struct.unpack_from("<III", data, 0)
    coord = struct.unpack_from("<16f", data, 12)
    (root_scene,) = struct.unpack_from("<I", data, 76)
    sizes = struct.unpack_from("<6Q", data, 80)
    off, images = 128, []
    for _ in range(sizes[4]):
        ntypes, nchars = struct.unpack_from("<QQ", data, off)
        (itype,) = struct.unpack_from("<I", data, off + 16)
        name = data[off + 20 : off + 20 + nchars].decode()
        assert ntypes == 1
        images.append((itype, name))
        off += 20 + nchars
    assert off == len(data), "trailing bytes"
    return {"magic": magic, "version": version, "endianness": endian, "coord": coord, "root_scene": root_scene,
            "sizes": sizes, "images": images}


def test_rsy_image_table_bytes_for_gltf(tmp_path):
    p, names = _write_gltf(tmp_path, with_images=False)
    # <file>.rsy belongs to the reference's importer (Main.cpp:13-14, the complete asset): this tool must never touch it
    (tmp_path / "scene.rsy").write_bytes(b"the importer's complete asset")
    before = sorted(os.listdir(tmp_path))
    out = subprocess.run([CLI, "--dry-run", str(p)], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    assert sorted(os.listdir(tmp_path)) == before, "--dry-run must not write anything"
    out = subprocess.run([CLI, "--image-table", str(p)], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    assert (tmp_path / "scene.rsy").read_bytes() == b"the importer's complete asset"
    data = (tmp_path / "scene.images.rsy").read_bytes()
    # hand-computed vector for the first bytes: 'RSYF' little endian, version 1, endianness 1, identity, root scene 0,
    # sizes {0,0,0,0,4,0}
    import struct

    head = struct.pack("<III", 0x52535946, 1, 1) + struct.pack("<16f", *[1.0 if i % 5 == 0 else 0.0 for i in range(16)]) + \
        struct.pack("<I", 0) + struct.pack("<6Q", 0, 0, 0, 0, 4, 0)
    assert data[:128] == head and data[:4] == b"FYSR"
    r = _parse_rsy(data)
    # image i: name = <asset dir>/<uri up to the first '.'>.dds (Gltf.cpp:102-108); type from the list that conditions it:
    # 0 colour (image 0), 2 metallic-roughness (image 1: the metallic list runs after the colour list), 1 normal (image 2),
    # image 3 (emissive only) is never conditioned and keeps the default 0
    want_names = [str(tmp_path / (n.replace(" ", "%20").split(".")[0] + ".dds")) for n in names]
    assert [n for _, n in r["images"]] == want_names
    assert [t for t, _ in r["images"]] == [0, 2, 1, 0]
    # first record, byte for byte: size_t 1, size_t len, uint32 type, chars
    n0 = want_names[0].encode()
    assert data[128 : 148 + len(n0)] == struct.pack("<QQI", 1, len(n0), 0) + n0


def test_rsy_image_table_pinned_to_the_reference_writer_and_reader(tmp_path, oracle):
    """the image table this repo writes is the file the reference's own asset::write
    (src/Asset/Asset.cpp:39-113, :296-340, compiled as is into oracle/_ref/libasset_ref.so) writes for the same
    images, byte for byte, and the reference's own asset::read (:409-848) reads ours back"""
    if not oracle.have_ref_asset():
        pytest.skip("oracle/_ref/libasset_ref.so not built (make -C oracle ref)")
    p, names = _write_gltf(tmp_path, with_images=False)
    out = subprocess.run([CLI, "--image-table", str(p)], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    ours = (tmp_path / "scene.images.rsy").read_bytes()
    images = _parse_rsy(ours)["images"]
    assert len(images) == 4
    oracle.ref_asset_write_images(tmp_path / "ref.rsy", images)
    assert (tmp_path / "ref.rsy").read_bytes() == ours
    got = oracle.ref_asset_read_images(tmp_path / "scene.images.rsy")
    assert got is not None, "asset::read rejected the file"
    assert got[0] == images and got[1] == (0, 0, 0, 0, 0)
    # and the FBX route, whose names are full .dds paths and whose types include the mixmap (3)
    fbx, _ = _write_fbx_dir(tmp_path / "fbx")
    out = subprocess.run([CLI, "--image-table", str(fbx)], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    ours = (tmp_path / "fbx" / "rock.images.rsy").read_bytes()
    images = _parse_rsy(ours)["images"]
    assert sorted(t for t, _ in images) == [0, 0, 1, 3]
    oracle.ref_asset_write_images(tmp_path / "ref_fbx.rsy", images)
    assert (tmp_path / "ref_fbx.rsy").read_bytes() == ours
    assert oracle.ref_asset_read_images(tmp_path / "fbx" / "rock.images.rsy")[0] == images
    # an empty image list is a valid asset too
    oracle.ref_asset_write_images(tmp_path / "empty.rsy", [])
    assert _parse_rsy((tmp_path / "empty.rsy").read_bytes())["sizes"] == (0, 0, 0, 0, 0, 0)


def _write_fbx_dir(tmp_path, size=64):
    tmp_path.mkdir(parents=True, exist_ok=True)
    """a .fbx (never opened: the image section only scans its directory) and the TGA set of Fbx.cpp:426-430,
    plus decoys: wrong extension, wrong suffix, suffix match inside the name only"""
    from rosy_b200 import synth

    def tga(path, img):
        h, w = img.shape[:2]
        hdr = bytes([0, 0, 2, 0, 0, 0, 0, 0, 0, 0, 0, 0, w & 255, w >> 8, h & 255, h >> 8, 32, 0x28])
        path.write_bytes(hdr + img[..., [2, 1, 0, 3]].tobytes())

    imgs = {"rock_normal.tga": synth.normal(size, size, 1), "rock_mixmap.tga": synth.make("metal_rough", size, size, 2),
            "rock_albedo.tga": synth.albedo(size, size, 7), "albedo.tga": synth.albedo(size, size, 3)}
    for n, im in imgs.items():
        tga(tmp_path / n, im)
    tga(tmp_path / "rock_height.tga", imgs["albedo.tga"])          # unsupported suffix
    tga(tmp_path / "normal.tga.bak", imgs["albedo.tga"])           # not a .tga
    (tmp_path / "rock_albedo.png").write_bytes(b"not read")        # FBX import only supports .tga
    tga(tmp_path / "albedo_rock.tga", imgs["albedo.tga"])          # 'albedo' not at the end
    (tmp_path / "rock.fbx").write_bytes(b"Kaydara FBX Binary  \x00")
    return tmp_path / "rock.fbx", imgs


def test_fbx_image_routing_plan_and_rsy(tmp_path):
    fbx, imgs = _write_fbx_dir(tmp_path)
    out = subprocess.run([CLI, "--dry-run", str(fbx)], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    lines = [l for l in out.stdout.splitlines() if " -> " in l]
    # visited in file-name order; type 0 albedo, 3 mixmap, 1 normal (Fbx.cpp:450,459,468)
    want = [(0, "albedo.tga"), (0, "rock_albedo.tga"), (3, "rock_mixmap.tga"), (1, "rock_normal.tga")]
    assert [(int(l.split()[0]), os.path.basename(l.split(" -> ")[0].split(" ", 1)[1])) for l in lines] == want
    assert "image found:" in out.stdout + out.stderr
    assert not (tmp_path / "rock.rsy").exists() and not (tmp_path / "rock.images.rsy").exists()  # a dry run writes nothing
    out = subprocess.run([CLI, "--image-table", str(fbx)], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    r = _parse_rsy((tmp_path / "rock.images.rsy").read_bytes())
    assert r["sizes"] == (0, 0, 0, 0, 4, 0)
    assert r["images"] == [(t, str(tmp_path / n.replace(".tga", ".dds"))) for t, n in want]
    assert not (tmp_path / "rock.rsy").exists()
    # a complete asset (any count besides the image count non-zero) under the target name is never replaced
    import struct

    full = bytearray((tmp_path / "rock.images.rsy").read_bytes())
    struct.pack_into("<Q", full, 80 + 8 * 5, 3)  # pretend it holds 3 meshes
    (tmp_path / "rock.images.rsy").write_bytes(bytes(full))
    out = subprocess.run([CLI, "--image-table", str(fbx)], capture_output=True, text=True)
    assert out.returncode != 0 and "refusing to overwrite" in out.stdout + out.stderr
    assert (tmp_path / "rock.images.rsy").read_bytes() == bytes(full)


@pytest.mark.gpu
@pytest.mark.parametrize("serial", [False, True])
def test_fbx_packager_run(tmp_path, serial, tex):
    """`rosy_packager_tex rock.fbx`: the three TGA kinds conditioned like Fbx.cpp:443-473, byte-checked against the
    C ABI's single-texture entry points; a corrupt TGA aborts the asset (non-zero exit, reference error line)"""
    fbx, imgs = _write_fbx_dir(tmp_path)
    out = subprocess.run([CLI] + (["--serial"] if serial else []) + [str(fbx)], capture_output=True, text=True)
    assert out.returncode == 0, out.stdout + out.stderr
    for name, img in imgs.items():
        kind = tex.KIND_NORMAL_MAP if name.endswith("normal.tga") else (tex.KIND_MIXMAP if name.endswith("mixmap.tga") else tex.KIND_COLOR)
        got = np.fromfile(tmp_path / name.replace(".tga", ".dds"), dtype=np.uint8)
        assert (got == tex.encode_rgba8(img, kind)).all(), name
    assert not (tmp_path / "rock_height.dds").exists() and not (tmp_path / "albedo_rock.dds").exists()
    assert len(_parse_rsy((tmp_path / "rock.images.rsy").read_bytes())["images"]) == 4
    assert not (tmp_path / "rock.rsy").exists()
    (tmp_path / "bad_normal.tga").write_bytes(b"\x00" * 10)
    out = subprocess.run([CLI, "--serial", str(fbx)], capture_output=True, text=True)
    assert out.returncode != 0
    assert "error creating normal fbx image" in out.stdout + out.stderr
