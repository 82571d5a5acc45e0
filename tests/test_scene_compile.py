"""The host scene compiler (scene_parser.h:241-595 + main.cpp:86-104 semantics) against dumps of the reference's own
build_scene / setup_camera (tests/golden/scene_*.json, and a live ref_driver run for the 8192-instance scene)."""
import json
import os

import numpy as np
import pytest

import scene_match
from conftest import ROOT, scene_path
from oracle_lib import RefDriver, have_ref_driver

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
SMALL = ["cornell_box_image_textures", "three_orbs", "cornell_box", "cornell_box_small_lights", "cornell_box_with_volume", "cornell_box_with_volume2",
         "light_test"]
MAT_KIND = {"lambertian": 0, "metal": 1, "dielectric": 2, "diffuse_light": 3, "isotropic": 4}


def _flat(name, w=32, h=18):
    import pathtrace_b200 as pt
    return pt.Scene.from_file(scene_path(name), w, h, device=-1)


def _check_against_dump(name, dump):
    sc = _flat(name)
    flat = sc.describe()
    mapping, worst = scene_match.match_leaves(dump, flat)
    assert worst == 0.0  # transforms and primitive parameters bit-equal (values survive the JSON round trip exactly)
    assert sorted(mapping.tolist()) == list(range(len(flat["instances"])))
    # the instance table is stored in the reference tree's leaf order (tie rule): the mapping is the identity
    # except among identical twins
    for r, leaf in enumerate(dump["leaves"]):
        inst = flat["instances"][mapping[r]]
        assert bool(inst["flags"] & 2) == bool(leaf["bvh_duplicated"])
        assert inst["light_index"] == leaf["light_index"]
        mat = flat["materials"][inst["material"]]
        assert mat["kind"] == MAT_KIND[leaf["prim"]["material"]["type"]]
        if leaf["prim"]["type"] == "rect":
            assert bool(inst["flags"] & 4) == (not leaf["prim"]["normal"])  # rect::normal = !flipped (primitive.h:128)
    assert len(flat["lights"]) == dump["n_lights"]
    cam = dump["camera"]
    for k_ref, k in (("origin", "origin"), ("lower_left_corner", "lower_left"), ("horizontal", "horizontal"),
                     ("vertical", "vertical"), ("u", "u"), ("v", "v"), ("w", "w")):
        assert np.array_equal(np.asarray(cam[k_ref], np.float32), flat["camera"][k]), k
    assert np.float32(cam["lens_radius"]) == np.float32(flat["camera"]["lens_radius"])
    # Perlin tables replayed from the default-seeded mt19937 (texture.h:76-110,180-183)
    assert np.array_equal(np.asarray(dump["perlin"]["ranvec"], np.float32), flat["perlin_ranvec"])
    for k, row in enumerate(("perm_x", "perm_y", "perm_z")):
        assert np.array_equal(np.asarray(dump["perlin"][row]), flat["perlin_perm"][k])
    # instance boxes (primitive.h:266-297) bound the BVH
    lo = np.min([l["bbox_min"] for l in dump["leaves"]], axis=0)
    hi = np.max([l["bbox_max"] for l in dump["leaves"]], axis=0)
    b = sc.bvh_info()["bounds"]
    assert np.array_equal(b[:3], lo.astype(np.float32)) and np.array_equal(b[3:], hi.astype(np.float32))
    return flat


@pytest.mark.parametrize("name", SMALL)
def test_compile_matches_reference_dump(name):
    with open(os.path.join(GOLDEN, f"scene_{name}.json")) as f:
        dump = json.load(f)
    _check_against_dump(name, dump)


def test_compile_matches_live_reference_book2():
    if not have_ref_driver():
        pytest.skip("oracle/_ref/ref_driver not built")
    flat = _check_against_dump("book2_final", RefDriver.scene(scene_path("book2_final"), 32, 18))
    assert len(flat["instances"]) == 8192 and sum(1 for i in flat["instances"] if i["flags"] & 1) == 2


def test_scene_files_equal_reference_files():
    """scenes/*.json restate the reference's scene files; where the reference tree is mounted they are compared."""
    ref_dir = "/root/reference/scenes"
    if not os.path.isdir(ref_dir):
        pytest.skip("reference tree not mounted")
    for name in SMALL:
        if not os.path.exists(os.path.join(ref_dir, name + ".json")):
            continue  # cornell_box_image_textures is ours (the reference's image scene has its texture disabled)
        with open(scene_path(name)) as f, open(os.path.join(ref_dir, name + ".json")) as g:
            assert json.load(f) == json.load(g), name


def test_parser_quirks():
    """Q13: primitives without a material get the MAUVE lambertian; `skip` instances are not instanced; a volume is a
    light iff its boundary's material is; unknown material strings fall to lambertian."""
    import pathtrace_b200 as pt
    sj = {
        "camera": {"look_from": [0, 0, -5], "look_at": [0, 0, 0]},
        "world": {"color": [0.1, 0.2, 0.3]},
        "textures": [], "materials": [{"id": "l", "type": "diffuse_light", "data": {"color": [2, 2, 2]}},
                                      {"id": "odd", "type": "no-such-material", "data": {"color": [0.5, 0.5, 0.5]}}],
        "primitives": [{"id": "s", "type": "sphere"}, {"id": "ls", "type": "sphere", "material": {"id": "l"}},
                       {"id": "o", "type": "sphere", "material": {"id": "odd"}}],
        "instances": [{"type": "ref", "primitive": {"id": "s"}},
                      {"type": "ref", "primitive": {"id": "ls"}, "transform": {"translate": [3, 0, 0]}},
                      {"type": "ref", "primitive": {"id": "o"}, "transform": {"translate": [-3, 0, 0], "scale": 0.5}},
                      {"type": "ref", "primitive": {"id": "s"}, "skip": True}],
    }
    flat = pt.Scene.from_json(sj, 16, 16, device=-1).describe()
    assert len(flat["instances"]) == 3 and len(flat["lights"]) == 1
    kinds = sorted(flat["materials"][i["material"]]["kind"] for i in flat["instances"])
    assert kinds == [0, 0, 3]
    bg = flat["textures"][flat["background"]]
    assert np.allclose(bg["color"], [0.1, 0.2, 0.3])
    assert abs(flat["camera"]["lens_radius"]) == 0.0


def _png_scene(path):
    return {"camera": {"look_from": [0, 0, -5], "look_at": [0, 0, 0]}, "world": {"texture": "img"},
            "textures": [{"id": "img", "type": "png", "data": {"path": path}}],
            "materials": [{"id": "m", "type": "lambertian", "data": {"texture": "img"}}],
            "primitives": [{"id": "s", "type": "sphere", "material": {"id": "m"}}],
            "instances": [{"type": "ref", "primitive": {"id": "s"}}]}


def test_png_textures_decode_like_lodepng(tmp_path):
    """`png` textures (scene_parser.h:39-55,318-323; image.h:52-70): the reader in host/png.cpp must give the RGBA8
    pixels lodepng::decode gives, stored as byte / 255.0 in float.  Checked against Pillow for every colour type and
    scanline filter the reader claims, and on the committed assets (whose rows cycle through all five filters)."""
    Image = pytest.importorskip("PIL.Image")
    import pathtrace_b200 as pt
    rng = np.random.default_rng(3)
    h, w = 13, 21
    rgba = rng.integers(0, 256, size=(h, w, 4), dtype=np.uint8)
    rgba[::3, ::2, 3] = 0
    cases = {"rgba8": Image.fromarray(rgba, "RGBA"), "rgb8": Image.fromarray(rgba[..., :3].copy(), "RGB"),
             "grey8": Image.fromarray(rgba[..., 0].copy(), "L"), "greya8": Image.fromarray(rgba[..., :2].copy(), "LA"),
             "pal8": Image.fromarray(rgba[..., :3].copy(), "RGB").quantize(colors=17),
             "grey16": Image.fromarray(rgba[..., 0].astype(np.uint16) * 257),
             "bilevel": Image.fromarray(rgba[..., 0] > 127)}
    for name, im in cases.items():
        path = str(tmp_path / (name + ".png"))
        im.save(path)
        expect = np.asarray(Image.open(path).convert("RGBA"), dtype=np.uint8)
        if name == "grey16":  # Pillow has no direct I;16 -> RGBA: the high byte is what lodepng keeps
            g = (np.asarray(Image.open(path), dtype=np.uint16) >> 8).astype(np.uint8)
            expect = np.dstack([g, g, g, np.full_like(g, 255)])
        got = pt.Scene.from_json(_png_scene(path), 8, 8, device=-1).describe()["images"][0]
        assert got.shape == (h, w, 4), name
        assert np.array_equal(got, (expect / 255.0).astype(np.float32)), name
    for asset in ("panel", "sky"):
        path = os.path.join(ROOT, "scenes", "assets", asset + ".png")
        got = pt.Scene.from_json(_png_scene(path), 8, 8, device=-1).describe()["images"][0]
        expect = np.asarray(Image.open(path).convert("RGBA"), dtype=np.uint8)
        assert np.array_equal(got, (expect / 255.0).astype(np.float32)), asset
    with pytest.raises(Exception, match="png"):
        pt.Scene.from_json(_png_scene(str(tmp_path / "missing.png")), 8, 8, device=-1)
    # relative paths: the working directory first (the reference), then the scene file's directory and its parent
    flat = pt.Scene.from_file(scene_path("cornell_box_image_textures"), 8, 8, device=-1).describe()
    assert [im.shape for im in flat["images"]] == [(8, 16, 4), (16, 32, 4)]
    kinds = [t["kind"] for t in flat["textures"]]
    assert kinds.count(3) == 2 and flat["textures"][flat["background"]]["kind"] == 3


def test_json_numbers_are_correctly_rounded():
    """json.cpp reads plain decimals without strtod (mantissa and power of ten both exact in a double -> one
    correctly rounded operation); everything else goes to strtod.  Either way the value must be the one Python's
    float() -- also correctly rounded -- gives for the same literal."""
    import random
    import pathtrace_b200 as pt
    rnd = random.Random(5)
    lits = ["0", "-0.0", "1", "0.1", "0.7", "25.0", "1e-7", "1E+5", "123456789012345", "1234567890123456789",
            "0.000000000000000000001", "3.141592653589793238462643", "1e22", "1e23", "1e-22", "1e-23", "4.35e-320",
            "17976931348623157e292", "0.30000000000000004", "9007199254740993", "5e-1", "100000000000000000000000"]
    for _ in range(300):
        m, e = rnd.randrange(10 ** rnd.randrange(1, 19)), rnd.randrange(-30, 30)
        lits += [f"{m}e{e}", f"{m // 1000}.{m % 1000:03d}", f"-0.{m:018d}", f"{m}.{m}E{e:+d}"]
    lits = lits[:len(lits) // 3 * 3]
    mats = ",".join('{"id": "m%d", "type": "lambertian", "data": {"color": [%s, %s, %s]}}' % (k, *lits[3 * k:3 * k + 3])
                    for k in range(len(lits) // 3))
    text = ('{"camera": {"look_from": [0, 0, -5], "look_at": [0, 0, 0], "fov": 40, "aperture": 0, "dist_to_focus": 5},'
            ' "world": {"color": [0, 0, 0]}, "assets": [], "textures": [], "materials": [' + mats + '], "primitives": [],'
            ' "instances": [{"type": "direct", "primitive": {"type": "sphere", "material": {"id": "m0"}, "radius": 1}}]}')
    sc = pt.Scene.from_json(text, 8, 8, device=-1)
    d = sc.describe()  # a lambertian's colour becomes a constant texture
    got = np.array([d["textures"][m["texture"]]["color"] for m in d["materials"]], dtype=np.float32).ravel()
    sc.close()
    with np.errstate(over="ignore"):
        want = np.array([float(s) for s in lits], dtype=np.float64).astype(np.float32)
    assert got.shape == want.shape
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), [l for l, g, w in zip(lits, got, want) if g != w][:5]
