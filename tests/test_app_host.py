"""CPU: the host side of the headless application layer (include/gsb_app.h) against golden vectors generated from the
reference's OWN sources (tests/golden/app_project.json <- tools/make_golden_app.cpp: src/Project.h + nlohmann json,
src/Camera.cpp + glm) and against the reference's file formats (.gobj UF:333-450, OBJ RH:107-180, PNG via zlib).
Everything here is exact: strings byte for byte, floats bit for bit."""
import importlib
import json
import os
import struct
import zlib

import numpy as np
import pytest

import helpers as H

gsb = H.gsb
app = importlib.import_module("gaussian-splatterer_b200.app")
GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "app_project.json")))
f32 = np.float32


def test_library_exports_every_app_symbol():
    L = app.L()
    missing = [s for s in app.APP_SYMBOLS_HOST if not hasattr(L, s)]
    assert not missing, missing


# ------------------------------------------------------------------------------------------------ settings.json
def test_project_defaults_and_dump_match_nlohmann_byte_for_byte():
    p = app.Project()
    assert p.dumps() == GOLD["default_dump"]
    assert app.Project(new=True).dumps() == GOLD["new_dump"]          # UiFrame::initProject
    q = app.Project(new=True)
    m = json.loads(GOLD["modified_dump"])
    q.perspective, q.pathModel, q.pathTextureDiffuse = m["perspective"], m["pathModel"], m["pathTextureDiffuse"]
    q.lrLocation, q.lrSh, q.paramScaleMax, q.paramCullSize, q.lrScale = 1.25e-7, 123456.0, 1e15, 1e16, 0.1
    q.previewTimer, q.iterations, q.previewTruth, q.previewFreeOrbit = 1234.5678, 98765, 1, 0
    q.d.sphere1.rotX, q.d.sphere1.rotY, q.paramSplitScale, q.paramCloneDistance = 359.99, -0.0, 1e-5, 0.0001
    assert q.dumps() == GOLD["modified_dump"]


@pytest.mark.parametrize("case", GOLD["parse"], ids=lambda c: c["in"][:40])
def test_project_from_json_follows_with_default_rules(case):
    p = app.Project()
    p.iterations, p.lrSh, p.pathModel = 555, 9.0, "keep?"
    if "error" in case:
        with pytest.raises(gsb.GsbError) as e:
            p.loads(case["in"])
        assert str(e.value) == case["error"]
        assert p.iterations == 555 and p.pathModel == "keep?"          # a failed load leaves the project untouched
    else:
        p.loads(case["in"])
        assert p.dumps() == case["out"]


def test_settings_file_round_trip_and_missing_file(tmp_path):
    p = app.Project(new=True)
    p.pathModel, p.lrOpacity, p.intervalDensify = "models/büste.obj", 0.01, 75
    path = tmp_path / "settings.json"
    p.save(path)
    assert path.read_text(encoding="utf-8") == p.dumps()               # `file << j`: compact dump, no newline
    q = app.Project()
    q.load(path)
    assert q.dumps() == p.dumps()
    with pytest.raises(gsb.GsbError, match="Failed to load settings file"):
        q.load(tmp_path / "nope.json")
    tp = p.train_params()
    assert [getattr(tp, n) for n, _ in tp._fields_] == [getattr(p, n) for n, _ in tp._fields_]


# ------------------------------------------------------------------------------------------------------ cameras
def test_cameras_bit_exact_with_reference_camera_cpp():
    gold = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "cameras.json")))
    seen = 0
    p = app.Project()
    cache = {}
    for g in gold:
        key = (g["n1"], g["n2"], g["d1"], g["d2"], g["f1"], g["f2"], g["rx1"], g["ry1"], g["rx2"], g["ry2"])
        if key not in cache:
            s1, s2 = p.d.sphere1, p.d.sphere2
            s1.count, s1.distance, s1.fovDeg, s1.rotX, s1.rotY = g["n1"], g["d1"], g["f1"], g["rx1"], g["ry1"]
            s2.count, s2.distance, s2.fovDeg, s2.rotX, s2.rotY = g["n2"], g["d2"], g["f2"], g["rx2"], g["ry2"]
            cache[key] = p.cameras()
        cam = cache[key][g["index"]]
        assert np.array_equal(np.array(cam.location[:], f32), np.array(g["location"], f32))
        assert cam.fovDegY == f32(g["fov"])
        c = app.camera_constants(cam, g["W"], g["H"], square_fov=True)
        assert np.array_equal(np.array(c.view[:], f32), np.array(g["view"], f32))
        assert np.array_equal(np.array(c.projview[:], f32), np.array(g["projview"], f32))
        seen += 1
    assert seen == len(gold) > 40


def test_preview_camera_and_inverse_matrix_match_glm():
    for g in GOLD["preview"]:
        p = app.Project()
        p.loads(g["project"])
        c = p.preview_camera()
        assert np.array_equal(np.array(c.location[:], f32), np.array(g["location"], f32)), g["project"][:80]
        assert np.array_equal(np.array(c.target[:], f32), np.array(g["target"], f32)) and c.fovDegY == f32(g["fov"])
    for g in GOLD["inverse"]:
        inv = app.camera_inverse_projview(app.camera_desc(g["location"], fov=g["fov"]), g["aspect"])
        assert np.array_equal(inv, np.array(g["inverse"], f32))
    p = app.Project()
    p.previewTruth, p.previewTruthIndex = 1, 99                        # std::vector::at throws in the reference
    with pytest.raises(gsb.GsbError, match="range"):
        p.preview_camera()
    c = app.camera_constants(app.camera_desc((0, 0, 10), fov=60.0), 1920, 1080)     # Trainer::render's fov convention
    assert c.tan_fovy == f32(np.tan(f32(np.radians(f32(60.0))) * f32(0.5)))
    assert abs(c.tan_fovx - np.tan(np.radians(1920 * 60.0 / 1080) / 2)) < 1e-5


# ------------------------------------------------------------------------------- ModelSplatsHost, .gobj, fields
def test_model_host_container_semantics():
    m = app.ModelSplatsHostC.create(2, 1, 4)
    m.push_back((1, 2, 3), np.arange(12), (0.1, 0.2, 0.3), 0.5, (1, 0, 0, 0))
    m.push_back((4, 5, 6), np.zeros(12), (0.4, 0.5, 0.6), 0.25, (0, 1, 0, 0))
    with pytest.raises(gsb.GsbError, match="Model ran out of capacity!"):
        m.push_back((0, 0, 0), np.zeros(12), (1, 1, 1), 1.0, (1, 0, 0, 0))
    m.copy(1, 0)
    a = m.arrays()
    assert a["locations"].tolist() == [1, 2, 3, 1, 2, 3] and a["shs"][12:].tolist() == list(range(12)) and a["opacities"].tolist() == [0.5, 0.5]
    with pytest.raises(gsb.GsbError, match="Can't copy splat in model"):
        m.copy(2, 0)
    big = app.ModelSplatsHostC.from_arrays(np.zeros(9), np.zeros(3 * 16 * 3), np.ones(9), np.ones(3), np.zeros(12))
    v = big.view()
    assert (v.capacity, v.count, v.shCoeffs, v.shDegree) == (1000000, 3, 16, 5)       # (16 - 1) / 3, MH:38
    with pytest.raises(gsb.GsbError, match="Inconsistent feature dimensions"):
        app.ModelSplatsHostC.from_arrays(np.zeros(9), np.zeros(36), np.ones(9), np.ones(2), np.zeros(12))


def test_gobj_text_format_round_trip(tmp_path):
    cloud = gsb.synth.splat_cloud(257, 0x60B1)
    m = app.ModelSplatsHostC.from_arrays(*[cloud[k] for k in ("locations", "shs", "scales", "opacities", "rotations")])
    path = tmp_path / "splats.gobj"
    m.save_gobj(path)
    lines = path.read_text().split("\n")
    assert len(lines) == 257 * 5 + 1 and lines[-1] == ""
    # line grammar and ostream default float formatting (%g, 6 significant digits), UF:342-353
    g = lambda x: "%g" % x
    assert lines[0] == "v " + " ".join(g(x) for x in cloud["locations"][:3])
    assert lines[1] == "sh " + " ".join(g(x) for x in cloud["shs"][:12])
    assert lines[2] == "s " + " ".join(g(x) for x in cloud["scales"][:3])
    assert lines[3] == "a " + g(cloud["opacities"][0])
    assert lines[4] == "r " + " ".join(g(x) for x in cloud["rotations"][:4])
    back = app.ModelSplatsHostC.load_gobj(path)
    v, a = back.view(), back.arrays()
    assert (v.count, v.shCoeffs, v.shDegree, v.capacity) == (257, 4, 1, 1000000)
    for k in ("locations", "shs", "scales", "opacities", "rotations"):
        want = np.array([f32(float(g(x))) for x in np.asarray(cloud[k]).reshape(-1)], f32)     # 6 digits survive, no more
        assert np.array_equal(a[k], want), k
    # a second save of the loaded model is byte-identical (the format is a fixed point after one pass)
    back.save_gobj(tmp_path / "again.gobj")
    assert (tmp_path / "again.gobj").read_bytes() == path.read_bytes()
    (tmp_path / "bad.gobj").write_text("v 0 0 0\nsh 1 2 3\ns 1 1 1\na 1\nr 1 0 0 0\nv 1 1 1\nsh 1 2 3 4 5 6\n")
    with pytest.raises(gsb.GsbError, match="Inconsistent SH degree!"):
        app.ModelSplatsHostC.load_gobj(tmp_path / "bad.gobj")
    with pytest.raises(gsb.GsbError, match="Failed to load splats file"):
        app.ModelSplatsHostC.load_gobj(tmp_path / "missing.gobj")


def test_field_initialisers_match_reference():
    grid = app.ModelSplatsHostC.field_grid()
    v, a = grid.view(), grid.arrays()
    assert (v.count, v.capacity, v.shDegree, v.shCoeffs) == (17 ** 3, 1000000, 1, 4)
    loc = a["locations"].reshape(-1, 3)
    ax = np.arange(-4.0, 4.25, 0.5, dtype=f32)
    want = np.stack(np.meshgrid(ax, ax, ax, indexing="ij"), -1).reshape(-1, 3)          # x outermost, z innermost (UF:147-153)
    assert np.array_equal(loc, want)
    assert np.all(a["scales"] == f32(0.5) * f32(0.1)) and np.all(a["opacities"] == 1.0) and not a["shs"].any()
    ident = np.array(GOLD["field_identity_rotation"], f32)
    assert np.array_equal(a["rotations"].reshape(-1, 4), np.tile(ident, (v.count, 1)))
    mono = app.ModelSplatsHostC.field_mono()
    am = mono.arrays()
    assert mono.view().count == 1 and am["scales"].tolist() == [f32(0.3)] * 3 and am["locations"].tolist() == [0, 0, 0]
    assert np.array_equal(am["rotations"], ident)


def test_field_from_obj_model_matches_glm(tmp_path):
    tris = GOLD["field_model"]
    with open(tmp_path / "m.obj", "w") as f:
        f.write("# comment\nvt 0.5 0.5\n")
        for t in tris:
            for k in range(3):
                f.write("v %.9g %.9g %.9g\n" % tuple(t["tri"][3 * k:3 * k + 3]))
        for i in range(len(tris)):
            f.write("f %d/1/1 %d/1 %d\n" % (3 * i + 1, 3 * i + 2, 3 * i + 3))
        f.write("f 1 2 3 4\n")                                                          # a quad -> two more triangles (UF:223-225)
    m = app.ModelSplatsHostC.field_model(tmp_path / "m.obj")
    a = m.arrays()
    assert m.view().count == len(tris) + 2
    for i, t in enumerate(tris):
        assert np.array_equal(a["locations"][3 * i:3 * i + 3], np.array(t["location"], f32)), i
        assert np.array_equal(a["scales"][3 * i:3 * i + 3], np.array(t["scale"], f32)), i
        assert np.array_equal(a["rotations"][4 * i:4 * i + 4], np.array(t["rotation"], f32)), i
    (tmp_path / "five.obj").write_text("v 0 0 0\nv 1 0 0\nv 0 1 0\nv 1 1 0\nv 2 2 0\nf 1 2 3 4 5\n")
    with pytest.raises(gsb.GsbError, match=r"Unexpected vertex count in face list!5"):
        app.ModelSplatsHostC.field_model(tmp_path / "five.obj")
    with pytest.raises(gsb.GsbError, match="Failed to load the model file for splat field generation!"):
        app.ModelSplatsHostC.field_model(tmp_path / "none.obj")


# ---------------------------------------------------------------------------------------------------------- PNG
def _png_decode(path):
    """Independent PNG reader (zlib + the five row filters), RGBA8 / RGB8 / grey8 only: the checker for gsb_png_write."""
    b = open(path, "rb").read()
    assert b[:8] == b"\x89PNG\r\n\x1a\n"
    at, idat, hdr = 8, b"", None
    while at < len(b):
        n, t = struct.unpack(">I4s", b[at:at + 8])
        data = b[at + 8:at + 8 + n]
        assert zlib.crc32(t + data) == struct.unpack(">I", b[at + 8 + n:at + 12 + n])[0]
        if t == b"IHDR":
            hdr = struct.unpack(">IIBBBBB", data)
        elif t == b"IDAT":
            idat += data
        at += 12 + n
    w, h, depth, ctype = hdr[:4]
    ch = {0: 1, 2: 3, 6: 4}[ctype]
    raw = zlib.decompress(idat)
    stride = w * ch
    out = np.zeros((h, stride), np.uint8)
    prev = np.zeros(stride, np.int32)
    for y in range(h):
        ft = raw[y * (stride + 1)]
        cur = np.frombuffer(raw[y * (stride + 1) + 1:(y + 1) * (stride + 1)], np.uint8).astype(np.int32)
        rec = np.zeros(stride, np.int32)
        for i in range(stride):
            a = rec[i - ch] if i >= ch else 0
            bb, c = prev[i], (prev[i - ch] if i >= ch else 0)
            if ft == 0: pr = 0
            elif ft == 1: pr = a
            elif ft == 2: pr = bb
            elif ft == 3: pr = (a + bb) >> 1
            else:
                pp = a + bb - c
                pa, pb, pc = abs(pp - a), abs(pp - bb), abs(pp - c)
                pr = a if (pa <= pb and pa <= pc) else (bb if pb <= pc else c)
            rec[i] = (cur[i] + pr) & 255
        out[y] = rec
        prev = rec
    return out.reshape(h, w, ch), depth


def test_png_write_is_flipped_rgba8_and_decodes_exactly(tmp_path):
    rng = np.random.default_rng(5)
    w, h = 37, 23
    img = rng.integers(0, 256, (h, w, 4), dtype=np.uint8)
    img[5:15, 3:30] = img[5, 3]                       # flat area: exercises Sub/Up filters
    fb = img.reshape(-1).view(np.uint32)              # make_rgba packing: R | G << 8 | B << 16 | A << 24
    app.png_write(tmp_path / "a.png", fb, w, h, flip_vertically=True)
    dec, depth = _png_decode(tmp_path / "a.png")
    assert depth == 8 and np.array_equal(dec, img[::-1])              # stbi_flip_vertically_on_write(true), TV:254
    app.png_write(tmp_path / "b.png", fb, w, h, flip_vertically=False)
    assert np.array_equal(_png_decode(tmp_path / "b.png")[0], img)
    assert np.array_equal(app.png_read(tmp_path / "b.png"), img)      # own reader agrees too


def _png_encode(path, arr, ctype, depth=8, palette=None, trns=None):
    h, w = arr.shape[:2]
    rows = b"".join(b"\x00" + arr[y].tobytes() for y in range(h))
    def chunk(t, d):
        return struct.pack(">I", len(d)) + t + d + struct.pack(">I", zlib.crc32(t + d))
    out = b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w if depth >= 8 else arr.shape[1] * 8 // depth, h, depth, ctype, 0, 0, 0))
    if palette is not None:
        out += chunk(b"PLTE", bytes(palette))
    if trns is not None:
        out += chunk(b"tRNS", bytes(trns))
    out += chunk(b"IDAT", zlib.compress(rows)) + chunk(b"IEND", b"")
    open(path, "wb").write(out)


def test_png_read_expands_like_stbi_load_rgb_alpha(tmp_path):
    rng = np.random.default_rng(6)
    rgb = rng.integers(0, 256, (5, 7, 3), dtype=np.uint8)
    _png_encode(tmp_path / "rgb.png", rgb, 2)
    got = app.png_read(tmp_path / "rgb.png")
    assert np.array_equal(got[..., :3], rgb) and np.all(got[..., 3] == 255)
    grey = rng.integers(0, 256, (4, 6), dtype=np.uint8)
    _png_encode(tmp_path / "g.png", grey, 0)
    got = app.png_read(tmp_path / "g.png")
    assert all(np.array_equal(got[..., k], grey) for k in range(3)) and np.all(got[..., 3] == 255)
    ga = rng.integers(0, 256, (3, 5, 2), dtype=np.uint8)
    _png_encode(tmp_path / "ga.png", ga, 4)
    got = app.png_read(tmp_path / "ga.png")
    assert np.array_equal(got[..., 0], ga[..., 0]) and np.array_equal(got[..., 3], ga[..., 1])
    idx = rng.integers(0, 4, (4, 4), dtype=np.uint8)
    pal = [255, 0, 0, 0, 255, 0, 0, 0, 255, 9, 8, 7]
    _png_encode(tmp_path / "p.png", idx, 3, palette=pal, trns=[0, 128])
    got = app.png_read(tmp_path / "p.png")
    lut = np.array(pal, np.uint8).reshape(4, 3)
    assert np.array_equal(got[..., :3], lut[idx]) and np.array_equal(got[..., 3], np.array([0, 128, 255, 255], np.uint8)[idx])
    rgb16 = rng.integers(0, 65536, (2, 3, 3)).astype(">u2")
    _png_encode(tmp_path / "16.png", rgb16, 2, depth=16)
    assert np.array_equal(app.png_read(tmp_path / "16.png")[..., :3], (rgb16 >> 8).astype(np.uint8))     # stb keeps the high byte
    bits = np.array([[0b10110000]], np.uint8)                           # 1-bit grey, 4 pixels wide
    _png_encode(tmp_path / "1.png", bits, 0, depth=1)
    # width from the helper is 8 pixels for one byte
    assert app.png_read(tmp_path / "1.png")[0, :4, 0].tolist() == [255, 0, 255, 255]
    (tmp_path / "x.jpg").write_bytes(b"\xff\xd8\xff\xe0 not a png")
    with pytest.raises(gsb.GsbError, match="not a PNG"):
        app.png_read(tmp_path / "x.jpg")
