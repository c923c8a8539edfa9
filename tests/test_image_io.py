"""Image file I/O of the C++ host layer (vision_b200/host/vh_image.cpp): Radiance .hdr read / write, PNG write,
"gamma unless exr / hdr" (reference src/base/mgr/pipeline.cpp:337-354), .hdr = linear on load (image_pool.cpp:18-24)."""
import struct
import zlib

import numpy as np
import pytest

from vision_b200 import host_api, scene_py, capi


def test_rgbe_known_answers():
    # Ward's RGBE: shared exponent of the largest component; (1, .5, .25) = (128, 64, 32) * 2^(129 - 136)
    assert host_api.rgbe_encode([1.0, 0.5, 0.25]) == bytes([128, 64, 32, 129])
    assert host_api.rgbe_encode([0.0, 0.0, 0.0]) == bytes([0, 0, 0, 0])
    assert host_api.rgbe_encode([300.0, 1.0, 0.0]) == bytes([150, 0, 0, 137])          # 300 = 150 * 2, exponent 2^9 -> 128 + 9
    assert host_api.rgbe_encode([1e-3, 2e-3, 4e-3])[3] == 128 - 7                        # 4e-3 in [2^-8, 2^-7)


def test_hdr_round_trip_and_twin_reader(tmp_path):
    rng = np.random.default_rng(3)
    img = np.concatenate([rng.random((17, 33, 3), np.float32) * np.float32(50.0), np.ones((17, 33, 1), np.float32)], -1)
    img[0, 0, :3] = 0.0
    p = tmp_path / "a.hdr"
    host_api.save_image(p, img)
    back = host_api.load_image(p)
    assert back.shape == img.shape
    mx = img[..., :3].max(-1, keepdims=True)
    assert np.all(np.abs(back[..., :3] - img[..., :3]) <= mx / 128.0 + 1e-30)           # 8-bit mantissa of the shared exponent
    assert np.array_equal(back, scene_py.read_hdr(p))                                   # Python twin reads the same floats
    again = tmp_path / "b.hdr"
    host_api.save_image(again, back)                                                    # re-encoding decoded values is lossless
    assert np.array_equal(host_api.load_image(again), back)


def _rle_hdr(path, img8):
    """writes a new-style run-length Radiance file from (h, w, 4) RGBE bytes (what other tools produce)"""
    h, w, _ = img8.shape
    out = bytearray(b"#?RADIANCE\n# made by a test\nFORMAT=32-bit_rle_rgbe\nEXPOSURE=1.0\n\n-Y %d +X %d\n" % (h, w))
    for y in range(h):
        out += bytes([2, 2, w >> 8, w & 255])
        for c in range(4):
            row = img8[y, :, c]
            x = 0
            while x < w:
                run = 1
                while x + run < w and run < 127 and row[x + run] == row[x]:
                    run += 1
                if run >= 4:
                    out += bytes([128 + run, int(row[x])]); x += run
                else:
                    n = min(w - x, 16)
                    out += bytes([n]) + bytes(int(v) for v in row[x:x + n]); x += n
    path.write_bytes(bytes(out))


def test_hdr_run_length_scanlines(tmp_path):
    rng = np.random.default_rng(5)
    img8 = rng.integers(0, 256, (9, 40, 4), dtype=np.uint8)
    img8[:, 5:30, 0] = 77; img8[:, :, 3] = 130                                          # long runs in two channels
    p = tmp_path / "rle.hdr"
    _rle_hdr(p, img8)
    a = host_api.load_image(p)
    want = img8[..., :3].astype(np.float32) * np.float32(2.0 ** (130 - 136))
    assert np.array_equal(a[..., :3], want) and np.array_equal(a, scene_py.read_hdr(p))


def _read_png(path):
    data = path.read_bytes()
    assert data[:8] == b"\x89PNG\r\n\x1a\n"
    pos, chunks = 8, []
    while pos < len(data):
        n, tag = struct.unpack(">I4s", data[pos:pos + 8])
        body = data[pos + 8:pos + 8 + n]
        assert struct.unpack(">I", data[pos + 8 + n:pos + 12 + n])[0] == zlib.crc32(tag + body)       # chunk CRC
        chunks.append((tag, body)); pos += 12 + n
    w, h, depth, ctype = struct.unpack(">IIBB", chunks[0][1][:10])
    assert (depth, ctype) == (8, 2) and chunks[-1][0] == b"IEND"
    raw = zlib.decompress(b"".join(b for t, b in chunks if t == b"IDAT"))                  # checks the Adler-32 too
    rows = np.frombuffer(raw, np.uint8).reshape(h, 1 + 3 * w)
    assert (rows[:, 0] == 0).all()
    return rows[:, 1:].reshape(h, w, 3)


def test_png_is_valid_and_gamma_encoded(tmp_path):
    w, h = 300, 250                                                                      # > 65535 raw bytes: several stored blocks
    x = np.linspace(0, 1, w, dtype=np.float32)[None, :].repeat(h, 0)
    img = np.stack([x, x * 0.5, np.full_like(x, 0.2140411), np.ones_like(x)], -1)
    p = tmp_path / "o.png"
    host_api.save_image(p, img)
    got = _read_png(p)
    lin = img[..., :3].astype(np.float64)
    want = np.where(lin <= 0.0031308, 12.92 * lin, 1.055 * lin ** (1 / 2.4) - 0.055) * 255.0
    assert np.abs(got.astype(np.float64) - want).max() <= 0.51
    assert got[0, 0, 2] == 127 or got[0, 0, 2] == 128                                    # linear 0.214 = sRGB 0.5
    try:
        from PIL import Image
        assert np.array_equal(np.asarray(Image.open(p).convert("RGB")), got)
    except ImportError:
        pass


def test_linear_outputs_and_unknown_extension(tmp_path):
    img = np.full((4, 6, 4), 0.25, np.float32)
    host_api.save_image(tmp_path / "o.pfm", img)
    assert np.allclose(scene_py.read_pfm(tmp_path / "o.pfm")[..., :3], 0.25)             # float formats stay linear
    host_api.save_image(tmp_path / "o.hdr", img)
    assert np.allclose(host_api.load_image(tmp_path / "o.hdr")[..., :3], 0.25)
    with pytest.raises(capi.VbError):
        host_api.save_image(tmp_path / "o.tiff", img)


def test_hdr_environment_map_loads_natively(tmp_path):
    """a spherical light over a real .hdr file (no side-car): host layer == Python twin records"""
    import json
    src = scene_py.load_json_with_comments(scene_py.Path(__file__).resolve().parent.parent / "scenes" / "cbox" / "cbox_lights.json")
    env = scene_py.read_pfm(scene_py.Path(__file__).resolve().parent.parent / "scenes" / "cbox" / "env_small.pfm")
    host_api.save_image(tmp_path / "env.hdr", env)
    for l in src["light_sampler"]["param"]["lights"]:
        if l["type"] == "spherical":
            l["param"]["color"]["fn"] = "env.hdr"
    (tmp_path / "scene.json").write_text(json.dumps(src))
    hp = host_api.HostPipeline(tmp_path / "scene.json")
    hp.prepare_records()
    a, b = hp.records(), scene_py.build_scene(tmp_path / "scene.json")
    assert len(a.textures) == len(b.textures) == 1 and np.array_equal(a.textures[0], b.textures[0])
    assert a.alias_func == b.alias_func and [bytes(x) for x in a.lights] == [bytes(x) for x in b.lights]


# ---- decoders written in the host layer (vh_codecs.cpp): JPEG, PNG, EXR --------------------------------------------------------
ROOT = __import__("pathlib").Path(__file__).resolve().parent.parent
JPEGS = ["scenes/cbox_stock/checker.jpg", "scenes/cbox_stock/Painting3.jpg", "scenes/classroom/textures/blackboard.jpg",
         "scenes/classroom/textures/wood1.jpg", "scenes/classroom/textures/wood2.jpg"]


@pytest.mark.parametrize("rel", JPEGS)
def test_jpeg_decoder_equals_an_independent_decoder(rel):
    """the scenes' textures (baseline JPEG: 4:4:4 and 4:2:0, odd sizes): every decoded sample equals libjpeg's (through PIL) - the
    decoder follows the IJG integer IDCT, fancy upsampling and fixed-point colour conversion"""
    from PIL import Image
    a = host_api.load_image(ROOT / rel, False)
    b = np.asarray(Image.open(ROOT / rel).convert("RGB"), np.uint8)
    assert a.shape[:2] == b.shape[:2] and np.all(a[..., 3] == 1.0)
    assert np.array_equal(np.round(a[..., :3] * 255.0).astype(np.uint8), b)
    assert np.array_equal(a[..., :3], b.astype(np.float32) / np.float32(255.0))      # exactly value / 255
    # colour-space rule (image_pool.cpp:18-24): "srgb" sources are linearised on load
    lin = host_api.load_image(ROOT / rel, True)
    c = a[..., :3].astype(np.float64)
    want = np.where(c <= 0.04045, c / 12.92, ((c + 0.055) / 1.055) ** 2.4)
    assert np.abs(lin[..., :3] - want).max() < 1e-6


def test_jpeg_gray_restart_and_422(tmp_path):
    """grayscale, restart intervals, 4:2:2 and 4:4:0-free paths, written by an independent encoder"""
    from PIL import Image
    rng = np.random.default_rng(5)
    yy, xx = np.mgrid[0:83, 0:117]
    img = np.stack([(xx * 2) % 256, (yy * 3) % 256, ((xx + yy) * 2) % 256], -1).astype(np.uint8)
    img = (img // 2 + rng.integers(0, 100, img.shape)).astype(np.uint8)
    for name, kw, mode in [("gray.jpg", {}, "L"), ("s422.jpg", {"subsampling": "4:2:2"}, "RGB"), ("s420.jpg", {"subsampling": "4:2:0"}, "RGB"),
                           ("s444.jpg", {"subsampling": "4:4:4", "quality": 95}, "RGB")]:
        p = tmp_path / name
        Image.fromarray(img if mode == "RGB" else img[..., 0]).save(p, quality=kw.pop("quality", 80), **kw)
        a = host_api.load_image(p, False)
        b = np.asarray(Image.open(p).convert("RGB"), np.uint8)
        assert np.array_equal(np.round(a[..., :3] * 255.0).astype(np.uint8), b), name
    prog = tmp_path / "prog.jpg"
    Image.fromarray(img).save(prog, progressive=True)
    with pytest.raises(capi.VbError, match="progressive"):
        host_api.load_image(prog, False)


def test_png_decoder_all_colour_types(tmp_path):
    from PIL import Image
    rng = np.random.default_rng(7)
    rgb = rng.integers(0, 256, (37, 53, 3), dtype=np.uint8)
    cases = {"rgb": Image.fromarray(rgb), "rgba": Image.fromarray(np.dstack([rgb, rgb[..., :1]])), "gray": Image.fromarray(rgb[..., 0]),
             "la": Image.fromarray(np.dstack([rgb[..., 0], rgb[..., 1]]), "LA"), "pal": Image.fromarray(rgb).quantize(16),
             "bit1": Image.fromarray((rgb[..., 0] > 127))}
    for name, im in cases.items():
        p = tmp_path / f"{name}.png"
        im.save(p)
        a = host_api.load_image(p, False)
        b = np.asarray(Image.open(p).convert("RGBA"), np.uint8)
        assert np.array_equal(np.round(a * 255.0).astype(np.uint8), b), name
    p16 = tmp_path / "g16.png"
    g16 = rng.integers(0, 65536, (9, 11), dtype=np.uint16)
    Image.fromarray(g16).save(p16)
    a = host_api.load_image(p16, False)
    assert np.allclose(a[..., 0], g16 / 65535.0, atol=1e-7)
    # the host's own PNG writer round-trips through its reader (display gamma applied on write)
    img = np.concatenate([rng.random((20, 30, 3), np.float32), np.ones((20, 30, 1), np.float32)], -1)
    out = tmp_path / "w.png"
    host_api.save_image(out, img)
    back = host_api.load_image(out, True)          # srgb -> linear undoes the writer's gamma up to 8-bit quantisation
    assert np.abs(back[..., :3] - img[..., :3]).max() < 0.012


def test_exr_round_trip_and_refusals(tmp_path):
    rng = np.random.default_rng(11)
    img = np.concatenate([(rng.random((41, 29, 3), np.float32) * np.float32(100.0)) ** np.float32(2.0), np.ones((41, 29, 1), np.float32)], -1)
    p = tmp_path / "a.exr"
    host_api.save_image(p, img)                    # FLOAT channels, ZIP blocks of 16 lines, linear (no gamma: pipeline.cpp:337-354)
    back = host_api.load_image(p, True)            # "srgb" is ignored for .exr (image_pool.cpp:18-24)
    assert np.array_equal(back, img)
    raw = p.read_bytes()
    assert raw[:4] == bytes([0x76, 0x2F, 0x31, 0x01]) and b"compression\0compression\0\x01\0\0\0\x03" in raw
    # an uncompressed HALF file assembled by hand (what other tools write), 3 x 2 pixels, channels B G R
    h = np.array([[0.5, 1.0, 2.0], [4.0, 0.25, 1024.0]], np.float16)
    hdr = bytearray(struct.pack("<II", 20000630, 2))

    def attr(name, typ, val):
        hdr.extend(name + b"\0" + typ + b"\0" + struct.pack("<I", len(val)) + val)
    ch = b"".join(n + b"\0" + struct.pack("<IIII", 1, 0, 1, 1) for n in (b"B", b"G", b"R")) + b"\0"
    attr(b"channels", b"chlist", ch); attr(b"compression", b"compression", b"\0")
    attr(b"dataWindow", b"box2i", struct.pack("<iiii", 0, 0, 2, 1)); attr(b"displayWindow", b"box2i", struct.pack("<iiii", 0, 0, 2, 1))
    attr(b"lineOrder", b"lineOrder", b"\0"); attr(b"pixelAspectRatio", b"float", struct.pack("<f", 1.0))
    attr(b"screenWindowCenter", b"v2f", struct.pack("<ff", 0, 0)); attr(b"screenWindowWidth", b"float", struct.pack("<f", 1.0))
    hdr.append(0)
    line = 3 * 3 * 2
    off0 = len(hdr) + 16
    body = struct.pack("<QQ", off0, off0 + 8 + line)
    for y in range(2):
        body += struct.pack("<iI", y, line) + (h[y] * 1).tobytes() + (h[y] * 2).astype(np.float16).tobytes() + (h[y] * 0 + 3).astype(np.float16).tobytes()
    q = tmp_path / "half.exr"
    q.write_bytes(bytes(hdr) + body)
    a = host_api.load_image(q, False)
    assert a.shape == (2, 3, 4)
    assert np.array_equal(a[..., 2], h.astype(np.float32)) and np.all(a[..., 0] == 3.0)                 # B plane first, R last
    assert np.array_equal(a[..., 1], (h * 2).astype(np.float16).astype(np.float32))
    piz = bytes(hdr).replace(b"compression\0compression\0\x01\0\0\0\0", b"compression\0compression\0\x01\0\0\0\x04")
    (tmp_path / "piz.exr").write_bytes(piz + body)
    with pytest.raises(capi.VbError, match="PIZ"):
        host_api.load_image(tmp_path / "piz.exr", False)
