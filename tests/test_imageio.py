"""CPU tier: the host API's own image I/O (pathtracing_b200/host/ImageIO.hpp: PNG / PPM / BMP / TGA / HDR / PFM decoders, inflate,
baseline JPEG encoder) and the Film writers of host/PathTracing.hpp, against Pillow / OpenCV and against the REFERENCE's writers
(Film::WritePPM / WritePNG / WriteJPG, Film.hpp:154-217, called through oracle/ref_driver.cpp ref_film_write):
PPM byte for byte, PNG pixel for pixel, JPEG to encoder accuracy.  Also the texture-file constructors of the host API
(Texture.hpp:139,170) and the bake of an arbitrary FunctionInfiniteLight sky (Light.hpp:74)."""
import ctypes as C
import struct
import subprocess
from pathlib import Path

import numpy as np
import pytest

from conftest import ROOT, needs_ref

HOST = ROOT / "pathtracing_b200" / "host"


@pytest.fixture(scope="module")
def tool(tmp_path_factory):
    d = tmp_path_factory.mktemp("imageio")
    exe = d / "imageio_tool"
    subprocess.run(["g++", "-std=c++20", "-O1", "-I", str(HOST), "-I", str(HOST / "glm_min"), "-I", str(ROOT / "include"),
                    str(ROOT / "tests" / "host" / "imageio_tool.cpp"), "-o", str(exe), "-L", str(ROOT / "pathtracing_b200"), "-lptb200",
                    f"-Wl,-rpath,{ROOT / 'pathtracing_b200'}"], check=True)
    return exe, d


def _run(exe, *args):
    r = subprocess.run([str(exe), *map(str, args)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return r.stdout


def _raw(path, dtype):
    b = Path(path).read_bytes()
    head, _, body = b.partition(b"\n")
    w, h, c = map(int, head.split())
    return np.frombuffer(body, dtype=dtype).reshape(h, w, c)


def _picture(h, w, c, seed):
    rng = np.random.default_rng(seed)
    y, x = np.mgrid[0:h, 0:w]
    base = np.stack([(x * 7 + y * 3) % 256, (x * x // 5 + y) % 256, (x + y * y // 3) % 256, (x * y) % 256], -1)[..., :c]
    noise = rng.integers(0, 40, (h, w, c))
    return ((base + noise) % 256).astype(np.uint8)


def test_png_decoder_against_pillow(tool):
    """every colour type and bit depth Pillow writes, compressed by zlib (dynamic and fixed Huffman blocks, all five filters)"""
    from PIL import Image
    exe, d = tool
    cases = {"rgb": ("RGB", 3), "rgba": ("RGBA", 4), "grey": ("L", 1), "grey_alpha": ("LA", 2)}
    for name, (mode, c) in cases.items():
        for k, (h, w) in enumerate(((37, 53), (1, 1), (64, 200))):
            a = _picture(h, w, c, k)
            f = d / f"{name}_{k}.png"
            Image.fromarray(a[..., 0] if c == 1 else a, mode).save(f, compress_level=[9, 1, 6][k])
            got = _raw((_run(exe, "decode8", f, d / "o.raw"), d / "o.raw")[1], np.uint8)
            assert got.shape == (h, w, c) and np.array_equal(got, a), name
    # palette (with and without transparency), 16 bit, 1-bit grey
    a = _picture(40, 60, 3, 5)
    Image.fromarray(a, "RGB").quantize(64).save(d / "pal.png")
    want = np.array(Image.open(d / "pal.png").convert("RGB"))
    _run(exe, "decode8", d / "pal.png", d / "o.raw")
    assert np.array_equal(_raw(d / "o.raw", np.uint8), want)
    g16 = (np.arange(50 * 30, dtype=np.uint32).reshape(30, 50) * 37 % 65536).astype(np.uint16)
    Image.fromarray(g16).save(d / "g16.png")
    _run(exe, "decode8", d / "g16.png", d / "o.raw")
    assert np.array_equal(_raw(d / "o.raw", np.uint8)[..., 0], (g16 >> 8).astype(np.uint8))      # the high byte, as stb_image converts 16 -> 8
    bw = (_picture(33, 41, 1, 2)[..., 0] > 127)
    Image.fromarray(bw).save(d / "bw.png")
    _run(exe, "decode8", d / "bw.png", d / "o.raw")
    assert np.array_equal(_raw(d / "o.raw", np.uint8)[..., 0], bw.astype(np.uint8) * 255)
    # stored (uncompressed) deflate blocks: what Film::WritePNG of the host API emits
    raw = b"".join(b"\x00" + a[y].tobytes() for y in range(a.shape[0]))
    import zlib
    z = zlib.compressobj(0).compress(raw) + zlib.compressobj(0).flush()
    z = zlib.compress(raw, 0)

    def chunk(t, data):
        return struct.pack(">I", len(data)) + t + data + struct.pack(">I", zlib.crc32(t + data))
    (d / "stored.png").write_bytes(b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", a.shape[1], a.shape[0], 8, 2, 0, 0, 0)) + chunk(b"IDAT", z) + chunk(b"IEND", b""))
    _run(exe, "decode8", d / "stored.png", d / "o.raw")
    assert np.array_equal(_raw(d / "o.raw", np.uint8), a)


def test_other_ldr_decoders_against_pillow(tool):
    from PIL import Image
    exe, d = tool
    a = _picture(45, 70, 3, 7)
    Image.fromarray(a, "RGB").save(d / "a.ppm")
    Image.fromarray(a[..., 0], "L").save(d / "a.pgm")
    Image.fromarray(a, "RGB").save(d / "a.bmp")
    Image.fromarray(a, "RGB").save(d / "a.tga")
    Image.fromarray(a, "RGB").save(d / "a_rle.tga", compression="tga_rle")
    rgba = _picture(31, 29, 4, 8)
    Image.fromarray(rgba, "RGBA").save(d / "b.tga")
    for f, want in (("a.ppm", a), ("a.pgm", a[..., :1]), ("a.bmp", a), ("a.tga", a), ("a_rle.tga", a), ("b.tga", rgba)):
        _run(exe, "decode8", d / f, d / "o.raw")
        assert np.array_equal(_raw(d / "o.raw", np.uint8), want), f
    # a JPEG says what to do instead of decoding garbage
    Image.fromarray(a, "RGB").save(d / "a.jpg")
    r = subprocess.run([str(exe), "decode8", str(d / "a.jpg"), str(d / "o.raw")], capture_output=True, text=True)
    assert r.returncode == 1 and "JPEG decoding is not implemented" in r.stderr
    # ImageTexture(file, gammaCorrection) of the host API constructs from it
    assert "channels 3" in _run(exe, "texture8", d / "a.ppm", 1)


def test_hdr_and_pfm_decoders_against_opencv(tool):
    import cv2
    exe, d = tool
    rng = np.random.default_rng(3)
    for k, (h, w) in enumerate(((20, 40), (33, 7), (64, 128))):      # widths below 8 are written flat, the others run-length encoded
        a = (rng.random((h, w, 3)) ** 3 * 50).astype(np.float32)
        a[:4] = 0.25                                                 # long runs
        assert cv2.imwrite(str(d / f"e{k}.hdr"), a[..., ::-1])
        want = cv2.imread(str(d / f"e{k}.hdr"), cv2.IMREAD_UNCHANGED)[..., ::-1]
        _run(exe, "decodef", d / f"e{k}.hdr", d / "o.raw")
        got = _raw(d / "o.raw", np.float32)
        assert got.shape == want.shape
        # RGBE mantissa convention: OpenCV decodes (m + 0.5) / 256 * 2^e, stb_image (and this decoder) m / 256 * 2^e
        step = want.max(axis=2, keepdims=True) / 128.0               # one shared exponent per pixel: a mantissa step of the largest channel
        assert (np.abs(got - want) <= step + 1e-6).all() and (np.abs(got - a) <= 1.5 * step + 1e-6).all()
    a = rng.random((12, 17, 3)).astype(np.float32)
    (d / "p.pfm").write_bytes(b"PF\n17 12\n-1.0\n" + a[::-1].tobytes())
    _run(exe, "decodef", d / "p.pfm", d / "o.raw")
    assert np.array_equal(_raw(d / "o.raw", np.float32), a)
    # an LDR file through the float loader: (v / 255) ^ 2.2 (stbi_loadf's default)
    from PIL import Image
    u = _picture(9, 11, 3, 4)
    Image.fromarray(u, "RGB").save(d / "l.png")
    _run(exe, "decodef", d / "l.png", d / "o.raw")
    assert np.allclose(_raw(d / "o.raw", np.float32), (u / 255.0) ** 2.2, rtol=1e-5, atol=1e-7)


def _film_values(h, w, seed):
    rng = np.random.default_rng(seed)
    a = (rng.random((h, w, 3)) ** 4 * 8).astype(np.float32)           # well beyond 1: the tone mapper matters
    a[0, :5] = 0.0
    a[1, :5] = 1e4
    return a


@needs_ref
def test_film_writers_against_the_reference_writers(tool):
    """PPM: the same bytes.  PNG: the same pixels (the reference's bytes come from stb's deflate, ours from stored blocks).
    JPEG: both decode, and to the same picture within what two encoders at the same quality differ by."""
    from PIL import Image
    from oracle import refapi
    exe, d = tool
    ref = refapi.load("strict")
    ref.ref_film_write.restype = C.c_int
    ref.ref_film_write.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_char_p, C.c_int, C.c_int]
    for k, (h, w) in enumerate(((48, 64), (17, 23))):
        a = _film_values(h, w, k)
        (d / "film.f32").write_bytes(a.tobytes())
        for fmt_id, fmt in enumerate(("ppm", "png", "jpg")):
            assert ref.ref_film_write(a.ctypes.data, w, h, str(d / f"ref{k}").encode(), fmt_id, 90) == 0
            _run(exe, "film", d / "film.f32", w, h, d / f"our{k}", fmt, 90)
        assert (d / f"ref{k}.ppm").read_bytes() == (d / f"our{k}.ppm").read_bytes()
        rp, op = np.array(Image.open(d / f"ref{k}.png")), np.array(Image.open(d / f"our{k}.png"))
        assert rp.shape == (h, w, 3) and np.array_equal(rp, op)
        # the PNG pixels are the PPM pixels
        assert np.array_equal(np.frombuffer((d / f"our{k}.ppm").read_bytes()[-h * w * 3:], np.uint8).reshape(h, w, 3), op)
        rj, oj = np.array(Image.open(d / f"ref{k}.jpg")).astype(int), np.array(Image.open(d / f"our{k}.jpg")).astype(int)
        assert rj.shape == oj.shape == (h, w, 3)
        err_ref, err_our = np.abs(rj - rp).mean(), np.abs(oj - op).mean()
        assert err_our <= 1.25 * err_ref + 0.5, (err_our, err_ref)    # as faithful to the 8-bit image as the reference's encoder
        assert np.abs(rj - oj).mean() <= err_ref + err_our + 0.5


def test_jpeg_encoder_decodes_at_every_quality(tool):
    from PIL import Image
    exe, d = tool
    a = _film_values(40, 56, 9)
    (d / "f.f32").write_bytes(a.tobytes())
    _run(exe, "film", d / "f.f32", 56, 40, d / "q", "png")
    want = np.array(Image.open(d / "q.png")).astype(int)
    last = None
    for q in (5, 50, 90, 100):
        _run(exe, "film", d / "f.f32", 56, 40, d / f"q{q}", "jpg", q)
        got = np.array(Image.open(d / f"q{q}.jpg")).astype(int)
        err = np.abs(got - want).mean()
        assert last is None or err <= last + 0.2, (q, err, last)         # better quality, smaller error
        last = err
    assert last < 2.5


def test_arbitrary_sky_function_is_baked_with_a_reported_error(tool):
    exe, d = tool
    out = _run(exe, "sky")
    line = [l for l in out.splitlines() if l.startswith("baked")][0].split()
    assert line[1] == "1" and float(line[3]) < 0.02 and float(line[5]) < 1e-3, out       # smooth sky: max 2 %, mean 0.1 % of its mean luminance
    assert "gradient baked 0" in out                                                     # the GradientSky functor stays parametric
