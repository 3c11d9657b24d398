"""Host-side C++ scene API (path-trace_b200/host): same results as the reference's classes for
matrices and vectors (bit-exact, golden + live), the flattener's program shape, quirk Q6, error
behaviour, and the Radiance HDR reader/writer either side of the path."""
import ctypes as C
import math
import os

import numpy as np
import pytest

import common
import golden_cases as gc
from ptb200 import capi, device, scenes
from ptb200.builder import product_builder

KAT = np.load(os.path.join(common.GOLDEN, "kat.npz"))
PUSH, UNION, DIFF, INTER = 0, 1, 2, 3


def kinds(flat):
    return [op & 0xFF for op in flat.ops()]


def test_matrix_known_answers_bit_exact():
    b = product_builder()
    for k, m in gc.kat_matrices(b).items():
        assert np.array(m, np.float32).tobytes() == KAT["mat_" + k].tobytes(), k


def test_vector_known_answers_bit_exact():
    b = product_builder()
    v, n, ior = gc.kat_vectors()
    for i in range(len(v)):
        assert np.array(b.reflect(v[i], n[i]), np.float32).tobytes() == KAT["reflect"][i].tobytes()
        assert np.array(b.refract(v[i], float(ior[i]), n[i]), np.float32).tobytes() == KAT["refract"][i].tobytes()
        assert np.float32(b.refract_strength(v[i], float(ior[i]), n[i])).tobytes() == KAT["refract_strength"][i].tobytes()


def test_singular_matrix_raises_like_reference():
    b = product_builder()
    with pytest.raises(ArithmeticError):
        b.inverse(b.scale(1, 0, 1))
    world = b.transformed(b.scale(1, 0, 1), b.sphere((0, 0, 0), 1, b.material()))
    with pytest.raises(RuntimeError) as e:  # std::domain_error from flatten(), as invert(m) in makeSpanIterator()
        device.Flat(world)
    assert "singular" in str(e.value)


def test_flatten_program_is_postfix_with_balanced_unions():
    b = product_builder()
    flat = device.Flat(scenes.config2(b))
    assert flat.desc.n_prims == 8 and flat.desc.n_materials == 2
    # unionArray over 8 objects: ((0 1)(2 3))((4 5)(6 7))
    assert kinds(flat) == [PUSH, PUSH, UNION, PUSH, PUSH, UNION, UNION, PUSH, PUSH, UNION, PUSH, PUSH, UNION, UNION, UNION]
    p = flat.prims()
    assert p[0].type == 0 and list(p[0].p) == [0.0, 0.0, -4.0, 1.0]
    assert p[1].type == 1 and list(p[1].p) == [0.0, 1.0, 0.0, 1.0]  # d = -dot(n, pos)


def test_nested_transforms_stay_separate_chains():
    b = product_builder()
    m1, m2 = b.translate(1, 0, 0), b.scale(2)
    inner = b.transformed(m2, b.union(b.sphere((0, 0, 0), 1, b.material()), b.sphere((1, 0, 0), 1, b.material())))
    world = b.union(b.transformed(m1, inner), b.sphere((5, 0, 0), 1, b.material()))
    flat = device.Flat(world)
    d = flat.desc
    assert d.n_xforms == 2 and d.n_chains == 1 and d.chains[0].count == 2
    assert [d.chain_xf[i] for i in range(2)] == [0, 1]  # outermost first
    assert [p.chain for p in flat.prims()] == [0, 0, -1]
    assert list(d.xforms[0].m) == m1 and list(d.xforms[1].inv) == b.inverse(m2)


def test_combinator_transform_drops_b_quirk_q6():
    """Union::transform(m) builds both operands from `a` (reference include/union.h:19-22)."""
    b = product_builder()
    u = b.union(b.sphere((0, 0, 0), 1, b.material()), b.plane((0, 1, 0), 3, b.material()))
    t = b.transform(b.translate(0, 0, 1), u)
    flat = device.Flat(t)
    assert kinds(flat) == [PUSH, PUSH, UNION] and [p.type for p in flat.prims()] == [0, 0]


def test_transformed_material_and_texture_chain():
    b = product_builder()
    img = b.image(scenes.pattern_image(8, 8, 1))
    mat = b.material_transform(b.rotate_x(math.pi / 2), b.material(b.tex_multiply((1, 2, 3), b.tex_spherical(b.tex_image(img))), b.color(0.5)))
    flat = device.Flat(b.sphere((0, 0, 0), 1, mat))
    d = flat.desc
    chain = []
    i = d.materials[0].tex[0]
    while True:
        chain.append(d.texops[i].kind)
        if d.texops[i].kind < 8:
            break
        i += 1
    assert chain == [8, 9, 12, 1]  # AFFINE, MULTIPLY, SPHERICAL, IMAGE
    assert d.texops[d.materials[0].tex[1]].kind == 0  # ColorTexture::transform stays a ColorTexture
    assert d.n_images == 1 and d.images[0].width == 8


def test_hdr_round_trip_and_bytes(tmp_path):
    b = product_builder()
    rs = np.random.RandomState(3)
    img = np.ones((23, 150, 4), np.float32)
    img[..., :3] = rs.uniform(0, 4, (23, 150, 3)) * rs.choice([1e-3, 1.0, 300.0], (23, 150, 1))
    img[5:9, 20:140, :3] = 0.75  # long runs for the RLE paths
    img[10] = 0
    img[10, :, 3] = 1
    mine = str(tmp_path / "mine.hdr")
    b.write_hdr(img, mine)
    back = b.image_pixels(b.image_load(mine))
    assert back.shape == img.shape
    ok = img[..., :3].max(-1) > 1e-3
    rel = np.abs(back[..., :3] - img[..., :3])[ok].max(-1) if False else np.abs(back[..., :3] - img[..., :3]).max(-1)[ok] / img[..., :3].max(-1)[ok]
    assert rel.max() < 1 / 128.0  # 8-bit mantissa of the brightest channel, floor rounding
    if common.have_ref():
        rb = common.ref_builder()
        theirs = str(tmp_path / "theirs.hdr")
        rb.write_hdr(img, theirs)
        assert open(mine, "rb").read() == open(theirs, "rb").read()  # byte-identical writer
        assert rb.image_pixels(rb.image_load(mine)).tobytes() == back.tobytes()  # identical reader


def test_image_errors():
    b = product_builder()
    with pytest.raises(RuntimeError):
        b.image_load("/nonexistent/file.hdr")
    with pytest.raises(RuntimeError):
        b.image_load("/tmp/noextension")


# ---- PNG ingestion (host/ptb200_png.cpp; reference src/png_decoder.cpp:40-121 + src/image.cpp:60-73) ----
def _png_chunk(kind, body):
    import struct
    import zlib
    return struct.pack(">I", len(body)) + kind + body + struct.pack(">I", zlib.crc32(kind + body) & 0xFFFFFFFF)


def _png_filter(kind, row, prev, bpp):
    out = bytearray(len(row))
    for i in range(len(row)):
        a = row[i - bpp] if i >= bpp else 0
        b = prev[i]
        c = prev[i - bpp] if i >= bpp else 0
        if kind == 0:
            pred = 0
        elif kind == 1:
            pred = a
        elif kind == 2:
            pred = b
        elif kind == 3:
            pred = (a + b) >> 1
        else:
            p = a + b - c
            pa, pb, pc = abs(p - a), abs(p - b), abs(p - c)
            pred = a if pa <= pb and pa <= pc else b if pb <= pc else c
        out[i] = (row[i] - pred) & 0xFF
    return bytes(out)


def _png_file(w, h, depth, ctype, rows, palette=None, interlace=0, pieces=3):
    """rows: unfiltered scanline bytes; every filter type is used in turn; IDAT split into `pieces`."""
    import struct
    import zlib
    channels = {0: 1, 2: 3, 3: 1, 4: 2, 6: 4}[ctype]
    bpp = max(1, channels * depth // 8)
    raw, prev = b"", bytes(len(rows[0]))
    for y, row in enumerate(rows):
        raw += bytes([y % 5]) + _png_filter(y % 5, row, prev, bpp)
        prev = row
    z = zlib.compress(raw, 6)
    step = max(1, (len(z) + pieces - 1) // pieces)
    out = b"\x89PNG\r\n\x1a\n" + _png_chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, depth, ctype, 0, 0, interlace))
    out += _png_chunk(b"tEXt", b"Comment\0made by the test")
    if palette is not None:
        out += _png_chunk(b"PLTE", bytes(palette))
    for k in range(0, len(z), step):
        out += _png_chunk(b"IDAT", z[k:k + step])
    return out + _png_chunk(b"IEND", b"")


def test_png_ingestion_matches_libpng_transforms_of_the_reference(tmp_path):
    b = product_builder()
    rs = np.random.RandomState(5)
    w, h = 37, 11

    def load(data, name):
        path = str(tmp_path / name)
        open(path, "wb").write(data)
        return b.image_pixels(b.image_load(path))

    # 8-bit RGB: alpha is the 0 filler (png_set_filler(png_ptr, 0, PNG_FILLER_AFTER))
    rgb = rs.randint(0, 256, (h, w, 3)).astype(np.uint8)
    rgb[3] = rgb[2]  # rows where Up predicts exactly
    got = load(_png_file(w, h, 8, 2, [rgb[y].tobytes() for y in range(h)]), "rgb8.png")
    assert got.shape == (h, w, 4)
    assert np.array_equal(got[..., :3], rgb.astype(np.float32) / np.float32(255.0)) and (got[..., 3] == 0).all()
    # 8-bit RGBA
    rgba = rs.randint(0, 256, (h, w, 4)).astype(np.uint8)
    got = load(_png_file(w, h, 8, 6, [rgba[y].tobytes() for y in range(h)], pieces=1), "rgba8.png")
    assert np.array_equal(got, rgba.astype(np.float32) / np.float32(255.0))
    # 16-bit RGB(A): the high byte (png_set_strip_16)
    deep = rs.randint(0, 65536, (h, w, 4)).astype(">u2")
    got = load(_png_file(w, h, 16, 6, [deep[y].tobytes() for y in range(h)]), "rgba16.png")
    assert np.array_equal(got, (deep >> 8).astype(np.float32) / np.float32(255.0))
    # palette, 8 / 4 / 2 / 1 bits per index
    for depth in (8, 4, 2, 1):
        n = 1 << depth
        palette = rs.randint(0, 256, (n, 3)).astype(np.uint8)
        index = rs.randint(0, n, (h, w)).astype(np.uint8)
        rows = []
        for y in range(h):
            bits = "".join(format(int(v), "0%db" % depth) for v in index[y])
            bits += "0" * (-len(bits) % 8)
            rows.append(bytes(int(bits[k:k + 8], 2) for k in range(0, len(bits), 8)))
        got = load(_png_file(w, h, depth, 3, rows, palette=palette.reshape(-1).tolist()), "pal%d.png" % depth)
        assert np.array_equal(got[..., :3], palette[index].astype(np.float32) / np.float32(255.0)) and (got[..., 3] == 0).all()
    # what the reference cannot load either, or mangles: errors, not garbage
    gray = [bytes(w) for _ in range(h)]
    for data, name in ((_png_file(w, h, 8, 0, gray), "gray.png"), (_png_file(w, h, 8, 2, [rgb[y].tobytes() for y in range(h)], interlace=1), "adam7.png"),
                       (b"\x89PNG\r\n\x1a\n" + b"\0" * 40, "junk.png")):
        with pytest.raises(RuntimeError):
            load(data, name)
    good = bytearray(_png_file(w, h, 8, 2, [rgb[y].tobytes() for y in range(h)]))
    good[len(good) // 2] ^= 0x40  # a flipped bit inside IDAT: CRC error
    with pytest.raises(RuntimeError):
        load(bytes(good), "crc.png")
    # and a file of the reference's own kind (8-bit RGB written by a stock encoder), when PIL is here
    PIL = pytest.importorskip("PIL.Image")
    path = str(tmp_path / "pil.png")
    PIL.fromarray(rgb, "RGB").save(path, optimize=True)
    got = b.image_pixels(b.image_load(path))
    assert np.array_equal(got[..., :3], rgb.astype(np.float32) / np.float32(255.0)) and (got[..., 3] == 0).all()
