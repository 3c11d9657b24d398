"""the tile wire text of the reference's render farm — writer RenderBlock::copyToSocket
(reference src/test.cpp:372-391), framing :688-716, reader NetRenderBlock::run :596-630.  Host only."""
import ctypes as C

import numpy as np

from ptb200 import capi


def encode(frame, x, y, size, sent=None, final=False):
    lib = capi.product()
    h, w, _ = frame.shape
    n = C.c_size_t()
    sp = sent.ctypes.data if sent is not None else None
    before = sent.copy() if sent is not None else None
    assert lib.ptb200_tile_encode(frame.ctypes.data, w, h, x, y, size, sp, int(final), None, 0, C.byref(n)) == 0
    tiny = C.create_string_buffer(2)
    if n.value + 1 > 2:  # a buffer that is too small: PTB_ENOMEM, and nothing counts as sent
        assert lib.ptb200_tile_encode(frame.ctypes.data, w, h, x, y, size, sp, int(final), tiny, 2, C.byref(n)) == -3
    assert sent is None or (sent == before).all()  # measuring (out == NULL) and failing leave sent[] alone
    buf = C.create_string_buffer(n.value + 1)
    assert lib.ptb200_tile_encode(frame.ctypes.data, w, h, x, y, size, sp, int(final), buf, len(buf), C.byref(n)) == 0
    return buf.value.decode()


def decode(text, x, y, size):
    lib = capi.product()
    rgb = np.zeros((size, size, 3), np.float32)
    valid = np.zeros((size, size), np.uint8)
    fin, used = C.c_int(), C.c_size_t()
    rc = lib.ptb200_tile_decode(text.encode(), len(text), x, y, size, rgb.ctypes.data, valid.ctypes.data, C.byref(fin), C.byref(used))
    return rc, rgb, valid, fin.value, used.value


def test_writer_is_the_reference_printf():
    frame = np.zeros((4, 6, 4), np.float32)
    frame[1, 2] = (0.5, 1e-5, 123456789.0, 1)
    frame[2, 3] = (1.0, 0.1, 2.5, 3)      # alpha = finished samples: any positive value counts as valid
    frame[0, 0] = (9, 9, 9, 1)            # outside the tile
    text = encode(frame, 2, 1, 2, final=True)
    assert text == "P2,1=0.5,1e-05,1.23457e+08\nP3,2=1,0.1,2.5\nE\nF\n"
    # wrotePixel: a pixel goes out once
    sent = np.zeros((2, 2), np.uint8)
    assert encode(frame, 2, 1, 2, sent) == "P2,1=0.5,1e-05,1.23457e+08\nP3,2=1,0.1,2.5\nE\n"
    lib = capi.product()
    n = C.c_size_t()
    buf = C.create_string_buffer(64)
    lib.ptb200_tile_encode(frame.ctypes.data, 6, 4, 2, 1, 2, sent.ctypes.data, 0, buf, 64, C.byref(n))
    assert sent.sum() == 2
    lib.ptb200_tile_encode(frame.ctypes.data, 6, 4, 2, 1, 2, sent.ctypes.data, 0, buf, 64, C.byref(n))
    assert buf.value == b"E\n"
    # a tile that sticks out of the screen sends what exists
    assert encode(frame, 4, 2, 4) == "E\n"


def test_reader_round_trip_and_rejections():
    rs = np.random.RandomState(3)
    frame = np.concatenate([rs.rand(40, 50, 3).astype(np.float32) * 3, np.ones((40, 50, 1), np.float32)], 2)
    frame[5:9, 20:30, 3] = 0  # unfinished pixels are not sent
    x, y, size = 16, 0, 32
    text = "S\n" + encode(frame, x, y, size) + encode(frame, x, y, size, final=True)
    rc, rgb, valid, fin, used = decode(text, x, y, size)
    assert rc == 0 and fin == 1 and used == len(text) - 1  # stops at F like the reference (src/test.cpp:626-628)
    want_valid = frame[y:y + size, x:x + size, 3] > 0
    assert (valid.astype(bool) == want_valid).all()
    # %g keeps 6 significant digits: that is the protocol's precision
    np.testing.assert_allclose(rgb[want_valid], frame[y:y + size, x:x + size, :3][want_valid], rtol=1e-5)
    # stream cut in the middle of a record: consumed stops before it, nothing is lost
    cut = text.index("P", 200) + 4
    rc, _, v1, fin, used = decode(text[:cut], x, y, size)
    assert rc == 0 and fin == 0 and text[used] == "P" and used < cut
    # pixel outside the tile, unknown record (src/test.cpp:610-614, :629-631)
    assert decode("P0,0=1,1,1\n", x, y, size)[0] == -1
    assert decode("Q\n", x, y, size)[0] == -1
    # %i reads hex and octal like the reference's fscanf
    rc, rgb, valid, _, _ = decode("P0x11,010=1,2,3\n", x, y, size)
    assert rc == 0 and valid[8, 1] == 1 and tuple(rgb[8, 1]) == (1, 2, 3)


def test_request_line():
    lib = capi.product()
    x, y, s = C.c_int(), C.c_int(), C.c_int()
    assert lib.ptb200_tile_request_parse(b"64 32 32\n", 1920, 1080, C.byref(x), C.byref(y), C.byref(s)) == 0
    assert (x.value, y.value, s.value) == (64, 32, 32)
    for bad in (b"-1 0 32", b"0 0 0", b"0 0 2000", b"1921 0 32", b"0 1081 32", b"1 2", b"a b c"):
        assert lib.ptb200_tile_request_parse(bad, 1920, 1080, C.byref(x), C.byref(y), C.byref(s)) == -1
