"""SURVEY.md §8 rows F3 and F4 pinned to the REFERENCE's own code (no GPU needed).

oracle/_ref/libptref_block.so is the reference's RenderBlock / NetRenderBlock / serverThreadFn compiled from
src/test.cpp:1-676, 681-720 (extracted at build time by oracle/Makefile; oracle/ref_block_shim.cpp explains the
one substitution: tracePixel becomes a look-up into a lattice of pixel values).  Checked here:

* F3: tests/adaptive_ref.py — the Python restatement that tests/test_gpu_adaptive.py holds the GPU's
  ptb200_render_adaptive to — reproduces the reference's frame bit for bit on whole rows of tiles, including the
  ragged last row, for a lattice with flat, ramped, noisy, black and HDR regions;
* F4: ptb200_tile_encode writes the bytes RenderBlock::copyToSocket writes; ptb200_tile_decode reads what
  NetRenderBlock::run reads (a live TCP exchange with the reference's client on 127.0.0.1);
  ptb200_tile_request_parse accepts and refuses the request lines serverThreadFn accepts and refuses, and the
  stream serverThreadFn sends decodes to the block's pixels.

Where /root/reference is absent (the GPU box) the same checks run against tests/golden/blocks.npz, written by
tests/golden/make_golden.py from this library."""
import ctypes as C
import os
import socket
import threading

import numpy as np
import pytest

import block_cases as bc
import common
from ptb200 import capi

GOLD = np.load(os.path.join(common.GOLDEN, "blocks.npz")) if os.path.exists(os.path.join(common.GOLDEN, "blocks.npz")) else None
LIB = bc.block_lib()
need_ref = pytest.mark.skipif(LIB is None, reason="oracle/_ref/libptref_block.so not built (no /root/reference here)")


def rgba_frame(rgb):
    """A ptb200 frame (alpha = 1 where a block wrote) from the reference's screen buffer."""
    out = np.zeros(rgb.shape[:2] + (4,), np.float32)
    ok = ~np.isnan(rgb[..., 0])
    out[..., :3] = np.where(ok[..., None], rgb, 0)
    out[..., 3] = ok
    return out


def encode(frame, x, y, size, sent=None, final=False):
    lib = capi.product()
    h, w, _ = frame.shape
    n = C.c_size_t()
    sp = sent.ctypes.data if sent is not None else None
    assert lib.ptb200_tile_encode(frame.ctypes.data, w, h, x, y, size, sp, int(final), None, 0, C.byref(n)) == 0
    buf = C.create_string_buffer(n.value + 1)
    assert lib.ptb200_tile_encode(frame.ctypes.data, w, h, x, y, size, sp, int(final), buf, len(buf), C.byref(n)) == 0
    return buf.value.decode()


def decode(text, x, y, size):
    rgb = np.zeros((size, size, 3), np.float32)
    valid = np.zeros((size, size), np.uint8)
    fin, used = C.c_int(), C.c_size_t()
    rc = capi.product().ptb200_tile_decode(text.encode(), len(text), x, y, size, rgb.ctypes.data, valid.ctypes.data, C.byref(fin), C.byref(used))
    return rc, rgb, valid, fin.value, used.value


def on_screen(text):
    """The reference's blocks also hold (and send) the pixels of a ragged tile that lie beyond the screen
    (renderSquare only stops at x > ScreenWidth, src/test.cpp:468); its client drops them in copyToBuffer.  A
    ptb200 frame has no such pixels."""
    keep = []
    for line in text.splitlines(True):
        if line.startswith("P"):
            px, py = line[1:line.index("=")].split(",")
            if int(px) >= bc.W or int(py) >= bc.H:
                continue
        keep.append(line)
    return "".join(keep)


# ---- F3 ------------------------------------------------------------------------------------------------------
@need_ref
def test_reference_constants_are_the_ones_the_library_defaults_to():
    k = (C.c_int * 8)()
    LIB.ref_block_constants(k)
    assert list(k)[:4] == [bc.W, bc.H, bc.BLOCK, bc.MAX_SAMPLE]
    assert np.frombuffer(bytes(k), np.float32)[6] == np.float32(bc.DELTA)


@need_ref
def test_python_restatement_equals_the_reference_render_block():
    lat = bc.lattice()
    rgb, missing = bc.ref_frame(LIB, lat)
    assert missing == 0 and not np.isnan(rgb).any()
    interpolated = (rgb != lat[:bc.H, :bc.W, :3]).any(2)
    assert 0.1 < interpolated.mean() < 0.9  # both outcomes of the quadtree occur
    for tr in bc.TILE_ROWS:
        rows, source = bc.restated_tile_row(lat, tr)
        y0 = tr * bc.BLOCK
        assert rows.tobytes() == rgb[y0:y0 + len(rows)].tobytes(), tr
        assert ((source == 2) >= interpolated[y0:y0 + len(rows)]).all()  # every changed pixel was an interpolated one


def test_python_restatement_equals_the_committed_reference_digests():
    assert GOLD is not None, "tests/golden/blocks.npz missing: run tests/golden/make_golden.py"
    lat = bc.lattice()
    for tr in bc.TILE_ROWS:
        rows, _ = bc.restated_tile_row(lat, tr)
        assert bc.digest(rows) == str(GOLD["frame_rows_sha256_%d" % tr]), tr


# ---- F4 ------------------------------------------------------------------------------------------------------
@need_ref
def test_tile_encode_writes_the_bytes_of_copy_to_socket():
    lat = bc.lattice()
    rgb, _ = bc.ref_frame(LIB, lat)
    frame = rgba_frame(rgb)
    for x, y in bc.TEXT_TILES:
        first, second = bc.ref_socket_text(LIB, lat, x, y)
        sent = np.zeros((bc.BLOCK, bc.BLOCK), np.uint8)
        assert encode(frame, x, y, bc.BLOCK, sent) == on_screen(first), (x, y)
        assert second == "E\n" and encode(frame, x, y, bc.BLOCK, sent) == second  # wrotePixel: a pixel goes out once
    x, y = bc.TEXT_TILES[1]
    assert len(on_screen(bc.ref_socket_text(LIB, lat, x, y)[0])) < len(bc.ref_socket_text(LIB, lat, x, y)[0])  # that tile does stick out


def test_tile_encode_writes_the_committed_reference_text():
    assert GOLD is not None, "tests/golden/blocks.npz missing: run tests/golden/make_golden.py"
    for k, (x, y) in enumerate(bc.TEXT_TILES):
        frame = np.zeros((bc.H, bc.W, 4), np.float32)
        tile = GOLD["tile_rgba_%d" % k]  # the reference's screen buffer over that tile (what its copyToBuffer pasted)
        frame[y:y + tile.shape[0], x:x + tile.shape[1]] = tile
        assert encode(frame, x, y, bc.BLOCK) == on_screen(bytes(GOLD["tile_text_%d" % k]).decode()), (x, y)


def serve_once(text, requests):
    """A one-connection server speaking the reference's protocol (src/test.cpp:686-716) with a canned stream."""
    srv = socket.socket()
    srv.bind(("127.0.0.1", 0))
    srv.listen(1)

    def run():
        conn, _ = srv.accept()
        conn.sendall(b"1\n")
        line = b""
        while not line.endswith(b"\n"):
            chunk = conn.recv(64)
            if not chunk:
                break
            line += chunk
        requests.append(line)
        conn.sendall(text.encode())
        conn.close()
        srv.close()

    threading.Thread(target=run, daemon=True).start()
    return srv.getsockname()[1]


@need_ref
def test_tile_decode_reads_what_the_reference_client_reads():
    lat = bc.lattice()
    for x, y in bc.TEXT_TILES:
        first, second = bc.ref_socket_text(LIB, lat, x, y)
        stream = "S\n" + first + second + "F\n"  # serverThreadFn: S, copyToSocket every second, a last one, F
        requests = []
        port = serve_once(stream, requests)
        want = np.zeros((bc.BLOCK, bc.BLOCK, 3), np.float32)
        want_valid = np.zeros((bc.BLOCK, bc.BLOCK), np.uint8)
        assert LIB.ref_net_block_fetch(("127.0.0.1:%d" % port).encode(), x, y, bc.BLOCK, want.ctypes.data, want_valid.ctypes.data, 20000) == 1
        assert requests == [("%d %d %d\n" % (x, y, bc.BLOCK)).encode()]  # the request line our parser must accept
        rc, rgb, valid, fin, used = decode(stream, x, y, bc.BLOCK)
        assert rc == 0 and fin == 1 and used == len(stream) - 1
        # NetRenderBlock::copyToBuffer clips to the screen it is given (x + size, y + size here: nothing clipped)
        assert np.array_equal(valid, want_valid) and rgb.tobytes() == want.tobytes()
        ex, ey, es = C.c_int(), C.c_int(), C.c_int()
        assert capi.product().ptb200_tile_request_parse(requests[0], bc.W, bc.H, C.byref(ex), C.byref(ey), C.byref(es)) == 0
        assert (ex.value, ey.value, es.value) == (x, y, bc.BLOCK)


def talk_to_reference_server(lat, request):
    """One connection to the reference's serverThreadFn: returns everything it sent."""
    fd = LIB.ref_server_open(lat.ctypes.data, bc.PITCH, bc.ROWS)
    assert fd >= 0
    s = socket.socket(fileno=fd)
    s.settimeout(30)
    s.sendall(request)
    try:
        s.shutdown(socket.SHUT_WR)
    except OSError:
        pass
    got = b""
    while True:
        chunk = s.recv(65536)
        if not chunk:
            break
        got += chunk
    s.close()
    return got.decode()


@need_ref
def test_request_check_and_framing_are_the_reference_servers():
    lat = bc.lattice()
    lib = capi.product()
    x, y, size = C.c_int(), C.c_int(), C.c_int()
    bad = [b"-1 0 32\n", b"0 0 0\n", b"0 0 2000\n", b"1921 0 32\n", b"0 1081 32\n", b"1 2\n", b"a b c\n", b"0 0 1000001\n"]
    good = [b"64 512 32\n", b" 1888  1056 32\n", b"0x40 01000 32\n"]  # %i: hex and octal too (64, 512)
    for line in bad:
        got = talk_to_reference_server(lat, line)
        assert got == "1\n", (line, got)  # greeted, then hung up without "S"
        assert lib.ptb200_tile_request_parse(line, bc.W, bc.H, C.byref(x), C.byref(y), C.byref(size)) == -1, line
    rgb, _ = bc.ref_frame(LIB, lat)
    for line in good:
        got = talk_to_reference_server(lat, line)
        assert lib.ptb200_tile_request_parse(line, bc.W, bc.H, C.byref(x), C.byref(y), C.byref(size)) == 0, line
        assert got.startswith("1\nS\n") and got.endswith("E\nF\n")
        rc, tile, valid, fin, used = decode(got[2:], x.value, y.value, size.value)
        assert rc == 0 and fin == 1
        want = rgb[y.value:y.value + size.value, x.value:x.value + size.value]
        h, w = want.shape[:2]
        assert valid[:h, :w].all()
        # %g keeps 6 significant digits: that is the protocol's precision (same check as the client's fscanf path above)
        np.testing.assert_allclose(tile[:h, :w], want, rtol=1e-5)
