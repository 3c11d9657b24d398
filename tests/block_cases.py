"""Seeded inputs for the tile-renderer / tile-text pins: one definition used by the golden
generator (tests/golden/make_golden.py, reference side), by tests/test_ref_blocks.py and by the GPU tests.

The reference's RenderBlock is compiled with its own constants (src/test.cpp:31-50): a 1920x1080 frame cut
into 32-pixel tiles, maximumSampleSize 9, minimumColorDelta 0.003; the lattice of pixel values it reads through
the substituted tracePixel is (34*32+1) rows of (60*32+1) points."""
import ctypes as C
import hashlib
import os

import numpy as np

W, H, BLOCK, MAX_SAMPLE, DELTA = 1920, 1080, 32, 9, 0.003
PITCH, ROWS = (W + BLOCK - 1) // BLOCK * BLOCK + 1, (H + BLOCK - 1) // BLOCK * BLOCK + 1
TILE_ROWS = (0, 16, 33)  # tile rows compared pixel by pixel with the Python restatement (33 = the ragged last one)
TEXT_TILES = ((64, 512), (1888, 1056))  # an interior tile; the bottom-right one, which sticks out of the screen


def lattice(seed=20261105):
    """A pixel-value function with everything the quadtree can meet: exactly flat areas (interpolated at
    size 8), gentle ramps below and above the 0.3 % threshold, noisy discs (subdivided to single pixels),
    a hard edge, black (|a| = 0: only an identical colour is "close") and HDR values."""
    rs = np.random.RandomState(seed)
    y, x = np.meshgrid(np.arange(ROWS, dtype=np.float64), np.arange(PITCH, dtype=np.float64), indexing="ij")
    rgb = np.empty((ROWS, PITCH, 3))
    rgb[..., 0] = 0.6
    rgb[..., 1] = 0.5 + 2e-5 * x            # far below the threshold
    rgb[..., 2] = 0.4 + 4e-4 * (y % 257)    # around it: 0.3 % of ~1 over 8 pixels is 3.75e-4 per pixel
    for _ in range(40):  # noisy discs
        cx, cy, r = rs.uniform(0, W), rs.uniform(0, H), rs.uniform(10, 90)
        m = (x - cx) ** 2 + (y - cy) ** 2 < r * r
        rgb[m] = rs.uniform(0, 3, (int(m.sum()), 3))
    rgb[:, 900:] *= 1.5                     # a hard vertical edge
    rgb[300:420, 200:500] = 0.0             # black
    rgb[600:700, 1200:1500] = (812.5, 40.25, 1e-3)  # HDR, exactly representable
    out = np.ones((ROWS, PITCH, 4), np.float32)
    out[..., :3] = rgb
    return out


def block_lib():
    """oracle/_ref/libptref_block.so: the reference's RenderBlock / NetRenderBlock / serverThreadFn."""
    import common
    path = os.path.join(common.ROOT, "oracle", "_ref", "libptref_block.so")
    if not os.path.exists(path):
        return None
    lib = C.CDLL(path)
    vp, i, sz = C.c_void_p, C.c_int, C.c_size_t
    lib.ref_block_constants.argtypes = [vp]
    lib.ref_block_frame.argtypes = [vp, i, i, vp]
    lib.ref_block_socket_text.argtypes = [vp, i, i, i, i, i, C.c_char_p, sz, C.c_char_p, sz]
    lib.ref_net_block_fetch.argtypes = [C.c_char_p, i, i, i, vp, vp, i]
    lib.ref_server_open.argtypes = [vp, i, i]
    return lib


def ref_frame(lib, lat):
    rgb = np.zeros((H, W, 3), np.float32)
    missing = lib.ref_block_frame(lat.ctypes.data, PITCH, ROWS, rgb.ctypes.data)
    return rgb, missing


def ref_socket_text(lib, lat, x, y, size=BLOCK):
    a, b = C.create_string_buffer(1 << 17), C.create_string_buffer(64)
    assert lib.ref_block_socket_text(lat.ctypes.data, PITCH, ROWS, x, y, size, a, len(a), b, len(b)) == 0
    return a.value.decode(), b.value.decode()


def digest(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def restated_tile_row(lat, tile_row):
    """tests/adaptive_ref.py (the Python restatement the GPU's adaptive renderer is checked against) on one
    row of tiles: (rows of the frame [32*tile_row, ...), source map)."""
    import adaptive_ref

    def pixel(x, y):
        return lat[y, x, :3].copy()

    y0 = tile_row * BLOCK
    screen = np.zeros((H, W, 3), np.float32)
    source = np.zeros((H, W), np.uint8)
    for tx in range(0, W, BLOCK):
        rb = adaptive_ref.RenderBlock(tx, y0, BLOCK, pixel, W, H, MAX_SAMPLE, DELTA)
        rb.run()
        rb.copy_to_buffer(screen, source, W, H)
    return screen[y0:min(H, y0 + BLOCK)], source[y0:min(H, y0 + BLOCK)]
