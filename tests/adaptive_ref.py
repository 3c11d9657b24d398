"""Test infrastructure: the reference's adaptive tile renderer restated on the CPU, depth-first and
per tile exactly as RenderBlock does it (reference src/test.cpp:325-518), over a pixel FUNCTION
supplied by the caller instead of tracePixel.  Used to check that ptb200_render_adaptive's
level-synchronous quadtree takes the same decisions and produces the same image for the same traced
values.  All arithmetic in numpy float32 scalars (IEEE binary32, no contraction), in the reference's
operation order."""
import numpy as np

F = np.float32


def abs_squared(v):
    """dot(v, v) = (x*x + y*y) + z*z  (reference include/vector3d.h:102-110)"""
    return F(F(F(v[0] * v[0]) + F(v[1] * v[1])) + F(v[2] * v[2]))


def color_close_enough(a, b, delta):
    """src/test.cpp:437-440"""
    d = np.array([F(a[0] - b[0]), F(a[1] - b[1]), F(a[2] - b[2])], F)
    return abs_squared(d) <= F(F(delta * delta) * abs_squared(a))


def lerp(a, b, f):
    """a + f * (b - a) per channel (src/test.cpp:428-434)"""
    return np.array([F(a[k] + F(f * F(b[k] - a[k]))) for k in range(3)], F)


class RenderBlock:
    """One tile: src/test.cpp:325-518.  pixel_fn(x, y) -> float32[3] stands in for tracePixel."""

    def __init__(self, x, y, size, pixel_fn, screen_w, screen_h, max_sample_size, delta):
        self.x0, self.y0, self.size = x, y, size
        self.buffer = np.zeros((size + 1, size + 1, 3), F)
        self.valid = np.zeros((size + 1, size + 1), bool)
        self.source = np.zeros((size + 1, size + 1), np.uint8)
        self.fn, self.sw, self.sh, self.mss, self.delta = pixel_fn, screen_w, screen_h, max_sample_size, F(delta)
        self.traced = 0

    def set_pixel(self, x, y, c, source):  # :408-421
        if x < self.x0 or x - self.x0 > self.size or y < self.y0 or y - self.y0 > self.size:
            return
        self.buffer[y - self.y0, x - self.x0] = c
        self.valid[y - self.y0, x - self.x0] = True
        self.source[y - self.y0, x - self.x0] = source

    def calc_pixel_color(self, x, y):  # :441-465
        if self.x0 <= x <= self.x0 + self.size and self.y0 <= y <= self.y0 + self.size and self.valid[y - self.y0, x - self.x0]:
            return self.buffer[y - self.y0, x - self.x0].copy()
        c = self.fn(x, y)
        self.traced += 1
        self.set_pixel(x, y, c, 1)
        return c

    def interpolate_square(self, x0, y0, size, tl, tr, bl, br):  # :423-436
        for y in range(size):
            fy = F(F(y) / F(size))
            l, r = lerp(tl, bl, fy), lerp(tr, br, fy)
            for x in range(size):
                self.set_pixel(x + x0, y + y0, lerp(l, r, F(F(x) / F(size))), 2)

    def render_square(self, x, y, size, tl, tr, bl, br):  # :466-499
        if x > self.sw or y > self.sh:
            return
        if size <= 1:
            self.set_pixel(x, y, tl, 1)
            return
        ce = lambda a, b: color_close_enough(a, b, self.delta)
        if ce(tl, tr) and ce(tl, bl) and ce(tl, br) and ce(tr, bl) and ce(tr, br) and ce(bl, br) and size <= self.mss:
            self.interpolate_square(x, y, size, tl, tr, bl, br)
            return
        half = size // 2
        cx, cy = x + half, y + half
        tc = self.calc_pixel_color(cx, y)
        cl = self.calc_pixel_color(x, cy)
        cc = self.calc_pixel_color(cx, cy)
        cr = self.calc_pixel_color(x + size, cy)
        bc = self.calc_pixel_color(cx, y + size)
        self.render_square(x, y, half, tl, tc, cl, cc)
        self.render_square(cx, y, half, tc, tr, cc, cr)
        self.render_square(x, cy, half, cl, cc, bl, bc)
        self.render_square(cx, cy, half, cc, cr, bc, br)

    def run(self):  # :501-507
        x, y, s = self.x0, self.y0, self.size
        self.render_square(x, y, s, self.calc_pixel_color(x, y), self.calc_pixel_color(x + s, y), self.calc_pixel_color(x, y + s), self.calc_pixel_color(x + s, y + s))

    def copy_to_buffer(self, screen, source, w, h):  # :359-371
        for y in range(self.y0, min(self.y0 + self.size, h)):
            for x in range(self.x0, min(self.x0 + self.size, w)):
                if self.valid[y - self.y0, x - self.x0]:
                    screen[y, x] = self.buffer[y - self.y0, x - self.x0]
                    source[y, x] = self.source[y - self.y0, x - self.x0]


def render_frame(pixel_fn, w, h, block_size, max_sample_size, delta):
    """The reference main loop's tiling (src/test.cpp:940-955): tiles at multiples of block_size."""
    screen = np.zeros((h, w, 3), F)
    source = np.zeros((h, w), np.uint8)
    traced = 0
    for ty in range(0, h, block_size):
        for tx in range(0, w, block_size):
            rb = RenderBlock(tx, ty, block_size, pixel_fn, w, h, max_sample_size, delta)
            rb.run()
            rb.copy_to_buffer(screen, source, w, h)
            traced += rb.traced
    return screen, source, traced
