"""Seeded synthetic document pages for tests and benchmarks.

A page is document-like rather than white noise, because the cost of Canny's
hysteresis and ACE's saturation rate depend on content: noisy paper
background, a smooth illumination fall-off, rows of dark glyph-like strokes
with soft edges, and one colour "photo" block.  Pages are numpy uint8 arrays
of shape (height, width, 4), RGBA, alpha 255.
"""
import numpy as np

A4_300DPI = (2480, 3508)   # width, height
A4_600DPI = (4960, 7016)
PAGE_SEED_BASE = 0x5EED0000


def synthetic_page(width=A4_300DPI[0], height=A4_300DPI[1], page_index=0):
    rng = np.random.Generator(np.random.PCG64(PAGE_SEED_BASE + page_index))
    scale = max(1.0, width / 2480.0)
    yy = np.linspace(0.0, 1.0, height, dtype=np.float32)[:, None]
    xx = np.linspace(0.0, 1.0, width, dtype=np.float32)[None, :]

    page = np.empty((height, width, 3), dtype=np.float32)
    page[..., 0] = 235.0
    page[..., 1] = 230.0
    page[..., 2] = 215.0

    # "text": rows of dark strokes with one soft pixel on each side
    ink = np.zeros((height, width), dtype=np.float32)
    n_lines = 60
    margin_x = int(0.08 * width)
    margin_y = int(0.06 * height)
    pitch = (height - 2 * margin_y) / n_lines
    for line in range(n_lines):
        y0 = int(margin_y + line * pitch)
        glyph_h = max(4, int(pitch * 0.55))
        x = margin_x + int(rng.integers(0, int(40 * scale)))
        x_end = width - margin_x - int(rng.integers(0, int(400 * scale)))
        while x < x_end:
            sw = int(rng.integers(2, 13) * scale)
            gap = int(rng.integers(3, 15) * scale)
            top = y0 + int(rng.integers(0, max(1, glyph_h // 3)))
            bot = y0 + glyph_h - int(rng.integers(0, max(1, glyph_h // 4)))
            ink[top:bot, x:x + sw] = 1.0
            ink[top:bot, x - 1:x] = np.maximum(ink[top:bot, x - 1:x], 0.5)
            ink[top:bot, x + sw:x + sw + 1] = np.maximum(
                ink[top:bot, x + sw:x + sw + 1], 0.5)
            ink[top - 1:top, x:x + sw] = np.maximum(ink[top - 1:top, x:x + sw], 0.4)
            ink[bot:bot + 1, x:x + sw] = np.maximum(ink[bot:bot + 1, x:x + sw], 0.4)
            x += sw + gap
            if rng.random() < 0.12:
                x += int(25 * scale)    # word space
    page = page * (1.0 - ink[..., None]) + 30.0 * ink[..., None]

    # colour photo block of low-frequency sinusoids
    bw, bh = int(600 * scale), int(400 * scale)
    bx, by = int(0.55 * width), int(0.62 * height)
    bw = min(bw, width - bx)
    bh = min(bh, height - by)
    if bw > 0 and bh > 0:
        u = np.linspace(0.0, 1.0, bw, dtype=np.float32)[None, :]
        v = np.linspace(0.0, 1.0, bh, dtype=np.float32)[:, None]
        page[by:by + bh, bx:bx + bw, 0] = 128 + 100 * np.sin(6.3 * u + 2.1 * v)
        page[by:by + bh, bx:bx + bw, 1] = 128 + 100 * np.sin(4.4 * v + 1.3 * u + 1.0)
        page[by:by + bh, bx:bx + bw, 2] = 128 + 100 * np.cos(5.2 * (u + v))

    # illumination falling to 70 % toward the bottom-right corner
    light = 1.0 - 0.3 * (xx * yy)
    page *= light[..., None]

    # sensor noise, +-6 uniform per channel
    page += rng.integers(-6, 7, size=page.shape).astype(np.float32)

    out = np.empty((height, width, 4), dtype=np.uint8)
    out[..., :3] = np.clip(page, 0.0, 255.0).astype(np.uint8)
    out[..., 3] = 255
    return out


def uniform_random_page(width, height, seed=0):
    """Content-independent control page: i.i.d. uniform RGB, alpha 255."""
    rng = np.random.Generator(np.random.PCG64(seed))
    out = rng.integers(0, 256, size=(height, width, 4), dtype=np.uint8)
    out[..., 3] = 255
    return out
