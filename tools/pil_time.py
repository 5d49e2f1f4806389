"""Times the Python entry points on a PIL image (what a user of upstream's
pillowfight module calls), step by step."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import libpillowfight_b200 as pf  # noqa: E402
from libpillowfight_b200.pages import A4_300DPI, synthetic_page  # noqa: E402
from PIL import Image  # noqa: E402
import PIL  # noqa: E402

w, h = A4_300DPI
arr = synthetic_page(w, h, 0)
img = Image.fromarray(arr, "RGBA")
rgb = img.convert("RGB")


def best(fn, n=7):
    fn()
    ts = []
    for _ in range(n):
        t = time.perf_counter()
        fn()
        ts.append(time.perf_counter() - t)
    return min(ts) * 1e3


print("Pillow", PIL.__version__)
print("sobel(numpy array)      %.2f ms" % best(lambda: pf.sobel(arr)))
print("sobel(PIL RGBA)         %.2f ms" % best(lambda: pf.sobel(img)))
print("sobel(PIL RGB)          %.2f ms" % best(lambda: pf.sobel(rgb)))
print("ace(PIL RGBA)           %.2f ms" % best(lambda: pf.ace(img, seed=1)))
print("  img.convert('RGBA')   %.2f ms" % best(lambda: img.convert("RGBA")))
print("  rgb.convert('RGBA')   %.2f ms" % best(lambda: rgb.convert("RGBA")))
print("  img.tobytes()         %.2f ms" % best(lambda: img.tobytes()))
b = img.tobytes()
print("  Image.frombytes       %.2f ms" % best(lambda: Image.frombytes("RGBA", (w, h), b)))
print("  Image.frombuffer      %.2f ms" % best(lambda: Image.frombuffer("RGBA", (w, h), b, "raw", "RGBA", 0, 1)))
out = pf.sobel(img)
assert np.array_equal(np.asarray(out), pf.sobel(arr))
print("  pf._to_array(RGBA)    %.2f ms" % best(lambda: pf._to_array(img)))
a2, _ = pf._to_array(img)
print("  pf._run on that array %.2f ms" % best(lambda: pf._run(pf.lib.pf_b200_sobel, a2)))
res = pf._run(pf.lib.pf_b200_sobel, a2)
print("  pf._from_array        %.2f ms" % best(lambda: pf._from_array(res, True)))
print("  np.frombuffer(tobytes)%.2f ms" % best(lambda: np.frombuffer(img.tobytes(), dtype=np.uint8)))
