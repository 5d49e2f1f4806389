"""Wall-clock of the Python calls on ordinary (pageable) numpy pages."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import libpillowfight_b200 as pf  # noqa: E402
from libpillowfight_b200.pages import A4_300DPI, synthetic_page  # noqa: E402

page = synthetic_page(*A4_300DPI, page_index=0)
calls = {"ace": lambda: pf.ace(page, seed=1), "gaussian": lambda: pf.gaussian(page),
         "sobel": lambda: pf.sobel(page), "canny": lambda: pf.canny(page)}
for name, fn in calls.items():
    for _ in range(3):
        fn()
    t0 = time.perf_counter()
    for _ in range(10):
        fn()
    print("%-9s %.2f ms per call (pageable numpy in, new numpy out)" % (name, (time.perf_counter() - t0) * 100))
