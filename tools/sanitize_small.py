"""Small, odd-shaped pages through every entry point -- the program
compute-sanitizer (memcheck / racecheck) is run on."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import libpillowfight_b200 as pf  # noqa: E402
from libpillowfight_b200.pages import synthetic_page, uniform_random_page  # noqa: E402

for (w, h) in [(1, 1), (5, 4), (33, 17), (130, 67), (256, 200), (620, 300), (124, 513)]:
    page = uniform_random_page(w, h, seed=w + h)
    pf.gaussian(page, 2.0, 5)
    pf.gaussian(page, 1.0, 3)
    pf.gaussian(page, 3.0, 7)
    pf.gaussian(page, 2.0, 4)
    pf.sobel(page)
    pf.canny(page)
    if h >= 5:
        pf.ace(page, 10, 1000, 20, seed=1)
        pf.ace(page, 7.5, 1000, 20, seed=1)
doc = synthetic_page(480, 640, 0)
pf.canny(doc)
pf.gaussian(doc)
pf.ace(doc, samples=600, seed=2)
print("ok")
