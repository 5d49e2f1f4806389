"""Regenerates tests/golden/oracle_small.npz from the CPU oracle.

These are regression fixtures of OUR oracle (oracle/pf_oracle.c) on a seeded
96 x 128 synthetic page.  They are not upstream golden images: the reference
mount holds no sources, tests or images (top).  Run from the repo
root:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from libpillowfight_b200.pages import synthetic_page  # noqa: E402
from oracle import pf_oracle as o  # noqa: E402

page = synthetic_page(96, 128, 3)
xy = o.ace_draw_samples(24, 96, 128, 12345)
np.savez_compressed(
    os.path.join(ROOT, "tests", "golden", "oracle_small.npz"),
    page=page, xy=xy,
    gaussian=o.gaussian(page, 2.0, 5), sobel=o.sobel(page), canny=o.canny(page),
    ace=o.ace_with_samples(page, xy, 10.0, 1000.0))
print("written")
