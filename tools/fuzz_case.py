"""Replays one case of fuzz_parity.py's ACE check and describes the differences.
Usage: fuzz_case.py <seed> <case> [...]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import fuzz_parity as f  # noqa: E402
import libpillowfight_b200 as pf  # noqa: E402
from oracle import pf_oracle as o  # noqa: E402

seed = int(sys.argv[1])
want = set(int(a) for a in sys.argv[2:])
for case, img, g, ace in f.gen_cases(max(want) + 1, seed):
    if case not in want or ace is None:
        continue
    slope, limit, ns, sd = ace
    h, w = img.shape[:2]
    xy = o.ace_draw_samples(ns, w, h, sd)
    ref = o.ace_with_samples(img, xy, slope, limit, nb_threads=4)
    got = pf.ace(img, slope, limit, ns, seed=sd)
    sc = o.ace_scores(img, xy, slope, limit)
    print("case", case, img.shape, ace)
    for c in range(3):
        s = sc[c]
        d = got[..., c].astype(int) - ref[..., c].astype(int)
        idx = np.nonzero(d)
        print(" ch", c, "differ", len(idx[0]), "min/max", np.nanmin(s), np.nanmax(s))
        if len(idx[0]):
            pairs, cnt = np.unique(np.stack([ref[..., c][idx], got[..., c][idx], img[..., c][idx]], 1), axis=0,
                                   return_counts=True)
            for p, n in list(zip(pairs, cnt))[:8]:
                print("    ref %d got %d (pixel value %d): %d pixels" % (p[0], p[1], p[2], n))
            ss = s[idx]
            print("    scores of those: %d distinct, range %.17g .. %.17g" % (len(np.unique(ss)), ss.min(), ss.max()))
