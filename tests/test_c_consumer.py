"""The drop-in from C: tests/c/consumer.c includes the header as plain C99,
links libpillowfight_b200.so and calls the reference's own entry points
(void return, struct pf_bitmap), like a program written against libpillowfight."""
import os
import subprocess

import numpy as np
import pytest

from libpillowfight_b200.pages import synthetic_page

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIBDIR = os.path.join(ROOT, "libpillowfight_b200")


@pytest.fixture(scope="module")
def consumer(tmp_path_factory):
    exe = str(tmp_path_factory.mktemp("c") / "consumer")
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Wextra", "-Werror", "-pedantic",
                    "-I", os.path.join(ROOT, "include"), "-o", exe,
                    os.path.join(ROOT, "tests", "c", "consumer.c"),
                    "-L", LIBDIR, "-lpillowfight_b200", "-Wl,-rpath," + LIBDIR],
                   check=True, capture_output=True, text=True)
    return exe


def _run(exe, op, page, tmp_path):
    h, w = page.shape[:2]
    fin, fout = str(tmp_path / "in.raw"), str(tmp_path / "out.raw")
    page.tofile(fin)
    res = subprocess.run([exe, op, str(w), str(h), fin, fout], capture_output=True, text=True, timeout=300)
    out = np.fromfile(fout, dtype=np.uint8).reshape(h, w, 4) if res.returncode == 0 else None
    return res, out


def test_header_is_plain_c_and_the_library_links(consumer, tmp_path):
    """Builds with -std=c99 -pedantic -Werror; without a GPU the call reports
    the missing device (there is no CPU path) and leaves a non-zero status."""
    res, out = _run(consumer, "sobel", synthetic_page(16, 12, 0), tmp_path)
    assert res.returncode in (0, 2), res.stderr
    if res.returncode == 2:
        assert "no CPU path" in res.stderr


@pytest.mark.gpu
def test_c_program_gets_the_references_results(consumer, oracle, tmp_path):
    page = synthetic_page(180, 131, 3)
    for op, ref in (("sobel", oracle.sobel(page)), ("gaussian", oracle.gaussian(page, 2.0, 5)),
                    ("canny", oracle.canny(page))):
        res, out = _run(consumer, op, page, tmp_path)
        assert res.returncode == 0, res.stderr
        assert np.array_equal(out, ref), op
    # pf_apply_ace draws its samples from libc rand(): srand(12345) in the C
    # program, the same table from the oracle's draw
    res, out = _run(consumer, "ace", page, tmp_path)
    assert res.returncode == 0, res.stderr
    xy = oracle.ace_draw_samples(40, 180, 131, 12345)
    ref = oracle.ace_with_samples(page, xy, 10.0, 1000.0, nb_threads=4)
    d = np.abs(out.astype(np.int16) - ref.astype(np.int16))
    assert d.max() <= 1 and (d[..., :3] != 0).mean() <= 1e-3
