"""The only upstream artefact that pins rendered output: out.bmp, an 800x800 / 1024 spp render of test_scene.txt
(src/main.rs:13-60) made with an OS-seeded RNG -- a STATISTICAL golden.  Its 8x8-block linear means are committed
in tests/golden/outbmp_block_means.npz (tests/golden/make_golden.py).  The oracle renders the same scene at 100x100
(one pixel per block, same footprint) with the reference's sequential xorshift RNG and must agree in the mean.

What this pins: BMP row 0 = image bottom, the camera matrix orientation, the un-normalised hemisphere sample
(1 - r1^2), the weight diffuse*dot(n,dir)/(samples*0.5), the depth cut `> 4`, to_srgb threshold semantics.
The flip controls show the test has teeth.
"""
import hashlib
import os

import numpy as np
import pytest

from conftest import REFERENCE, needs_reference
from test_gpu_parity import REF_TEXT_SCENE

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _block_stats(golden, lin_bgr):
    """ratio statistics of rendered linear BGR (100,100,3) against the golden block means on unsaturated 10x10 super-blocks"""
    g = golden["linear_bgr"].astype(np.float64)
    sat = golden["saturated"]
    ok = sat.reshape(10, 10, 10, 10, 3).max(axis=(1, 3, 4)) < 0.02  # super-blocks with (almost) no clipped samples
    gs = g.reshape(10, 10, 10, 10, 3).mean(axis=(1, 3))
    rs = lin_bgr.reshape(10, 10, 10, 10, 3).mean(axis=(1, 3))
    m = ok[:, :, None] & (gs > 2e-3)  # channels the red/green walls do not reflect are exactly 0 on both sides: no ratio
    return rs[m] / gs[m], int(ok.sum())


@pytest.fixture(scope="module")
def cornell_render(orc):
    from oracle import scene_text
    desc = scene_text.deserialize(REF_TEXT_SCENE)
    r = orc.OracleScene(desc).render(width=100, height=100, antialias=1024, max_depth=4, seed=2026, rng_mode=orc.RNG_SEQUENTIAL)
    return r


def test_statistical_match_with_upstream_render(orc, cornell_render):
    golden = np.load(os.path.join(GOLDEN, "outbmp_block_means.npz"))
    assert int(golden["width"]) == 800 and int(golden["height"]) == 800
    acc = cornell_render["accum"]                       # (100,100,3) RGB linear, row 0 = bottom
    # out.bmp stores clipped 8-bit codes: compare what a BMP would hold, i.e. the decoded quantised value
    v, _ = orc.srgb_tables()
    lin_bgr = v[orc.quantize_bgr(acc.reshape(-1, 3))].reshape(100, 100, 3)
    ratio, n = _block_stats(golden, lin_bgr)
    assert n >= 40, "too few unsaturated super-blocks (%d)" % n
    assert abs(np.median(ratio) - 1.0) < 0.015, np.median(ratio)
    assert np.abs(ratio - 1.0).max() < 0.12, (ratio.min(), ratio.max())
    assert np.abs(ratio - 1.0).mean() < 0.03
    # controls: a vertically / horizontally flipped image must be clearly worse
    for flipped in (lin_bgr[::-1], lin_bgr[:, ::-1]):
        r2, _ = _block_stats(golden, flipped)
        assert np.abs(r2 - 1.0).mean() > 2.0 * np.abs(ratio - 1.0).mean()


def test_counter_rng_renders_the_same_distribution(orc, cornell_render):
    """the counter RNG the GPU shares must sample the same distribution as the sequential reference RNG"""
    from oracle import scene_text
    desc = scene_text.deserialize(REF_TEXT_SCENE)
    r = orc.OracleScene(desc).render(width=100, height=100, antialias=1024, max_depth=4, seed=7)
    a = np.minimum(cornell_render["accum"], 1.0).reshape(10, 10, 10, 10, 3).mean(axis=(1, 3))  # clip like the 8-bit output does
    b = np.minimum(r["accum"], 1.0).reshape(10, 10, 10, 10, 3).mean(axis=(1, 3))
    m = a > 0.02
    ratio = b[m] / a[m]
    assert np.abs(ratio - 1.0).max() < 0.15, (ratio.min(), ratio.max())
    assert abs(np.median(ratio) - 1.0) < 0.015


@needs_reference
def test_ref_text_scene_restates_upstream_scene():
    """REF_TEXT_SCENE (restated values) parses to the same scene as upstream's test_scene.txt; fixtures are current."""
    from oracle import scene_text
    raw = open(os.path.join(REFERENCE, "test_scene.txt"), "rb").read()
    assert hashlib.sha256(raw).hexdigest() == open(os.path.join(GOLDEN, "test_scene_sha256.txt")).read().strip()
    a = scene_text.deserialize(raw.decode())
    b = scene_text.deserialize(REF_TEXT_SCENE)
    assert a == b
