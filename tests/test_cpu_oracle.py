"""Pins the oracle (TEST INFRASTRUCTURE: unmodified reference sources + oracle/embree3_shim.cpp,
built into oracle/_ref/ by oracle/Makefile) before anything is compared against it:

  * the reference's own seven known-answer tests (CMakeLists.txt:94-129; src/tests/*.cpp) print
    SUCCESS when linked against the shim (tests/intersection.cpp is the one that pins the seam);
  * pcg32 known-answer vector (pcg.h:32-67) restated in integer arithmetic here;
  * the committed golden images are reproduced bit for bit by the oracle at any thread count;
  * mode A (per-tile streams, render.cpp:82) and mode B (per-(pixel,sample) streams) agree
    statistically, and spp slices of mode B add up to the full render (the multi-GPU partition).

oracle/_ref is git-ignored; __graft_entry__.build() creates it where /root/reference exists.
"""
import json
import os
import subprocess

import numpy as np
import pytest

from conftest import GOLDEN, ROOT

REFDIR = os.path.join(ROOT, "oracle", "_ref")
DRIVER = os.path.join(REFDIR, "ref_driver")
SCENES = os.path.join(REFDIR, "scenes")

needs_oracle = pytest.mark.skipif(not (os.path.exists(DRIVER) and os.path.isdir(SCENES)),
                                  reason="oracle/_ref not built (needs /root/reference; run __graft_entry__.build())")


def oracle_render(xml, res, mode, spp, threads=8, extra=()):
    from lajolla_public_b200.api import read_pfm
    out = f"/tmp/lj_test_oracle_{os.getpid()}.pfm"
    cmd = [DRIVER, "render", os.path.join(SCENES, xml), "-o", out, "--mode", mode, "--spp", str(spp),
           "--res", str(res[0]), str(res[1]), "-t", str(threads)] + list(extra)
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:]
    stats = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
    img = read_pfm(out)
    os.remove(out)
    return img, stats


@needs_oracle
@pytest.mark.parametrize("name", ["filter", "frame", "image", "intersection", "materials", "matrix", "mipmap"])
def test_reference_known_answer_tests(name):
    exe = os.path.join(REFDIR, "test_" + name)
    assert os.path.exists(exe)
    r = subprocess.run([exe], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300, cwd=REFDIR)
    assert r.returncode == 0, r.stdout[-1000:]
    assert "SUCCESS" in r.stdout and "FAIL" not in r.stdout


def pcg32_reals(stream, n):
    """pcg.h:32-67 in Python integers: init_pcg32(stream, seed 0x853c49e6748fea9b), next_pcg32_real<double>"""
    M = (1 << 64) - 1
    mult = 6364136223846793005
    inc = ((stream << 1) | 1) & M
    state = 0
    state = (state * mult + inc) & M          # first next_pcg32
    state = (state + 0x853c49e6748fea9b) & M
    state = (state * mult + inc) & M          # second next_pcg32
    out = []
    for _ in range(n):
        old = state
        state = (old * mult + inc) & M
        xorshifted = (((old >> 18) ^ old) >> 27) & 0xFFFFFFFF
        rot = old >> 59
        u32 = ((xorshifted >> rot) | (xorshifted << ((-rot) & 31))) & 0xFFFFFFFF
        bits = (u32 << 20) | 0x3FF0000000000000
        out.append(np.frombuffer(np.uint64(bits).tobytes(), dtype=np.float64)[0] - 1.0)
    return out


@needs_oracle
def test_pcg32_known_answer():
    r = subprocess.run([DRIVER, "rng", "64"], stdout=subprocess.PIPE, text=True)
    ref = [float(x) for x in r.stdout.split()]
    mine = pcg32_reals(7, 64)
    assert ref == mine
    assert all(0.0 <= x < 1.0 for x in ref)


@needs_oracle
@pytest.mark.parametrize("name,key", [("cbox", "A1"), ("cbox", "B1"), ("veach_mi", "A1"), ("veach_mi", "B1")])
def test_oracle_reproduces_committed_goldens(name, key):
    """the fixtures under tests/golden are outputs of the oracle, independent of thread count"""
    manifest = json.load(open(os.path.join(GOLDEN, "manifest.json")))
    e = manifest[name]
    gold = np.load(os.path.join(GOLDEN, f"{name}_{key}.npy"))
    for threads in (1, 5):
        img, st = oracle_render(e["xml"], e["res"], key[0], int(key[1:]), threads=threads)
        assert np.array_equal(img, gold)
        assert st["intersect_rays"] == e["images"][key]["intersect_rays"]
        assert st["occluded_rays"] == e["images"][key]["occluded_rays"]


@needs_oracle
def test_mode_b_slices_add_up():
    """spp slices [0,2) and [2,4) of per-sample streams average to the [0,4) render: the property the
    multi-GPU spp partition relies on (bench.py: rank r renders [r*S, (r+1)*S))"""
    full, _ = oracle_render("cbox/cbox.xml", (64, 64), "B", 4)
    a, _ = oracle_render("cbox/cbox.xml", (64, 64), "B", 4, extra=("--sample-range", "0", "2"))
    b, _ = oracle_render("cbox/cbox.xml", (64, 64), "B", 4, extra=("--sample-range", "2", "4"))
    np.testing.assert_allclose(0.5 * (a.astype(np.float64) + b), full, rtol=2e-6, atol=1e-7)


@needs_oracle
def test_mode_a_and_b_agree_statistically():
    """different stream layouts, same estimator: image means agree within Monte Carlo noise"""
    a, _ = oracle_render("cbox/cbox.xml", (64, 64), "A", 64)
    b, _ = oracle_render("cbox/cbox.xml", (64, 64), "B", 64)
    assert abs(a.mean() - b.mean()) < 0.02 * a.mean()
    rmse = float(np.sqrt(np.mean((a - b) ** 2)))
    assert rmse < 0.5 * float(a.mean()) + 0.2      # two independent 64-spp estimates


def test_golden_manifest_is_consistent():
    manifest = json.load(open(os.path.join(GOLDEN, "manifest.json")))
    for name, e in manifest.items():
        if not e["committed"]:
            continue
        assert os.path.exists(os.path.join(GOLDEN, name + ".ljs"))
        for key, meta in e["images"].items():
            img = np.load(os.path.join(GOLDEN, f"{name}_{key}.npy"))
            assert img.shape == (e["res"][1], e["res"][0], 3) and img.dtype == np.float32
            assert abs(float(img.mean()) - meta["mean"]) < 1e-6
            assert np.isfinite(img).all()


def test_imwrite_golden_is_the_references_output(tmp_path):
    """tests/golden/imwrite_kat.{exr,pfm} are what the reference's imwrite() (image.cpp:135-173) writes for
    imwrite_kat.f32img: re-minted here when the oracle binary is present, and in any case decoded and compared
    with the numpy restatement of tinyexr's float -> half conversion (tinyexr.h:889-924)."""
    import struct
    import subprocess
    from exr_util import half_bits_tinyexr, read_exr_half
    gold = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
    raw = open(os.path.join(gold, "imwrite_kat.f32img"), "rb").read()
    w, h = struct.unpack_from("<ii", raw)
    img = np.frombuffer(raw, dtype="<f4", offset=8).reshape(h, w, 3)
    exr = open(os.path.join(gold, "imwrite_kat.exr"), "rb").read()
    pfm = open(os.path.join(gold, "imwrite_kat.pfm"), "rb").read()
    ch, attrs = read_exr_half(exr)
    want = half_bits_tinyexr(img)
    assert attrs["compression"][1] == b"\x03"            # ZIP, as SaveEXR picks for images >= 16 px
    for ci, c in enumerate("RGB"):
        assert np.array_equal(ch[c], want[:, :, ci])
    assert pfm == b"PF\n%d %d\n-1\n" % (w, h) + img.tobytes()
    drv = os.path.join(os.path.dirname(gold), "..", "oracle", "_ref", "ref_driver")
    if os.path.exists(drv):
        for ext, committed in (("exr", exr), ("pfm", pfm)):
            out = tmp_path / ("kat." + ext)
            subprocess.run([drv, "imwrite", os.path.join(gold, "imwrite_kat.f32img"), str(out)], check=True)
            assert out.read_bytes() == committed
