"""CPU: the oracle (oracle/liboracle.so, the restatement every GPU parity test is judged by) against
the golden vectors produced by the reference's own compiled code (tests/golden/make_golden.py), and
- where /root/reference is present and oracle/_ref was built - against that code live.

Bar: bit-exact on every array (NaNs must sit in the same places).  The light-ray estimator of
src/graph is compared after its float sum, also bit-exact.
"""
import os

import numpy as np
import pytest

from cpu_checker import Checker, have_ref
from golden_cases import CASES

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_vectors.npz")


def bits_equal(a, b):
    a = np.ascontiguousarray(a)
    b = np.ascontiguousarray(b)
    if a.shape != b.shape or a.dtype != b.dtype:
        return False
    return bool((a.view(np.uint8) == b.view(np.uint8)).all())


def same_values(a, b):
    """bit-equal, or equal as numbers with NaN == NaN (different NaN payloads are not a parity matter)"""
    if bits_equal(a, b):
        return True
    a = np.asarray(a)
    b = np.asarray(b)
    if a.shape != b.shape or a.dtype != b.dtype or a.dtype.kind != "f":
        return False
    return bool(((a == b) | (np.isnan(a) & np.isnan(b))).all()) and bool((np.signbit(a) == np.signbit(b)).all())


@pytest.fixture(scope="module")
def golden():
    assert os.path.exists(GOLDEN), "tests/golden/reference_vectors.npz missing (tests/golden/make_golden.py)"
    return np.load(GOLDEN)


@pytest.mark.parametrize("case", sorted(CASES))
def test_oracle_matches_reference_vectors(orc, golden, case):
    got = CASES[case](orc)
    keys = [k for k in golden.files if k.startswith(case + "/")]
    assert len(keys) == len(got) and len(keys) > 0
    bad = [k for k in keys if not same_values(got[k.split("/", 1)[1]], golden[k])]
    assert not bad, f"{len(bad)} of {len(keys)} arrays differ from the reference: {bad[:5]}"


@pytest.mark.skipif(not have_ref(), reason="oracle/_ref not built (no /root/reference on this machine)")
@pytest.mark.parametrize("case", sorted(CASES))
def test_oracle_matches_reference_live(orc, case):
    ref = Checker("ref")
    got, want = CASES[case](orc), CASES[case](ref)
    bad = [k for k in want if not same_values(got[k], want[k])]
    assert not bad, bad[:5]


@pytest.mark.skipif(not have_ref(), reason="oracle/_ref not built (no /root/reference on this machine)")
def test_golden_file_is_current(golden):
    """The committed fixture is what the reference produces today (guards against a stale .npz)."""
    ref = Checker("ref")
    for case in ("scalar", "transform"):
        want = CASES[case](ref)
        for k, v in want.items():
            assert same_values(v, golden[f"{case}/{k}"]), k


def test_host_libm_is_the_reference_libm(orc):
    """The EXACT device path restates glibc's logf/sinf/cosf (avr_math.cuh); the oracle calls the host
    libm the way the reference does.  Pin a few known-answer values so a different libm is noticed."""
    x = np.array([0.5, 1.0, 2.0, 0.1, 7.25, 1e-20], dtype=np.float32)
    got = orc.libm(0, x)
    want = np.array([-0.6931472, 0.0, 0.6931472, -2.3025851, 1.9810015, -46.0517], dtype=np.float32)
    assert np.allclose(got, want, rtol=3e-7, atol=0)
    assert orc.libm(1, np.array([0.0], dtype=np.float32))[0] == 0.0
    assert orc.libm(2, np.array([0.0], dtype=np.float32))[0] == 1.0


# ---- VolPathIntegrator::Li as a whole: tolerance pin against the reference's own integrator ----------------
def _rel_mse(a, b):
    a = a.astype(np.float64); b = b.astype(np.float64)
    return float(np.mean((a - b) ** 2 / (b ** 2 + 1e-3)))


def _check_li(got, want):
    """got: the restatement (oracle), want: the reference's VolPathIntegrator::Li (oracle/refprobe_volpath.cpp).
    Bars: sampler-driven quantities (wavelengths, their pdfs, the host-side light direction and
    DistantLight::sceneRadius) bit-exact; "lockstep" scenes (every random number from the sampler streams):
    >= 99.5 % of pixel samples within 1e-3 relative radiance; "image" scenes (full estimator with NEE, whose
    shadow-ray RNG seeds hash ray coordinates): converged image within 3x the reference's own seed-to-seed
    relMSE (and within a tenth of it against the reference's image of the same seed, where the camera paths
    coincide), mean within max(5 %, 3x the reference's seed-to-seed mean difference)."""
    n_lock = n_img = 0
    for k in want:
        if k.endswith("_info"):
            assert same_values(got[k], want[k]), k
        elif k.endswith("_samples"):
            g, w = np.asarray(got[k]), np.asarray(want[k])
            assert same_values(g[:, 4:], w[:, 4:]), f"{k}: wavelengths / pdfs"
            L, Lw = g[:, :4].astype(np.float64), w[:, :4].astype(np.float64)
            assert (Lw.max(axis=1) > 0).mean() > 0.1
            scale = np.maximum(np.abs(Lw).max(axis=1), 1e-6)
            ok = np.abs(L - Lw).max(axis=1) <= 1e-3 * scale
            assert ok.mean() >= 0.995, f"{k}: only {ok.mean():.5f} of samples within 1e-3"
            n_lock += 1
        elif k.endswith("_mono0"):
            w = [np.asarray(want[k]), np.asarray(want[k[:-1] + "1"])]
            for seed in (0, 1):
                g = np.asarray(got[k[:-1] + str(seed)])
                other = w[1 - seed]
                # against the reference's image of the OTHER seed: no worse than the reference itself is
                assert _rel_mse(g, other) <= 3 * _rel_mse(w[seed], other), (k, seed)
                # against the reference's image of the SAME seed the camera paths coincide: far below the noise
                assert _rel_mse(g, w[seed]) <= 0.1 * min(_rel_mse(w[0], w[1]), _rel_mse(w[1], w[0])), (k, seed)
                assert abs(g.mean() - w[seed].mean()) <= max(0.05 * w[seed].mean(), 3 * abs(w[0].mean() - w[1].mean())), (k, seed)
            n_img += 1
    assert n_lock >= 2 and n_img >= 2


def _check_camera(orc, want):
    from golden_cases import camera_configs, camera_rays_restated
    for i in range(len(camera_configs())):
        rays, own = camera_rays_restated(orc, want[f"matrices{i}"], i)
        # GetCameraSample + GenerateRayDifferential + RenderFromCamera on the reference's matrices: bit-exact
        assert same_values(rays, want[f"rays{i}"]), f"camera rays {i}"
        # scenes.py composes the two matrices in float64, pbrt's constructor in fp32: equal to a few ulp
        assert np.allclose(own, want[f"matrices{i}"], rtol=2e-6, atol=1e-6), f"camera matrices {i}"


def test_oracle_camera_matches_reference_vectors(orc, golden):
    want = {k.split("/", 1)[1]: golden[k] for k in golden.files if k.startswith("camera/")}
    assert want, "camera/* missing from the fixture (tests/golden/make_golden.py)"
    _check_camera(orc, want)


@pytest.mark.skipif(not have_ref(), reason="oracle/_ref not built (no /root/reference on this machine)")
def test_oracle_camera_matches_reference_live(orc):
    from golden_cases import case_camera_reference
    _check_camera(orc, case_camera_reference(Checker("ref")))


def test_oracle_volpath_li_matches_reference_vectors(orc, golden):
    from golden_cases import case_volpath_li
    want = {k.split("/", 1)[1]: golden[k] for k in golden.files if k.startswith("volpath_li/")}
    assert want, "volpath_li/* missing from the fixture (tests/golden/make_golden.py)"
    _check_li(case_volpath_li(orc), want)


@pytest.mark.skipif(not have_ref(), reason="oracle/_ref not built (no /root/reference on this machine)")
def test_oracle_volpath_li_matches_reference_live(orc):
    from golden_cases import case_volpath_li
    _check_li(case_volpath_li(orc), case_volpath_li(Checker("ref")))
