"""CPU: the C-ABI shared library loads without a GPU and exports every symbol include/avr.h declares;
struct layouts seen from Python match the header's; a missing library fails loudly."""
import ctypes
import os
import re

import pytest

from acceleratedvolrenderer_b200 import _abi, api

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "avr.h")


def header_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    src = re.sub(r"//[^\n]*", "", src)
    return sorted(set(re.findall(r"\b(avr_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_the_expected_surface():
    names = header_functions()
    for must in ("avr_ctx_create", "avr_medium_create_uniformgrid", "avr_medium_create_rgbgrid", "avr_sample_tmaj",
                 "avr_dda_segments", "avr_render_volpath", "avr_graph_propagate", "avr_graph_connect"):
        assert must in names


def test_library_exports_every_declared_symbol():
    lib = api.load_library()  # no compute call: works on a machine without a GPU
    missing = [n for n in header_functions() if not hasattr(lib, n)]
    assert not missing, missing
    assert set(api.ABI_SYMBOLS) <= set(header_functions())


def test_struct_sizes_match_the_header():
    # avr_tmaj_result is 80 bytes (T_maj, Tr, r_u, r_l: 4 floats each; t_event; event; n_fetch; n_seg)
    assert _abi.TMAJ_DTYPE.itemsize == 80
    assert ctypes.sizeof(_abi.avr_render_stats) % 8 == 0


def test_version_and_error_string_without_gpu():
    lib = api.load_library()
    lib.avr_version.restype = ctypes.c_char_p
    assert lib.avr_version().decode().startswith("avr")


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    monkeypatch.setattr(api, "_lib", None)
    monkeypatch.setattr(api, "LIB_PATH", str(tmp_path / "nope.so"))
    with pytest.raises(ImportError):
        api.load_library()


def test_film_resolve_matches_the_reference_film():
    """avr_film_resolve (host code of the product, no GPU) against RGBFilm::GetPixelRGB of the reference
    (film.h:257-275; fixture case "film", produced by the reference's own inline film code): bit-exact."""
    import sys
    import numpy as np
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from golden_cases import film_case_inputs
    golden = np.load(os.path.join(ROOT, "tests", "golden", "reference_vectors.npz"))
    rd, _, _ = film_case_inputs()
    film = np.ascontiguousarray(golden["film/film"], dtype=np.float64)
    rgb = np.zeros(film.shape[:2] + (3,), dtype=np.float32)
    lib = api.load_library()
    st = lib.avr_film_resolve(ctypes.byref(rd.c), film.ctypes.data_as(ctypes.POINTER(ctypes.c_double)),
                              rgb.ctypes.data_as(ctypes.POINTER(ctypes.c_float)))
    assert st == 0
    assert rgb.tobytes() == np.ascontiguousarray(golden["film/rgb"]).tobytes()
    assert np.abs(rgb).max() > 0
