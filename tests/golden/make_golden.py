#!/usr/bin/env python
"""Regenerate tests/golden/reference_vectors.npz.

Runs the cases of tests/golden_cases.py through oracle/_ref/libavr_refprobe.so, i.e. through the
reference's OWN templates and classes (pbrt GridMedium / RGBGridMedium constructors, SampleT_maj,
DDAMajorantIterator, Transform, RNG, Eigen SpMV, graph::SampleTransmittance, NanoVDBMedium over real nanovdb
grids, FreeGraph / LightingCalculator, VolPathIntegrator::Li) compiled in place from
/root/reference by oracle/Makefile.  That library only exists where /root/reference does; the .npz it
produces is the fixture that travels.  Usage:  make -C oracle && python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from cpu_checker import Checker, have_ref  # noqa: E402
from golden_cases import CASES, case_camera_reference, case_volpath_li  # noqa: E402


def main():
    if not have_ref():
        raise SystemExit("oracle/_ref/libavr_refprobe.so missing: run `make -C oracle` where /root/reference exists")
    ck = Checker("ref")
    out = {}
    for name, fn in CASES.items():
        for k, v in fn(ck).items():
            out[f"{name}/{k}"] = np.asarray(v)
    # the reference's own VolPathIntegrator::Li (compared by tolerance, tests/test_oracle_golden.py)
    for k, v in case_volpath_li(ck).items():
        out[f"volpath_li/{k}"] = np.asarray(v)
    # the reference's real PerspectiveCamera (matrices + rays); the restatement is run on these matrices
    for k, v in case_camera_reference(ck).items():
        out[f"camera/{k}"] = np.asarray(v)
    path = os.path.join(HERE, "reference_vectors.npz")
    np.savez_compressed(path, **out)
    print(f"wrote {path}: {len(out)} arrays, {os.path.getsize(path) / 1e6:.2f} MB")


if __name__ == "__main__":
    main()
