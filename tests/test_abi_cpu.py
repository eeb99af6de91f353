"""CPU (-m "not gpu"): the C-ABI library loads, exports every symbol include/orbb200.h declares, validates arguments,
and FAILS LOUDLY (ORBB_E_CUDA) instead of falling back when no CUDA device is present.  No compute calls here."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    txt = open(os.path.join(ROOT, "include", "orbb200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(orb[xmv]_[a-z0-9_]+|orbb200_[a-z0-9_]+)\s*\(", txt)))


def test_header_symbols_all_exported():
    import orbb200
    L = orbb200.lib()
    names = _declared_symbols()
    assert len(names) >= 30
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, missing
    assert sorted(orbb200.SYMBOLS) == names, "harness symbol list out of sync with include/orbb200.h"


def test_struct_layouts_match_header():
    import orbb200
    assert orbb200.KP_DTYPE.itemsize == 28            # cv::KeyPoint
    assert orbb200.QUERY_DTYPE.itemsize == 52 and orbb200.PROJ_QUERY_DTYPE.itemsize == 72
    assert C.sizeof(orbb200.OrbConfig) == 40


def test_strerror_and_host_only_entry_points():
    import orbb200
    L = orbb200.lib()
    for code, word in [(0, "ok"), (-1, "invalid"), (-2, "CUDA"), (-3, "capacity"), (-4, "shape"), (-5, "memory"), (-6, "order")]:
        assert word.lower() in L.orbb200_strerror(code).decode().lower()
    # ratio test (ORBmatcher.cpp:278-281) is host arithmetic on the results: d1 <= TH_LOW && (float)d1 < ratio*(float)d2
    d1 = np.array([10, 50, 51, 30, 0, 45], np.int32); d2 = np.array([20, 100, 100, 40, 0, 60], np.int32)
    acc = orbb200.ratio_test(d1, d2, th_low=50, ratio=0.75)
    assert acc.tolist() == [True, True, False, False, False, False]     # 30 < 0.75*40 = 30 is false; 45 < 45 is false
    assert L.orbm_ratio_test(None, None, 3, 50, C.c_float(0.75), None) == orbb200.E_INVALID


def test_argument_validation_without_device():
    import orbb200
    L = orbb200.lib()
    h = C.c_void_p()
    assert L.orbx_create(None, C.byref(h)) == orbb200.E_INVALID
    bad = orbb200.OrbConfig(1000, 1.2, 8, 5, 7, 640, 480, 1, 0, 0)          # iniTh < minTh
    assert L.orbx_create(C.byref(bad), C.byref(h)) == orbb200.E_INVALID
    bad = orbb200.OrbConfig(1000, 1.2, 40, 20, 7, 640, 480, 1, 0, 0)        # nlevels > ORBB_MAX_LEVELS
    assert L.orbx_create(C.byref(bad), C.byref(h)) == orbb200.E_INVALID
    bad = orbb200.OrbConfig(1000, 1.2, 8, 20, 7, 100000, 480, 1, 0, 0)
    assert L.orbx_create(C.byref(bad), C.byref(h)) == orbb200.E_BAD_SHAPE
    assert L.orbx_destroy(None) == orbb200.E_INVALID
    assert L.orbx_sync(None) == orbb200.E_INVALID
    assert L.orbm_db_rows(None) == -1
    n = C.c_int()
    assert L.orbx_extract(None, None, 0, 0, 0, None, None, 0, C.byref(n)) == orbb200.E_INVALID


def test_no_cpu_fallback_without_device():
    """Without a GPU every compute entry point reports ORBB_E_CUDA; nothing is computed on the host."""
    import orbb200
    L = orbb200.lib()
    if L.orbb200_device_count() > 0:
        pytest.skip("a CUDA device is present: the failure path is exercised on the CPU-only container")
    h = C.c_void_p()
    cfg = orbb200.OrbConfig(1000, 1.2, 8, 20, 7, 640, 480, 1, 0, 0)
    assert L.orbx_create(C.byref(cfg), C.byref(h)) == orbb200.E_CUDA and not h.value
    assert b"no CPU fallback" in L.orbb200_last_cuda_error()
    with pytest.raises(orbb200.OrbError) as e:
        orbb200.ORBextractor(1000)
    assert e.value.code == orbb200.E_CUDA
    a = np.zeros((4, 32), np.uint8); out = np.zeros(4, np.int32)
    assert L.orbm_distance_pairs(a.ctypes.data_as(C.c_void_p), a.ctypes.data_as(C.c_void_p), 4, out.ctypes.data_as(C.c_void_p), 0) == orbb200.E_CUDA
    db = C.c_void_p()
    assert L.orbm_db_create(C.byref(db), 100, 0, 0) == orbb200.E_CUDA
    with pytest.raises(orbb200.OrbError):
        orbb200.popc_peak(0)


def test_product_never_imports_the_oracle():
    """oracle/ is test infrastructure: nothing under the package or include/ may reference it."""
    pkg = os.path.join(ROOT, "orb-slam2-annotation_b200")
    offenders = []
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp", ".cc")) or f == "Makefile":
                txt = open(os.path.join(dp, f), errors="ignore").read()
                if re.search(r"pyoracle|liborb_oracle|liborb_ref|oracle/|orb_extract_oracle|cv_prims", txt):
                    offenders.append(os.path.relpath(os.path.join(dp, f), ROOT))
    assert not offenders, offenders
