"""The C-ABI library loads and exports every symbol include/ccm_orb.h declares; argument validation and the
"no CPU fallback" rule are checked without touching a GPU."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    names = []
    for fn in sorted(os.listdir(os.path.join(ROOT, "include"))):
        if fn.endswith(".h"):
            txt = open(os.path.join(ROOT, "include", fn)).read()
            txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
            names += re.findall(r"\b(ccm_[a-z0-9_]+)\s*\(", txt)
    return sorted(set(names))


def test_header_symbols_exported(cuda_lib):
    from motioncheck_ccm_slam_b200 import _lib
    syms = declared_symbols()
    assert len(syms) >= 14
    for s in syms:
        assert hasattr(cuda_lib, s), "missing export " + s
        assert s in _lib.SIGNATURES, "binding missing for " + s
    assert sorted(_lib.SIGNATURES) == syms


def test_bad_config_rejected(cuda_lib):
    from motioncheck_ccm_slam_b200._lib import CCM_E_BADARG, OrbConfig
    h = ctypes.c_void_p()
    for cfg in [OrbConfig(0, 1.2, 8, 20, 7, 752, 480, 1, 0), OrbConfig(1000, 1.0, 8, 20, 7, 752, 480, 1, 0),
                OrbConfig(1000, 1.2, 17, 20, 7, 752, 480, 1, 0), OrbConfig(1000, 1.2, 8, 0, 7, 752, 480, 1, 0),
                OrbConfig(1000, 1.2, 8, 20, 7, 5000, 480, 1, 0), OrbConfig(1000, 1.2, 8, 20, 7, 752, 480, 0, 0)]:
        assert cuda_lib.ccm_orb_create(ctypes.byref(cfg), ctypes.byref(h)) == CCM_E_BADARG
        assert not h.value
        assert b"bad ccm_orb_config" in cuda_lib.ccm_last_error(None)
    assert cuda_lib.ccm_orb_create(None, ctypes.byref(h)) == CCM_E_BADARG


def test_no_cpu_fallback_without_gpu(cuda_lib):
    """Without a CUDA device the product path must fail loudly, never compute on the host."""
    try:
        import torch
        if torch.cuda.is_available():
            pytest.skip("GPU present")
    except ImportError:
        pass
    from motioncheck_ccm_slam_b200 import ORBextractor
    from motioncheck_ccm_slam_b200._lib import CCM_E_CUDA, CcmError
    with pytest.raises(CcmError) as e:
        ORBextractor(1000, 1.2, 8, 20, 7)
    assert e.value.code == CCM_E_CUDA and "no CPU fallback" in str(e.value)


def test_product_does_not_import_oracle():
    """Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may touch oracle/."""
    pkg = os.path.join(ROOT, "motioncheck_ccm_slam_b200")
    for dirpath, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                txt = open(os.path.join(dirpath, fn), errors="ignore").read()
                assert "liborb_oracle" not in txt and "_oracle" not in txt and "oracle/" not in txt.replace("oracle/cv_primitives.cpp linear_table", ""), fn


def test_matcher_no_cpu_fallback_and_badargs(cuda_lib):
    from motioncheck_ccm_slam_b200._lib import CCM_E_BADARG, CCM_E_CUDA
    h = ctypes.c_void_p()
    assert cuda_lib.ccm_matcher_create(0, None) == CCM_E_BADARG
    assert cuda_lib.ccm_matcher_create(-1, ctypes.byref(h)) in (CCM_E_CUDA,)
    assert not h.value
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except ImportError:
        has_gpu = False
    if not has_gpu:
        assert cuda_lib.ccm_matcher_create(0, ctypes.byref(h)) == CCM_E_CUDA and not h.value
        assert len(cuda_lib.ccm_matcher_last_error(None)) > 0
    # null handles never dereference
    assert cuda_lib.ccm_matcher_sync(None) == CCM_E_BADARG
    assert cuda_lib.ccm_frame_pack_create(None, 10, ctypes.byref(h)) == CCM_E_BADARG
    assert cuda_lib.ccm_search_window(None, None, None, 0, None, None, None, None, None) == CCM_E_BADARG
    assert cuda_lib.ccm_match_allpairs_best2(None, None, 0, None, 0, 0, None) == CCM_E_BADARG
    assert cuda_lib.ccm_matcher_launch_count(None) == 0
