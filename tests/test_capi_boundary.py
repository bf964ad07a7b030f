"""CPU: the C-ABI library loads and exports every symbol include/mslam_b200.h declares; argument validation that needs
no device; the product refuses to run without a GPU (no CPU fallback)."""
import ctypes as C
import os
import re

import pytest

import __graft_entry__ as graft

HEADER = os.path.join(graft.ROOT, "include", "mslam_b200.h")


def declared_symbols():
    txt = open(HEADER).read()
    return sorted(set(re.findall(r"MSLAM_B200_API\s+[\w\s\*]+?\b(mslam_b200_\w+)\s*\(", txt)))


def test_header_declares_the_expected_surface():
    names = declared_symbols()
    for must in ("mslam_b200_create", "mslam_b200_destroy", "mslam_b200_detect", "mslam_b200_detect_batch",
                 "mslam_b200_detect_batch_dev", "mslam_b200_describe", "mslam_b200_match_knn_ratio", "mslam_b200_knn2",
                 "mslam_b200_match_crosscheck", "mslam_b200_match_batch", "mslam_b200_match_batch_dev",
                 "mslam_b200_db_scores_dev", "mslam_b200_db_topk_dev", "mslam_b200_db_merge_topk_dev", "mslam_b200_db_query",
                 "mslam_b200_stage_times", "mslam_b200_measure_int_peaks"):
        assert must in names
    assert len(names) >= 25


def test_library_exports_every_declared_symbol(pkg):
    lib = pkg.library()
    missing = [n for n in declared_symbols() if not hasattr(lib, n)]
    assert not missing, missing


def test_version_defaults_and_status_strings(pkg):
    lib = pkg.library()
    assert lib.mslam_b200_version() == 100
    cfg = pkg.default_config()
    # the reference's hard-coded values: cv::ORB::create(1000) (orb_feature.cpp:25), ratio 0.7 (orb_feature.cpp:101)
    assert (cfg.nfeatures, cfg.nlevels, cfg.fast_threshold, cfg.edge_threshold) == (1000, 8, 20, 31)
    assert abs(cfg.scale_factor - 1.2) < 1e-6 and cfg.ratio == 0.7 and cfg.struct_size == C.sizeof(pkg.Config)
    for s in (0, -1, -2, -3, -4):
        assert lib.mslam_b200_status_string(s)


def test_create_validates_arguments(pkg):
    lib = pkg.library()
    h = C.c_void_p()
    assert lib.mslam_b200_create(None, C.byref(h)) == pkg.E_INVALID
    for bad in (dict(nlevels=0), dict(nlevels=17), dict(scale_factor=1.0), dict(edge_threshold=30), dict(max_batch=0),
                dict(fast_threshold=0), dict(max_width=5000), dict(struct_size=4)):
        cfg = pkg.default_config(**bad)
        assert lib.mslam_b200_create(C.byref(cfg), C.byref(h)) == pkg.E_INVALID, bad
        assert not h.value


def test_no_cpu_fallback_without_a_device(pkg):
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:  # pragma: no cover
        has_gpu = False
    if has_gpu:
        pytest.skip("a CUDA device is present")
    with pytest.raises(pkg.MslamB200Error) as e:
        pkg.Context()
    assert e.value.status == pkg.E_CUDA


def test_product_never_touches_the_oracle():
    pkg_dir = graft.PKG_DIR
    for root, _, files in os.walk(pkg_dir):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp", ".hpp", "Makefile")):
                txt = open(os.path.join(root, f), errors="ignore").read()
                assert "orb_oracle" not in txt and "from oracle" not in txt and "import oracle" not in txt, f
                assert "import cv2" not in txt, f
