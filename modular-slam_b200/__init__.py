"""modular-slam_b200 -- B200-native ORB detect/describe + Hamming matching hot path of modular-slam.

Product code lives in csrc/ (sm_100a CUDA kernels + the C ABI of include/mslam_b200.h) and host/ (the C++17 plugin
adapter implementing the reference's IFeatureDetector / IFeatureMatcher).  This Python package is the ctypes
harness over the same C ABI used by tests/ and bench.py.  Imported as `modular_slam_b200` through
__graft_entry__.load_package() (the directory name carries a hyphen).
"""
from .capi import (Context, Config, Camera, PnpParams, pnp_params, MslamB200Error, KP_DTYPE, LIB_PATH, library, default_config, pinned_empty, db_shard_range, nccl_unique_id, NCCL_ID_BYTES,  # noqa: F401
                   OK, E_INVALID, E_CUDA, E_CAPACITY, E_NOMEM, STAGE_PYRAMID, STAGE_BLURRED, STAGE_N_FAST,
                   STAGE_N_CUT1, STAGE_N_FINAL, FLAG_SERIAL, FLAG_KNN_SCALAR, FLAG_KNN_TENSOR, DistContext, DistConfig, dist_default_config,
                   DIST_PYRAMID, DIST_BLURRED, DIST_N_FAST, DIST_N_TREE, DIST_PASSES, match_knn_ratio_noalloc_dev, NOALLOC_MAX_DESC)
from .frontend import OrbB200Detector, OrbB200Matcher, DistributedOrbB200Detector, RgbFrame  # noqa: F401
from . import synth  # noqa: F401
