"""Python mirror of the reference's feature interfaces, over the C ABI -- same names, argument meaning and error
behaviour as the C++ plugin (host/orb_b200_plugin.cpp), so that tests read like the reference's:

    IFeatureDetector<RgbFrame, uint8_t, 32>::detect(const RgbFrame&)            feature_interface.hpp:50-56
    IFeatureMatcher<uint8_t, 32>::match(firstDescriptors, secondDescriptors)     feature_interface.hpp:62-70
    (reference implementations: OrbOpenCvDetector / OrbOpenCvMatcher, src/lib/modular_slam/orb_feature.cpp)
"""
import numpy as np

from .capi import Context

# KeypointDescriptor<uint8_t, 32> as laid out by the reference on x86-64 (feature_interface.hpp:18-29):
# Keypoint{Id id (8) + pad (8) + Eigen::Vector2d coordinates (16)} + std::array<uint8_t, 32>  = 64 bytes
ORB_KEYPOINT_DTYPE = np.dtype({"names": ["id", "coordinates", "descriptor"],
                               "formats": ["<u8", ("<f8", (2,)), ("u1", (32,))],
                               "offsets": [0, 16, 32], "itemsize": 64})
DESCRIPTOR_MATCH_DTYPE = np.dtype([("fromIndex", "<u8"), ("toIndex", "<u8")])


class RgbFrame:
    """mslam::RgbFrame {std::vector<uint8_t> data; Size size;} (types/rgb_frame.hpp:12-16)"""

    def __init__(self, data, width, height):
        self.data = np.ascontiguousarray(data, np.uint8).reshape(-1)
        self.width, self.height = int(width), int(height)
        assert self.data.size == self.width * self.height * 3


class OrbB200Detector:
    """drop-in for mslam::OrbOpenCvDetector (orb_feature.cpp:19-72)"""

    def __init__(self, ctx=None, **cfg):
        self.ctx = ctx or Context(**cfg)

    def detect(self, frame: RgbFrame):
        kp, desc = self.ctx.detect(frame.data.reshape(frame.height, frame.width, 3))
        out = np.zeros(len(kp), ORB_KEYPOINT_DTYPE)
        out["coordinates"][:, 0] = kp["x"]  # float -> double, as orb_feature.cpp:57
        out["coordinates"][:, 1] = kp["y"]
        out["descriptor"] = desc
        return out


class OrbB200Matcher:
    """drop-in for mslam::OrbOpenCvMatcher (orb_feature.cpp:74-117)"""

    def __init__(self, ctx=None, **cfg):
        self.ctx = ctx or Context(**cfg)

    def match(self, first, second):
        # the descriptor view keeps the 64-byte record stride, exactly like orb_feature.cpp:88-91
        fi, ti, _ = self.ctx.match_knn_ratio(first["descriptor"], second["descriptor"])
        out = np.zeros(len(fi), DESCRIPTOR_MATCH_DTYPE)
        out["fromIndex"], out["toIndex"] = fi, ti
        return out


class DistributedOrbB200Detector:
    """drop-in for mslam::DistributedOrbOpenCvDetector (distributed_cv_feature.cpp:1181-1222): Keypoint::id = index"""

    def __init__(self, ctx=None, **cfg):
        from .capi import DistContext
        self.ctx = ctx or DistContext(**cfg)

    def detect(self, frame: RgbFrame):
        kp, desc = self.ctx.detect(frame.data.reshape(frame.height, frame.width, 3))
        out = np.zeros(len(kp), ORB_KEYPOINT_DTYPE)
        out["id"] = np.arange(len(kp), dtype=np.uint64)  # distributed_cv_feature.cpp:1201-1205
        out["coordinates"][:, 0] = kp["x"]
        out["coordinates"][:, 1] = kp["y"]
        out["descriptor"] = desc
        return out
