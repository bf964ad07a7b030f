"""The C++17 plugin adapter (modular-slam_b200/host): boost::dll-style factory aliases, plugin_config.json reader and --
on the GPU -- the reference's own call sequence (detect, detect, match) through the plugin, compared with the oracle."""
import json
import os
import subprocess

import numpy as np
import pytest

import __graft_entry__ as graft

LIB = os.path.join(graft.PKG_DIR, "lib")
PLUGIN = os.path.join(LIB, "libmslam_b200_plugin.so")
DRIVER = os.path.join(LIB, "plugin_host_test")


@pytest.fixture(scope="module")
def built():
    if not (os.path.exists(PLUGIN) and os.path.exists(DRIVER)):
        graft.build()
    return DRIVER


def test_factory_aliases_load_like_boost_dll_import_alias(built):
    out = subprocess.run([built, "symbols", PLUGIN], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    assert "orbDetectorFactory" in out.stdout and "orbMatcherFactory" in out.stdout
    assert "distributedOrbDetectorFactory" in out.stdout and "ransacPnpFactory" in out.stdout
    # the aliases live in section "boostdll", as BOOST_DLL_ALIAS puts them (boost/dll/alias.hpp; test/dummy.cpp:11)
    sec = subprocess.run(["readelf", "-S", "-W", PLUGIN], capture_output=True, text=True).stdout
    assert "boostdll" in sec


def test_plugin_config_reader(built, tmp_path):
    cfg = {"plugins": [{"library": "libmslam_b200_plugin.so",
                        "factoryMethods": {"featureDetector": "orbDetectorFactory", "featureMatcher": "orbMatcherFactory",
                                           "loopDetection": "", "loopClosure": ""},
                        "parameters": {"nfeatures": 2000, "nlevels": 8, "scaleFactor": 1.2, "fastThreshold": 20, "ratio": 0.7}}]}
    p = tmp_path / "plugin_config.json"
    p.write_text(json.dumps(cfg, indent=4))
    out = subprocess.run([built, "config", str(p)], capture_output=True, text=True)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "detector=orbDetectorFactory matcher=orbMatcherFactory nfeatures=2000 nlevels=8 scaleFactor=1.200" in out.stdout
    # the reference's own fixture schema (test/plugin_config.json:1-11) parses too: empty names, no parameters
    p.write_text('{"plugins": [{"library": "", "factoryMethods": {"loopDetection": "", "loopClosure": ""}}]}')
    out = subprocess.run([built, "config", str(p)], capture_output=True, text=True)
    assert out.returncode == 0 and "nfeatures=1000" in out.stdout and "ratio=0.700" in out.stdout
    for bad in ('{"plugins": 3}', '{"plugins": [{"parameters": {"nlevels": 99}}]}', '{"plugins": [', "[]"):
        p.write_text(bad)
        assert subprocess.run([built, "config", str(p)], capture_output=True, text=True).returncode == 2, bad


@pytest.mark.gpu
def test_reference_call_sequence_through_the_plugin(built, oracle, bundled, tmp_path):
    f0, f1 = tmp_path / "f0.raw", tmp_path / "f1.raw"
    bundled["bgr0"].tofile(f0)
    bundled["bgr1"].tofile(f1)
    cfgp = tmp_path / "plugin_config.json"
    cfgp.write_text(json.dumps({"plugins": [{"library": PLUGIN, "parameters": {"maxWidth": 640, "maxHeight": 480}}]}))
    outp = tmp_path / "out.bin"
    env = dict(os.environ, MSLAM_B200_PLUGIN_CONFIG=str(cfgp))
    out = subprocess.run([built, "run", PLUGIN, str(f0), str(f1), "640", "480", str(outp)], capture_output=True, text=True,
                         env=env)
    assert out.returncode == 0, out.stdout + out.stderr
    raw = outp.read_bytes()
    n0, n1, nm, nnone = np.frombuffer(raw[:32], np.uint64).astype(int)
    rec = np.dtype({"names": ["id", "coordinates", "descriptor"], "formats": ["<u8", ("<f8", (2,)), ("u1", (32,))],
                    "offsets": [0, 16, 32], "itemsize": 64})
    k0 = np.frombuffer(raw, rec, n0, 32)
    k1 = np.frombuffer(raw, rec, n1, 32 + 64 * n0)
    m = np.frombuffer(raw, np.dtype([("fromIndex", "<u8"), ("toIndex", "<u8")]), nm, 32 + 64 * (n0 + n1))
    assert nnone == 0
    for k, i in ((k0, 0), (k1, 1)):
        okp, odesc = oracle.orb_detect_rgb(bundled[f"bgr{i}"])
        assert len(k) == len(okp) == 1000
        assert (k["coordinates"][:, 0] == okp["x"].astype(np.float64)).all()
        assert (k["coordinates"][:, 1] == okp["y"].astype(np.float64)).all()
        assert (k["descriptor"] == odesc).all()
    fi, ti, _ = oracle.match_ratio(k0["descriptor"], k1["descriptor"])
    assert nm == len(fi) == 693
    assert (m["fromIndex"].astype(np.int64) == fi).all() and (m["toIndex"].astype(np.int64) == ti).all()


@pytest.mark.gpu
def test_distributed_detector_through_the_plugin(built, oracle, bundled, tmp_path):
    """the Viewer's wiring (viewer.cpp:182-184): DistributedOrb detector + Orb matcher, loaded by alias"""
    f0, f1 = tmp_path / "f0.raw", tmp_path / "f1.raw"
    bundled["bgr0"].tofile(f0)
    bundled["bgr1"].tofile(f1)
    cfgp = tmp_path / "plugin_config.json"
    cfgp.write_text(json.dumps({"plugins": [{"library": PLUGIN,
                                             "factoryMethods": {"featureDetector": "distributedOrbDetectorFactory",
                                                                "featureMatcher": "orbMatcherFactory"},
                                             "parameters": {"maxWidth": 640, "maxHeight": 480, "maxDescriptors": 4096}}]}))
    outp = tmp_path / "out.bin"
    env = dict(os.environ, MSLAM_B200_PLUGIN_CONFIG=str(cfgp))
    out = subprocess.run([built, "run", PLUGIN, str(f0), str(f1), "640", "480", str(outp), "distributedOrbDetectorFactory"],
                         capture_output=True, text=True, env=env)
    assert out.returncode == 0, out.stdout + out.stderr
    raw = outp.read_bytes()
    n0, n1, nm, nnone = np.frombuffer(raw[:32], np.uint64).astype(int)
    rec = np.dtype({"names": ["id", "coordinates", "descriptor"], "formats": ["<u8", ("<f8", (2,)), ("u1", (32,))],
                    "offsets": [0, 16, 32], "itemsize": 64})
    k0 = np.frombuffer(raw, rec, n0, 32)
    k1 = np.frombuffer(raw, rec, n1, 32 + 64 * n0)
    m = np.frombuffer(raw, np.dtype([("fromIndex", "<u8"), ("toIndex", "<u8")]), nm, 32 + 64 * (n0 + n1))
    for k, i in ((k0, 0), (k1, 1)):
        okp, odesc = oracle.dist_detect_rgb(bundled[f"bgr{i}"])
        assert len(k) == len(okp) > 1000
        assert (k["id"] == np.arange(len(k))).all()  # distributed_cv_feature.cpp:1201-1205
        assert (k["coordinates"][:, 0] == okp["x"].astype(np.float64)).all()
        assert (k["coordinates"][:, 1] == okp["y"].astype(np.float64)).all()
        assert (k["descriptor"] == odesc).all()
    fi, ti, _ = oracle.match_ratio(k0["descriptor"], k1["descriptor"])
    assert nm == len(fi) > 100
    assert (m["fromIndex"].astype(np.int64) == fi).all() and (m["toIndex"].astype(np.int64) == ti).all()


@pytest.mark.gpu
def test_ransac_pnp_through_the_plugin(built, tmp_path):
    """OpenCvRansacPnp's replacement through its factory alias, called the way RgbdFeatureFrontend::track does
    (setCameraParameters, solvePnp; rgbd_feature_frontend.cpp:345-361), against the cv2.solvePnPRansac golden"""
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "pnp.npz"))
    synth = graft.load_package().synth
    i = 6  # 693 correspondences, 40 % outliers
    X, x = g[f"obj{i}"].astype(np.float64), g[f"img{i}"].astype(np.float64)
    inp, outp = tmp_path / "pnp_in.bin", tmp_path / "pnp_out.bin"
    with open(inp, "wb") as f:
        f.write(np.int64(len(X)).tobytes())
        f.write(np.array(synth.TUM_K, np.float64).tobytes())
        f.write(X.tobytes())
        f.write(x.tobytes())
    out = subprocess.run([built, "pnp", PLUGIN, str(inp), str(outp)], capture_output=True, text=True)
    assert out.returncode == 0, out.stdout + out.stderr
    raw = outp.read_bytes()
    found, none = np.frombuffer(raw[:16], np.uint64)
    assert found == 1 and none == 0
    pose = np.frombuffer(raw, np.float64, 7, 16)
    inl = np.frombuffer(raw, np.uint8, len(X), 16 + 56)
    assert (inl == g[f"inl{i}"]).all() and f"pnp inliers {inl.sum()} / {len(X)}" in out.stdout
    # the reference's conversion of cv2's (rvec, tvec): orientation = AngleAxis(rvec)^-1, position = -(orientation * tvec)
    from oracle import pnp_oracle as PO
    R = PO.rodrigues_exp(g[f"rvec{i}"])
    qx, qy, qz, qw = pose[3:]
    Rq = np.array([[1 - 2 * (qy * qy + qz * qz), 2 * (qx * qy - qw * qz), 2 * (qx * qz + qw * qy)],
                   [2 * (qx * qy + qw * qz), 1 - 2 * (qx * qx + qz * qz), 2 * (qy * qz - qw * qx)],
                   [2 * (qx * qz - qw * qy), 2 * (qy * qz + qw * qx), 1 - 2 * (qx * qx + qy * qy)]])
    assert np.abs(Rq - R.T).max() < 1e-6 and np.abs(pose[:3] + R.T @ g[f"tvec{i}"]).max() < 1e-6
