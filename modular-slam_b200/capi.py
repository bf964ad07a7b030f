"""ctypes binding of the C ABI in include/mslam_b200.h (the same entry points the C++ plugin adapter calls).

This module is the Python test / bench harness's way into libmslam_b200.so.  There is no CPU fallback:
if the shared library is missing or no CUDA device is present, calls fail loudly.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# MSLAM_B200_LIB: load another build of the SAME library (A/B measurements of compile-time variants); default = the in-tree build
LIB_PATH = os.environ.get("MSLAM_B200_LIB") or os.path.join(_HERE, "lib", "libmslam_b200.so")

OK, E_INVALID, E_CUDA, E_CAPACITY, E_NOMEM = 0, -1, -2, -3, -4
STAGE_PYRAMID, STAGE_BLURRED, STAGE_N_FAST, STAGE_N_CUT1, STAGE_N_FINAL = 0, 1, 2, 3, 4

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
                     ("octave", "<i4")])


class Config(C.Structure):
    _fields_ = [("struct_size", C.c_int), ("device", C.c_int), ("max_width", C.c_int), ("max_height", C.c_int),
                ("max_batch", C.c_int), ("nfeatures", C.c_int), ("nlevels", C.c_int), ("scale_factor", C.c_float),
                ("fast_threshold", C.c_int), ("edge_threshold", C.c_int), ("max_keypoints", C.c_int),
                ("max_descriptors", C.c_int), ("ratio", C.c_double), ("stream", C.c_void_p), ("flags", C.c_int),
                ("reserved", C.c_int)]


FLAG_SERIAL = 1
FLAG_KNN_SCALAR = 2
FLAG_KNN_TENSOR = 4


class Camera(C.Structure):
    """mslam_b200_camera == mslam::CameraParameters (sensors/camera_parameters.hpp:7-12)"""
    _fields_ = [("cx", C.c_double), ("cy", C.c_double), ("fx", C.c_double), ("fy", C.c_double), ("factor", C.c_float),
                ("reserved", C.c_int)]


class PnpParams(C.Structure):
    """mslam_b200_pnp_params: the arguments of cv::solvePnPRansac at cv_ransac_pnp.cpp:56-57 plus the camera matrix"""
    _fields_ = [("struct_size", C.c_int), ("iterations", C.c_int), ("reproj_error", C.c_double), ("confidence", C.c_double),
                ("seed", C.c_uint), ("frame_base", C.c_int), ("refine_iterations", C.c_int), ("reserved", C.c_int),
                ("fx", C.c_double), ("fy", C.c_double), ("cx", C.c_double), ("cy", C.c_double)]


def pnp_params(fx, fy, cx, cy, **kw):
    p = PnpParams()
    library().mslam_b200_pnp_default_params(C.byref(p))
    p.fx, p.fy, p.cx, p.cy = fx, fy, cx, cy
    for k, v in kw.items():
        if not hasattr(p, k):
            raise TypeError(k)
        setattr(p, k, v)
    return p


class MslamB200Error(RuntimeError):
    def __init__(self, status, what, detail=""):
        self.status = status
        super().__init__(f"{what}: status {status} ({_status_string(status)}) {detail}")


_lib = None


def library():
    """Load libmslam_b200.so (built in-tree by __graft_entry__.build()).  Raises if it is missing."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                               "(there is no CPU fallback)")
        lib = C.CDLL(LIB_PATH)
        lib.mslam_b200_status_string.restype = C.c_char_p
        lib.mslam_b200_last_cuda_error.restype = C.c_char_p
        lib.mslam_b200_last_cuda_error.argtypes = [C.c_void_p]
        lib.mslam_b200_launch_count.restype = C.c_longlong
        lib.mslam_b200_launch_count.argtypes = [C.c_void_p]
        lib.mslam_b200_destroy.argtypes = [C.c_void_p]
        lib.mslam_b200_destroy.restype = None
        lib.mslam_b200_default_config.restype = None
        lib.mslam_b200_pnp_default_params.restype = None
        lib.mslam_b200_db_shard_range.restype = None
        _lib = lib
    return _lib


def _status_string(s):
    try:
        return library().mslam_b200_status_string(int(s)).decode()
    except Exception:  # pragma: no cover
        return "?"


def _ptr(a):
    """host numpy array, torch tensor (any device), int address or None -> c_void_p"""
    if a is None:
        return C.c_void_p(0)
    if isinstance(a, np.ndarray):
        return C.c_void_p(a.ctypes.data)
    if isinstance(a, int):
        return C.c_void_p(a)
    if hasattr(a, "data_ptr"):
        return C.c_void_p(a.data_ptr())
    raise TypeError(type(a))


NCCL_ID_BYTES = 128


def db_shard_range(n_keyframes, rank, world):
    """contiguous id range [k0, k1) of `rank` (mslam_b200_db_shard_range): the first n % world ranks hold one keyframe more"""
    if world < 1 or not (0 <= rank < world) or n_keyframes < 0:
        raise ValueError("bad shard request")
    k0, k1 = C.c_longlong(0), C.c_longlong(0)
    library().mslam_b200_db_shard_range(C.c_longlong(n_keyframes), rank, world, C.byref(k0), C.byref(k1))
    return k0.value, k1.value


def nccl_unique_id():
    """ncclGetUniqueId through the library (rank 0 calls it and distributes the bytes) -> bytes[128]"""
    buf = (C.c_char * NCCL_ID_BYTES)()
    rc = library().mslam_b200_nccl_unique_id(buf)
    if rc:
        raise MslamB200Error(rc, "mslam_b200_nccl_unique_id", "(is libnccl.so.2 loadable?)")
    return bytes(buf.raw)


def default_config(**kw):
    cfg = Config()
    library().mslam_b200_default_config(C.byref(cfg))
    for k, v in kw.items():
        if not hasattr(cfg, k):
            raise TypeError(f"unknown config field {k}")
        setattr(cfg, k, v)
    return cfg


NOALLOC_MAX_DESC = 4096


def match_knn_ratio_noalloc_dev(d_from, n_from, d_to, n_to, d_out_from, d_out_to, d_out_dist, d_n_out, ratio=0.7, device=0,
                                stream=None):
    """mslam_b200_match_knn_ratio_noalloc_dev: one (from, to) pair, device in / device out, no context and no allocation (the
    kernels' scratch is a static __device__ array of the library; calls that overlap on one GPU race on it).  Arguments are
    device addresses or torch tensors; asynchronous on `stream` (a CUDA stream handle; None = the legacy default stream)."""
    msg = C.create_string_buffer(512)
    rc = library().mslam_b200_match_knn_ratio_noalloc_dev(C.c_int(device), C.c_void_p(stream or 0), _ptr(d_from), C.c_int(n_from),
                                                          _ptr(d_to), C.c_int(n_to), C.c_double(ratio), _ptr(d_out_from),
                                                          _ptr(d_out_to), _ptr(d_out_dist), _ptr(d_n_out), msg,
                                                          C.c_size_t(len(msg)))
    if rc != OK:
        raise MslamB200Error(rc, "match_knn_ratio_noalloc_dev", msg.value.decode(errors="replace"))


def pinned_empty(shape, dtype):
    """numpy array backed by pinned host memory from mslam_b200_host_alloc (freed with the array)."""
    dtype = np.dtype(dtype)
    n = int(np.prod(shape)) * dtype.itemsize
    p = C.c_void_p()
    rc = library().mslam_b200_host_alloc(C.byref(p), C.c_size_t(max(n, 1)))
    if rc:
        raise MslamB200Error(rc, "mslam_b200_host_alloc")
    buf = (C.c_uint8 * max(n, 1)).from_address(p.value)
    arr = np.frombuffer(buf, dtype=np.uint8, count=n).view(dtype).reshape(shape)
    _PINNED[p.value] = buf
    return arr


_PINNED = {}


class Context:
    """Owns one mslam_b200_ctx (one CUDA device, one stream)."""

    def __init__(self, **cfg_kw):
        self._h = None
        self.cfg = default_config(**cfg_kw)
        h = C.c_void_p()
        rc = library().mslam_b200_create(C.byref(self.cfg), C.byref(h))
        if rc:
            raise MslamB200Error(rc, "mslam_b200_create")
        self._h = h
        self.kp_cap = self.cfg.max_keypoints or (self.cfg.nfeatures + self.cfg.nfeatures // 4 + 64)
        self.desc_cap = self.cfg.max_descriptors or self.kp_cap

    # -- plumbing ---------------------------------------------------------------------------------------------------
    def close(self):
        if self._h is not None:
            library().mslam_b200_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _chk(self, rc, what, allow=()):
        if rc and rc not in allow:
            raise MslamB200Error(rc, what, library().mslam_b200_last_cuda_error(self._h).decode())
        return rc

    def set_stream(self, stream_handle):
        self._chk(library().mslam_b200_set_stream(self._h, C.c_void_p(stream_handle or 0)), "set_stream")

    def set_flags(self, flags):
        self._chk(library().mslam_b200_set_flags(self._h, int(flags)), "set_flags")

    def synchronize(self):
        self._chk(library().mslam_b200_synchronize(self._h), "synchronize")

    @property
    def launch_count(self):
        return int(library().mslam_b200_launch_count(self._h))

    def level_geometry(self, width, height):
        n = self.cfg.nlevels
        w = np.zeros(n, np.int32)
        h = np.zeros(n, np.int32)
        s = np.zeros(n, np.float32)
        k = np.zeros(n, np.int32)
        self._chk(library().mslam_b200_level_geometry(self._h, width, height, _ptr(w), _ptr(h), _ptr(s), _ptr(k)),
                  "level_geometry")
        return w, h, s, k

    # -- detection (host buffers) -------------------------------------------------------------------------------------
    def detect(self, bgr, capacity=None):
        """IFeatureDetector::detect on one frame [H, W, 3] u8 -> (keypoints[KP_DTYPE], descriptors[N, 32])."""
        kp, desc, n = self.detect_batch(bgr[None], capacity)
        return kp[0, :n[0]].copy(), desc[0, :n[0]].copy()

    def detect_batch(self, frames, capacity=None, out=None, allow_capacity_error=False):
        """frames: [B, H, W, 3] u8 array (or list of [H, W, 3]).  -> (kp [B, cap], desc [B, cap, 32], counts [B])."""
        if isinstance(frames, (list, tuple)):
            frames = [np.ascontiguousarray(f, np.uint8) for f in frames]
            B = len(frames)
            H, W, _ = frames[0].shape
            ptrs = (C.c_void_p * B)(*[f.ctypes.data for f in frames])
            stride = W * 3
        else:
            assert frames.dtype == np.uint8 and frames.ndim == 4 and frames.shape[3] == 3
            B, H, W, _ = frames.shape
            assert frames.strides[3] == 1 and frames.strides[2] == 3
            stride = frames.strides[1]
            ptrs = (C.c_void_p * B)(*[frames.ctypes.data + i * frames.strides[0] for i in range(B)])
        cap = capacity or self.kp_cap
        if out is None:
            kp = np.zeros((B, cap), KP_DTYPE)
            desc = np.zeros((B, cap, 32), np.uint8)
            n = np.zeros(B, np.int32)
        else:
            kp, desc, n = out
        rc = library().mslam_b200_detect_batch(self._h, ptrs, B, W, H, stride, _ptr(kp), _ptr(desc), cap, _ptr(n))
        self._chk(rc, "detect_batch", allow=(E_CAPACITY,) if allow_capacity_error else ())
        self.last_status = rc
        return kp, desc, n

    def detect_batch_begin(self, frames, capacity, out):
        """pipelined detect_batch: enqueue one batch (frames [B, H, W, 3] u8, ideally pinned; out = (kp, desc, counts)
        preallocated, ideally pinned) and return at once; up to two batches in flight.  Pair with detect_batch_end()."""
        assert isinstance(frames, np.ndarray) and frames.dtype == np.uint8 and frames.ndim == 4 and frames.shape[3] == 3
        B, H, W, _ = frames.shape
        assert frames.strides[3] == 1 and frames.strides[2] == 3
        ptrs = (C.c_void_p * B)(*[frames.ctypes.data + i * frames.strides[0] for i in range(B)])
        kp, desc, n = out
        rc = library().mslam_b200_detect_batch_begin(self._h, ptrs, B, W, H, frames.strides[1], _ptr(kp), _ptr(desc), capacity,
                                                     _ptr(n))
        self._chk(rc, "detect_batch_begin")
        self._inflight = getattr(self, "_inflight", []) + [(frames, out)]  # keep the buffers alive until the end

    def detect_batch_end(self, allow_capacity_error=False):
        """wait for the oldest batch in flight -> its (kp, desc, counts)"""
        rc = library().mslam_b200_detect_batch_end(self._h)
        self._chk(rc, "detect_batch_end", allow=(E_CAPACITY,) if allow_capacity_error else ())
        self.last_status = rc
        _, out = self._inflight.pop(0)
        return out

    def match_batch_begin(self, mode, frm, n_from, to, n_to, out):
        P, cap, _ = frm.shape
        assert to.shape == frm.shape and frm.flags.c_contiguous and to.flags.c_contiguous
        oa, ob, od, no = out
        self._chk(library().mslam_b200_match_batch_begin(self._h, mode, _ptr(frm), _ptr(n_from), _ptr(to), _ptr(n_to), P, cap,
                                                         _ptr(oa), _ptr(ob), _ptr(od), _ptr(no)), "match_batch_begin")
        self._match_inflight = (frm, n_from, to, n_to, out)

    def match_batch_end(self):
        self._chk(library().mslam_b200_match_batch_end(self._h), "match_batch_end")
        out = self._match_inflight[4]
        self._match_inflight = None
        return out

    def describe(self, bgr, kps):
        """descriptors (and angles) for GIVEN keypoints (x, y at level-0 scale, octave).  -> (kps', desc [N, 32])."""
        bgr = np.ascontiguousarray(bgr, np.uint8)
        H, W, _ = bgr.shape
        kps = np.ascontiguousarray(kps, KP_DTYPE).copy()
        desc = np.zeros((max(len(kps), 1), 32), np.uint8)
        self._chk(library().mslam_b200_describe(self._h, _ptr(bgr), W, H, W * 3, _ptr(kps), len(kps), _ptr(desc)),
                  "describe")
        return kps, desc[:len(kps)]

    def debug_read(self, stage, frame, level, width=None, height=None):
        if stage in (STAGE_PYRAMID, STAGE_BLURRED):
            out = np.zeros((height, width), np.uint8)
        else:
            out = np.zeros(1, np.int32)
        self._chk(library().mslam_b200_debug_read(self._h, stage, frame, level, _ptr(out), C.c_size_t(out.nbytes)),
                  "debug_read")
        return out if out.ndim == 2 else int(out[0])

    # -- detection (device buffers; torch tensors or raw addresses) --------------------------------------------------
    def detect_batch_dev(self, d_bgr, n_frames, width, height, d_kp, d_desc, capacity, d_counts):
        self._chk(library().mslam_b200_detect_batch_dev(self._h, _ptr(d_bgr), n_frames, width, height, _ptr(d_kp),
                                                        _ptr(d_desc), capacity, _ptr(d_counts)), "detect_batch_dev")

    # -- matching -------------------------------------------------------------------------------------------------------
    @staticmethod
    def _desc_arg(d):
        d = np.asarray(d)
        if d.size == 0:
            return np.zeros((0, 32), np.uint8), 0, 32
        assert d.dtype == np.uint8 and d.ndim == 2 and d.shape[1] == 32
        if d.strides[1] != 1:
            d = np.ascontiguousarray(d)
        return d, d.shape[0], d.strides[0]

    def match_knn_ratio(self, frm, to, capacity=None):
        """IFeatureMatcher::match(from, to) -> (fromIndex u64, toIndex u64, dist i32)."""
        f, nf, fs = self._desc_arg(frm)
        t, nt, ts = self._desc_arg(to)
        cap = capacity if capacity is not None else max(nt, 1)
        fi = np.zeros(max(cap, 1), np.uint64)
        ti = np.zeros(max(cap, 1), np.uint64)
        di = np.zeros(max(cap, 1), np.int32)
        n = C.c_int(0)
        self._chk(library().mslam_b200_match_knn_ratio(self._h, _ptr(f), nf, fs, _ptr(t), nt, ts, _ptr(fi), _ptr(ti),
                                                       _ptr(di), cap, C.byref(n)), "match_knn_ratio")
        return fi[:n.value].copy(), ti[:n.value].copy(), di[:n.value].copy()

    def knn2(self, query, train):
        q, nq, qs = self._desc_arg(query)
        t, nt, ts = self._desc_arg(train)
        out = np.zeros((4, max(nq, 1)), np.int32)
        self._chk(library().mslam_b200_knn2(self._h, _ptr(q), nq, qs, _ptr(t), nt, ts, _ptr(out[0]), _ptr(out[1]),
                                            _ptr(out[2]), _ptr(out[3])), "knn2")
        return out[0, :nq], out[1, :nq], out[2, :nq], out[3, :nq]

    def match_crosscheck(self, query, train):
        q, nq, qs = self._desc_arg(query)
        t, nt, ts = self._desc_arg(train)
        cap = max(nq, 1)
        qi = np.zeros(cap, np.int32)
        ti = np.zeros(cap, np.int32)
        di = np.zeros(cap, np.int32)
        n = C.c_int(0)
        self._chk(library().mslam_b200_match_crosscheck(self._h, _ptr(q), nq, qs, _ptr(t), nt, ts, _ptr(qi), _ptr(ti),
                                                        _ptr(di), cap, C.byref(n)), "match_crosscheck")
        return qi[:n.value].copy(), ti[:n.value].copy(), di[:n.value].copy()

    def match_batch(self, mode, frm, n_from, to, n_to, out=None):
        """dense [P, cap, 32] blocks + [P] counts -> (out_a [P, cap], out_b, dist, n_out [P])."""
        frm = np.ascontiguousarray(frm, np.uint8) if not isinstance(frm, np.ndarray) else frm
        to = np.ascontiguousarray(to, np.uint8) if not isinstance(to, np.ndarray) else to
        P, cap, _ = frm.shape
        assert to.shape == frm.shape and frm.flags.c_contiguous and to.flags.c_contiguous
        nf = np.ascontiguousarray(n_from, np.int32)
        nt = np.ascontiguousarray(n_to, np.int32)
        if out is None:
            oa = np.zeros((P, cap), np.int32)
            ob = np.zeros((P, cap), np.int32)
            od = np.zeros((P, cap), np.int32)
            no = np.zeros(P, np.int32)
        else:
            oa, ob, od, no = out
        self._chk(library().mslam_b200_match_batch(self._h, mode, _ptr(frm), _ptr(nf), _ptr(to), _ptr(nt), P, cap,
                                                   _ptr(oa), _ptr(ob), _ptr(od), _ptr(no)), "match_batch")
        return oa, ob, od, no

    def match_batch_dev(self, mode, d_from, d_n_from, d_to, d_n_to, n_pairs, capacity, d_out_a, d_out_b, d_out_dist,
                        d_n_out):
        self._chk(library().mslam_b200_match_batch_dev(self._h, mode, _ptr(d_from), _ptr(d_n_from), _ptr(d_to),
                                                       _ptr(d_n_to), n_pairs, capacity, _ptr(d_out_a), _ptr(d_out_b),
                                                       _ptr(d_out_dist), _ptr(d_n_out)), "match_batch_dev")

    # -- keyframe database ------------------------------------------------------------------------------------------------
    def db_query(self, queries, q_counts, db, k_counts, k, id_base=0, want_scores=False):
        queries = np.ascontiguousarray(queries, np.uint8)
        db = np.ascontiguousarray(db, np.uint8)
        qc = np.ascontiguousarray(q_counts, np.int32)
        kc = np.ascontiguousarray(k_counts, np.int32)
        Q, qcap, _ = queries.shape
        K, kcap, _ = db.shape
        os_ = np.zeros((Q, k), np.int32)
        oi = np.zeros((Q, k), np.int32)
        allsc = np.zeros((Q, K), np.int32) if want_scores else None
        self._chk(library().mslam_b200_db_query(self._h, _ptr(queries), _ptr(qc), Q, qcap, _ptr(db), _ptr(kc), K, kcap,
                                                id_base, k, _ptr(os_), _ptr(oi), _ptr(allsc)), "db_query")
        return (os_, oi, allsc) if want_scores else (os_, oi)

    def db_scores_dev(self, d_queries, d_q_counts, n_q, q_cap, d_db, d_k_counts, n_k, k_cap, d_scores):
        self._chk(library().mslam_b200_db_scores_dev(self._h, _ptr(d_queries), _ptr(d_q_counts), n_q, q_cap, _ptr(d_db),
                                                     _ptr(d_k_counts), n_k, k_cap, _ptr(d_scores)), "db_scores_dev")

    def db_topk_dev(self, d_scores, n_q, n_k, id_base, k, d_out_score, d_out_id):
        self._chk(library().mslam_b200_db_topk_dev(self._h, _ptr(d_scores), n_q, n_k, id_base, k, _ptr(d_out_score),
                                                   _ptr(d_out_id)), "db_topk_dev")

    def db_merge_topk_dev(self, d_scores, d_ids, n_lists, n_q, k, d_out_score, d_out_id):
        self._chk(library().mslam_b200_db_merge_topk_dev(self._h, _ptr(d_scores), _ptr(d_ids), n_lists, n_q, k,
                                                         _ptr(d_out_score), _ptr(d_out_id)), "db_merge_topk_dev")

    # -- sharded keyframe database (one shard per rank; NCCL all-gather + merge inside the library) ---------------------------
    def db_attach(self, rank, world, unique_id=None):
        """join the database communicator (collective): unique_id = the 128 bytes rank 0 got from nccl_unique_id()"""
        buf = (C.c_char * NCCL_ID_BYTES).from_buffer_copy(bytes(unique_id)) if unique_id is not None else None
        self._chk(library().mslam_b200_db_attach(self._h, rank, world, buf), "db_attach")

    def db_attach_comm(self, nccl_comm, rank, world):
        self._chk(library().mslam_b200_db_attach_comm(self._h, C.c_void_p(nccl_comm), rank, world), "db_attach_comm")

    def db_detach(self):
        self._chk(library().mslam_b200_db_detach(self._h), "db_detach")

    def db_shard_info(self):
        r, w = C.c_int(0), C.c_int(0)
        self._chk(library().mslam_b200_db_shard_info(self._h, C.byref(r), C.byref(w)), "db_shard_info")
        return r.value, w.value

    def db_query_sharded_dev(self, d_queries, d_q_counts, n_q, q_cap, d_db, d_k_counts, n_local, k_cap, id_base, k, d_out_score,
                             d_out_id):
        self._chk(library().mslam_b200_db_query_sharded_dev(self._h, _ptr(d_queries), _ptr(d_q_counts), n_q, q_cap, _ptr(d_db),
                                                            _ptr(d_k_counts), n_local, k_cap, id_base, k, _ptr(d_out_score),
                                                            _ptr(d_out_id)), "db_query_sharded_dev")

    def db_query_sharded(self, queries, q_counts, d_db, d_k_counts, n_local, k_cap, id_base, k):
        """host query frames [Q, cap, 32] -> host (scores [Q, k], ids [Q, k]); the shard arguments are device pointers"""
        queries = np.ascontiguousarray(queries, np.uint8)
        qc = np.ascontiguousarray(q_counts, np.int32)
        Q, qcap, _ = queries.shape
        os_ = np.zeros((Q, k), np.int32)
        oi = np.zeros((Q, k), np.int32)
        self._chk(library().mslam_b200_db_query_sharded(self._h, _ptr(queries), _ptr(qc), Q, qcap, _ptr(d_db), _ptr(d_k_counts),
                                                        n_local, k_cap, id_base, k, _ptr(os_), _ptr(oi)), "db_query_sharded")
        return os_, oi

    # -- RGB-D association -----------------------------------------------------------------------------
    def backproject(self, xy, depth, camera):
        """pointsFromRgbdKeypoints: xy [n, 2] float64 rows (any row stride that is a multiple of 8), depth [H, W] u16
        -> (xyz [n, 3] f64, valid [n] u8)"""
        xy = np.asarray(xy, np.float64)
        assert xy.ndim == 2 and xy.shape[1] == 2 and (xy.shape[0] == 0 or xy.strides[1] == 8)
        depth = np.ascontiguousarray(depth, np.uint16)
        n = xy.shape[0]
        xyz = np.zeros((n, 3), np.float64)
        valid = np.zeros(n, np.uint8)
        self._chk(library().mslam_b200_backproject(self._h, _ptr(xy) if n else None, n, xy.strides[0] if n else 16, _ptr(depth),
                                                   depth.shape[1], depth.shape[0], C.byref(camera), _ptr(xyz), _ptr(valid)),
                  "backproject")
        return xyz, valid

    def backproject_dev(self, d_kp, kp_cap, d_index, d_counts, n_pairs, cap, d_depth, width, height, camera, d_xyz, d_valid):
        self._chk(library().mslam_b200_backproject_dev(self._h, _ptr(d_kp), kp_cap, _ptr(d_index), _ptr(d_counts), n_pairs,
                                                       cap, _ptr(d_depth), width, height, C.byref(camera), _ptr(d_xyz),
                                                       _ptr(d_valid)), "backproject_dev")

    # -- RANSAC PnP ------------------------------------------------------------------------------------
    def pnp_ransac_batch(self, obj, img, counts, params):
        """OpenCvRansacPnp::solvePnp for a batch of frames: obj [F, cap, 3] f32, img [F, cap, 2] f32, counts [F]
        -> (rvec_tvec [F, 6] f64, pose [F, 7] f64 (qw qx qy qz px py pz), inliers [F, cap] u8, info [F, 4] i32 =
        found, n_inliers, winning hypothesis, LM iterations)"""
        obj = np.ascontiguousarray(obj, np.float32)
        img = np.ascontiguousarray(img, np.float32)
        counts = np.ascontiguousarray(counts, np.int32)
        F, cap, _ = obj.shape
        assert img.shape == (F, cap, 2) and counts.shape == (F,)
        rt = np.zeros((F, 6), np.float64)
        pose = np.zeros((F, 7), np.float64)
        inl = np.zeros((F, cap), np.uint8)
        info = np.zeros((F, 4), np.int32)
        self._chk(library().mslam_b200_pnp_ransac_batch(self._h, _ptr(obj), _ptr(img), _ptr(counts), F, cap, C.byref(params),
                                                        _ptr(rt), _ptr(pose), _ptr(inl), _ptr(info)), "pnp_ransac_batch")
        return rt, pose, inl, info

    def pnp_ransac_batch_dev(self, d_obj, d_img, d_counts, n_frames, cap, params, d_rt, d_pose, d_inliers, d_info):
        self._chk(library().mslam_b200_pnp_ransac_batch_dev(self._h, _ptr(d_obj), _ptr(d_img), _ptr(d_counts), n_frames, cap,
                                                            C.byref(params), _ptr(d_rt), _ptr(d_pose), _ptr(d_inliers),
                                                            _ptr(d_info)), "pnp_ransac_batch_dev")

    # -- measurement ------------------------------------------------------------------------------------------------------
    STAGES = ("gray", "resize", "fast", "select", "blur", "describe", "knn", "compact", "db")

    def enable_stage_timing(self, on=True):
        self._chk(library().mslam_b200_enable_stage_timing(self._h, int(bool(on))), "enable_stage_timing")

    def stage_times(self):
        """-> {stage: (total_ms, intervals, kernel launches)} since the last call"""
        n = len(self.STAGES)
        ms = np.zeros(n, np.float64)
        iv = np.zeros(n, np.int64)
        ln = np.zeros(n, np.int64)
        self._chk(library().mslam_b200_stage_times(self._h, _ptr(ms), _ptr(iv), _ptr(ln)), "stage_times")
        return {s: (float(ms[i]), int(iv[i]), int(ln[i])) for i, s in enumerate(self.STAGES)}

    def measure_int_peaks(self):
        out = np.zeros(8, np.float64)
        self._chk(library().mslam_b200_measure_int_peaks(self._h, _ptr(out)), "measure_int_peaks")
        return {"popc32_per_s": float(out[0]), "lop3_per_s": float(out[1]), "plain_pairs_per_s": float(out[2]),
                "sm_mhz": float(out[3]), "csa_pairs_per_s": float(out[4]), "lop3_3reg_per_s": float(out[5])}


    def measure_tensor_peak(self):
        out = np.zeros(4, np.float64)
        self._chk(library().mslam_b200_measure_tensor_peak(self._h, _ptr(out)), "measure_tensor_peak")
        return {"int8_ops_per_s_n128_sustained": float(out[0]), "int8_ops_per_s_n256_sustained": float(out[1]),
                "int8_ops_per_s_n128_burst": float(out[2]), "int8_ops_per_s_n256_burst": float(out[3])}


# ---- second detector: DistributedOrbOpenCvDetector (include/mslam_b200.h, "second detector") ---------------------------
DIST_PYRAMID, DIST_BLURRED, DIST_N_FAST, DIST_N_TREE, DIST_PASSES = 0, 1, 2, 3, 4


class DistConfig(C.Structure):
    _fields_ = [("struct_size", C.c_int), ("device", C.c_int), ("max_width", C.c_int), ("max_height", C.c_int),
                ("max_batch", C.c_int), ("nlevels", C.c_int), ("scale_factor", C.c_float),
                ("ini_fast_threshold", C.c_int), ("min_fast_threshold", C.c_int), ("min_size", C.c_int),
                ("max_keypoints", C.c_int), ("stream", C.c_void_p), ("flags", C.c_int), ("reserved", C.c_int)]


def dist_default_config(**kw):
    cfg = DistConfig()
    lib = library()
    lib.mslam_b200_dist_default_config.restype = None
    lib.mslam_b200_dist_default_config(C.byref(cfg))
    for k, v in kw.items():
        if not hasattr(cfg, k):
            raise TypeError(f"unknown config field {k}")
        setattr(cfg, k, v)
    return cfg


class DistContext:
    """Owns one mslam_b200_dist handle (DistributedOrbOpenCvDetector on one CUDA device / stream)."""

    def __init__(self, **cfg_kw):
        self._h = None
        self.cfg = dist_default_config(**cfg_kw)
        lib = library()
        lib.mslam_b200_dist_destroy.argtypes = [C.c_void_p]
        lib.mslam_b200_dist_destroy.restype = None
        lib.mslam_b200_dist_last_cuda_error.restype = C.c_char_p
        lib.mslam_b200_dist_last_cuda_error.argtypes = [C.c_void_p]
        lib.mslam_b200_dist_launch_count.restype = C.c_longlong
        lib.mslam_b200_dist_launch_count.argtypes = [C.c_void_p]
        h = C.c_void_p()
        rc = lib.mslam_b200_dist_create(C.byref(self.cfg), C.byref(h))
        if rc:
            raise MslamB200Error(rc, "mslam_b200_dist_create")
        self._h = h
        self.kp_cap = lib.mslam_b200_dist_keypoint_capacity(h)

    def close(self):
        if self._h is not None:
            library().mslam_b200_dist_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _check(self, rc, what):
        if rc:
            raise MslamB200Error(rc, what, library().mslam_b200_dist_last_cuda_error(self._h).decode())

    @property
    def launch_count(self):
        return int(library().mslam_b200_dist_launch_count(self._h))

    def synchronize(self):
        self._check(library().mslam_b200_dist_synchronize(self._h), "mslam_b200_dist_synchronize")

    def level_geometry(self, W, H):
        n = self.cfg.nlevels
        w, h, s = np.zeros(n, np.int32), np.zeros(n, np.int32), np.zeros(n, np.float32)
        self._check(library().mslam_b200_dist_level_geometry(self._h, W, H, _ptr(w), _ptr(h), _ptr(s)),
                    "mslam_b200_dist_level_geometry")
        return w, h, s

    def detect(self, bgr, capacity=None):
        """one frame [H, W, 3] u8 -> (keypoints[KP_DTYPE], descriptors [n, 32]) in the reference's order"""
        bgr = np.ascontiguousarray(bgr, np.uint8)
        H, W, _ = bgr.shape
        cap = capacity or self.kp_cap
        kp = np.zeros(cap, KP_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        n = C.c_int(0)
        rc = library().mslam_b200_dist_detect(self._h, _ptr(bgr), W, H, W * 3, _ptr(kp), _ptr(desc), cap, C.byref(n))
        self._check(rc, "mslam_b200_dist_detect")
        return kp[:n.value].copy(), desc[:n.value].copy()

    def detect_batch(self, frames, capacity=None, out=None):
        """frames [B, H, W, 3] u8 (host) -> (kp [B, cap], desc [B, cap, 32], counts [B]); out = preallocated triple"""
        if not (isinstance(frames, np.ndarray) and frames.dtype == np.uint8 and frames.flags.c_contiguous):
            frames = np.ascontiguousarray(frames, np.uint8)
        B, H, W, _ = frames.shape
        cap = capacity or self.kp_cap
        if out is None:
            kp = np.zeros((B, cap), KP_DTYPE)
            desc = np.zeros((B, cap, 32), np.uint8)
            counts = np.zeros(B, np.int32)
        else:
            kp, desc, counts = out
            assert kp.shape == (B, cap) and desc.shape == (B, cap, 32) and counts.shape == (B,)
        base, fs = frames.ctypes.data, frames.strides[0]
        ptrs = (C.c_void_p * B)(*[base + i * fs for i in range(B)])
        rc = library().mslam_b200_dist_detect_batch(self._h, ptrs, B, W, H, W * 3, _ptr(kp), _ptr(desc), cap, _ptr(counts))
        self._check(rc, "mslam_b200_dist_detect_batch")
        return kp, desc, counts

    def detect_batch_begin(self, frames, capacity, out):
        """pipelined detect_batch (two batches in flight); frames / out should be pinned.  Pair with detect_batch_end()."""
        assert isinstance(frames, np.ndarray) and frames.dtype == np.uint8 and frames.flags.c_contiguous
        B, H, W, _ = frames.shape
        kp, desc, counts = out
        base, fs = frames.ctypes.data, frames.strides[0]
        ptrs = (C.c_void_p * B)(*[base + i * fs for i in range(B)])
        rc = library().mslam_b200_dist_detect_batch_begin(self._h, ptrs, B, W, H, W * 3, _ptr(kp), _ptr(desc), capacity, _ptr(counts))
        self._check(rc, "mslam_b200_dist_detect_batch_begin")
        self._inflight = getattr(self, "_inflight", []) + [(frames, out)]

    def detect_batch_end(self):
        self._check(library().mslam_b200_dist_detect_batch_end(self._h), "mslam_b200_dist_detect_batch_end")
        return self._inflight.pop(0)[1]

    def detect_batch_dev(self, d_bgr, n_frames, W, H, d_kp, d_desc, capacity, d_counts):
        """device tensors in / out, asynchronous on the handle's stream"""
        rc = library().mslam_b200_dist_detect_batch_dev(self._h, _ptr(d_bgr), n_frames, W, H, _ptr(d_kp), _ptr(d_desc),
                                                        capacity, _ptr(d_counts))
        self._check(rc, "mslam_b200_dist_detect_batch_dev")

    def debug_image(self, stage, frame, level, W, H):
        w, h, _ = self.level_geometry(W, H)
        out = np.zeros((h[level], w[level]), np.uint8)
        self._check(library().mslam_b200_dist_debug_read(self._h, stage, frame, level, _ptr(out), C.c_size_t(out.size)),
                    "mslam_b200_dist_debug_read")
        return out

    def debug_count(self, stage, frame, level):
        v = np.zeros(1, np.int32)
        self._check(library().mslam_b200_dist_debug_read(self._h, stage, frame, level, _ptr(v), C.c_size_t(4)),
                    "mslam_b200_dist_debug_read")
        return int(v[0])

    def stage_times(self, enable=True):
        ms = np.zeros(6, np.float64)
        self._check(library().mslam_b200_dist_stage_times(self._h, int(enable), _ptr(ms)), "mslam_b200_dist_stage_times")
        return dict(zip(("gray", "resize", "fast", "tree", "blur", "describe"), ms.tolist()))
