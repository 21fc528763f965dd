"""ctypes binding of libcorsair_b200.so (the C ABI in include/corsair_b200.h).

PyTorch is used only as the device allocator / stream provider: every kernel lives in the
shared library.  There is no CPU fallback: if the library or a CUDA device is missing, calls
raise ``RuntimeError``.
"""
import ctypes
import os
import threading

import numpy as np
import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("CORSAIR_B200_LIB") or os.path.join(_HERE, "lib", "libcorsair_b200.so")  # env: tuning builds

_lib = None
_lib_lock = threading.Lock()

c_void_p, c_int, c_double, c_u64, c_size_t = (ctypes.c_void_p, ctypes.c_int, ctypes.c_double,
                                              ctypes.c_uint64, ctypes.c_size_t)

# name -> (restype, argtypes); mirrors include/corsair_b200.h
SIGNATURES = {
    "cb_version": (c_int, []),
    "cb_create": (c_int, [ctypes.POINTER(c_void_p), c_int, c_size_t]),
    "cb_destroy": (c_int, [c_void_p]),
    "cb_last_error": (ctypes.c_char_p, [c_void_p]),
    "cb_launch_count": (ctypes.c_int64, [c_void_p]),
    "cb_knn_variant": (ctypes.c_char_p, [c_void_p]),
    "cb_set_knn_variant": (c_int, [c_void_p, c_int]),
    "cb_ransac_variant": (ctypes.c_char_p, [c_void_p]),
    "cb_set_ransac_variant": (c_int, [c_void_p, c_int]),
    "cb_ransac_debug_dump": (c_int, [c_void_p, c_void_p, c_int]),
    "cb_knn_last_fallback_rows": (c_int, [c_void_p, c_void_p]),
    "cb_copy_bytes": (c_int, [c_void_p, ctypes.POINTER(ctypes.c_int64), ctypes.POINTER(ctypes.c_int64)]),
    "cb_timing_enable": (c_int, [c_void_p, c_int]),
    "cb_timing_read": (c_int, [c_void_p, c_int, ctypes.POINTER(c_double), ctypes.POINTER(c_double),
                               ctypes.POINTER(ctypes.c_int64)]),
    "cb_timing_reset": (c_int, [c_void_p]),
    "cb_knn": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_int, c_int, c_int, c_int, c_void_p,
                       c_void_p, c_void_p, c_void_p]),
    "cb_knn_ragged": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p,
                              c_int, c_void_p, c_void_p, c_void_p]),
    "cb_pdist": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p]),
    "cb_dot_matrix": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_int, c_int, c_void_p, c_void_p]),
    "cb_retrieval_topk": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_int, c_int, c_int, c_int,
                                  c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "cb_ransac": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_int, c_void_p, c_void_p, c_int,
                          c_void_p, c_u64, c_int, c_int, c_double, c_int, c_double, c_void_p,
                          c_void_p, c_void_p, c_void_p, c_void_p]),
    "cb_draw_minimal_sets": (c_int, [c_void_p, c_u64, c_int, c_int, c_int, c_void_p, c_void_p]),
    "cb_fit_minimal": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_int, c_int, c_void_p,
                               c_void_p]),
    "cb_chamfer": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_int, c_void_p, c_int, c_void_p,
                           c_void_p, c_void_p]),
    "cb_chamfer_static": (c_int, [c_void_p, c_int, c_void_p, c_int, c_void_p, c_int, c_void_p, c_void_p,
                                  ctypes.c_char_p, c_size_t, c_void_p]),  # no context argument
    "cb_chamfer_matrix": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p]),
    "cb_chamfer_matrix_half": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p]),
    "cb_self_hausdorff": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_int, c_void_p, c_void_p]),
    "cb_chamfer_cross": (c_int, [c_void_p, c_void_p, c_int, c_int, c_void_p, c_int, c_int, c_void_p, c_void_p]),
    "cb_nn_max": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_int, c_void_p, c_void_p, c_void_p]),
    "cb_feature_rows_topk": (c_int, [c_void_p, c_void_p, c_int, c_int, c_void_p, c_int, c_int,
                                     c_void_p, c_void_p]),
    "cb_kmeans_batch": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_int, c_int, c_int, c_void_p,
                                c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "cb_assign_labels": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_int, c_void_p, c_void_p,
                                 c_void_p]),
    "cb_symmetry_segments": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_int, c_int, c_double, c_void_p,
                                     c_void_p, c_void_p, c_void_p]),
    "cb_register_pairings": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_int, c_int, c_int,
                                     c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_int, c_double, c_u64,
                                     c_int, c_int, c_double, c_void_p, c_void_p, c_void_p, c_void_p]),
    "cb_np_seeded_choices": (c_int, [ctypes.c_uint32, c_void_p, c_int, c_int, c_void_p]),  # no context argument
    "cb_voxel_unique": (c_int, [c_void_p, c_void_p, c_int, ctypes.c_float, c_void_p, c_void_p, c_void_p, c_void_p]),
    "cb_sym_pose": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_int, c_int, c_int, c_int,
                            c_void_p, c_void_p, c_int, ctypes.c_int64, c_int, c_void_p, c_void_p, c_int, c_int,
                            c_double, c_double, c_u64, c_int, c_int, c_double, c_void_p, c_void_p]),
    "cb_irls_pose": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p]),
    "cb_sym_pose_many": (c_int, [c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p]),
    "cb_ransac_gemm_k": (c_int, []),
}


class SymJob(ctypes.Structure):
    """struct cb_sym_job of include/corsair_b200.h."""
    _fields_ = [("base_f", c_void_p), ("base_xyz", c_void_p), ("pos_f", c_void_p), ("pos_xyz", c_void_p),
                ("n0", ctypes.c_int32), ("n1", ctypes.c_int32), ("d", ctypes.c_int32), ("pos_sym", ctypes.c_int32),
                ("k_nn", ctypes.c_int32), ("n_samples", ctypes.c_int32), ("n_nn", ctypes.c_int32),
                ("n_init", ctypes.c_int32), ("n_trials", ctypes.c_int32), ("h_total", ctypes.c_int32),
                ("adaptive", ctypes.c_int32), ("reserved", ctypes.c_int32), ("np_seed", ctypes.c_int64),
                ("h_samples_base", c_void_p), ("h_samples_pos", c_void_p), ("h_km_first", c_void_p),
                ("h_km_unif", c_void_p), ("min_center", c_double), ("max_dist", c_double), ("confidence", c_double),
                ("seed", c_u64), ("h_out", c_void_p)]


def load_library():
    """dlopen the extension (no GPU needed for this step) and declare every signature."""
    global _lib
    with _lib_lock:
        if _lib is None:
            if not os.path.exists(LIB_PATH):
                raise RuntimeError(
                    f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; "
                    "g.build()'` (there is no CPU fallback)")
            lib = ctypes.CDLL(LIB_PATH)
            for name, (res, args) in SIGNATURES.items():
                fn = getattr(lib, name)
                fn.restype = res
                fn.argtypes = args
            _lib = lib
    return _lib


class CorsairError(RuntimeError):
    pass


_STATUS_EXC = {-1: ValueError, -4: NotImplementedError}


class Context:
    """One cb_ctx (device workspace) bound to a device; use one per host thread."""

    def __init__(self, device=0, workspace_bytes=256 << 20):
        if not torch.cuda.is_available():
            raise RuntimeError("corsair_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
        self.lib = load_library()
        self.device = torch.device("cuda", device)
        torch.cuda.init()
        with torch.cuda.device(self.device):
            torch.zeros(1, device=self.device)  # make sure the primary context exists
            h = c_void_p()
            rc = self.lib.cb_create(ctypes.byref(h), device, workspace_bytes)
        if rc != 0 or not h:
            raise RuntimeError(f"cb_create failed with status {rc}")
        self.handle = h
        self._last_stream = None
        self.h2d_bytes = 0  # counted where the copies are made: dev() / host()
        self.d2h_bytes = 0

    def close(self):
        if getattr(self, "handle", None):
            self.lib.cb_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def stream(self):
        """torch's current stream on this device.  The cb_ctx arena is rewound by every public call and
        is only safe for one stream at a time: when the caller has switched streams since the last call
        (``with torch.cuda.stream(s2):``), the new stream is made to wait for the work the old one still
        has in flight before it may overwrite that scratch."""
        cur = torch.cuda.current_stream(self.device)
        last = self._last_stream
        if last is not None and last.cuda_stream != cur.cuda_stream:
            ev = torch.cuda.Event()
            ev.record(last)
            cur.wait_event(ev)
        self._last_stream = cur
        return c_void_p(cur.cuda_stream)

    def call(self, name, *args):
        rc = getattr(self.lib, name)(self.handle, *args)
        if rc != 0:
            msg = self.lib.cb_last_error(self.handle).decode("utf-8", "replace")
            raise _STATUS_EXC.get(rc, CorsairError)(f"{name}: {msg} (status {rc})")

    def launches(self):
        return int(self.lib.cb_launch_count(self.handle))

    def copy_bytes(self):
        """(host->device, device->host) bytes moved so far: the marshalling of this Context (dev /
        host below) plus the staging copies the C library issued itself (cb_copy_bytes)."""
        a, b = ctypes.c_int64(), ctypes.c_int64()
        self.lib.cb_copy_bytes(self.handle, ctypes.byref(a), ctypes.byref(b))
        return self.h2d_bytes + a.value, self.d2h_bytes + b.value

    def knn_variant(self):
        return self.lib.cb_knn_variant(self.handle).decode()

    def set_knn_variant(self, v):
        self.lib.cb_set_knn_variant(self.handle, int(v))

    def ransac_variant(self):
        return self.lib.cb_ransac_variant(self.handle).decode()

    def ransac_k(self):
        return int(self.lib.cb_ransac_gemm_k())

    def set_ransac_variant(self, v):
        self.lib.cb_set_ransac_variant(self.handle, int(v))

    def ransac_debug_dump(self, tensor=None):
        """Route the raw tensor-core values of the next cb_ransac calls into `tensor` [H, ld] f32."""
        if tensor is None:
            self.lib.cb_ransac_debug_dump(self.handle, None, 0)
        else:
            self.lib.cb_ransac_debug_dump(self.handle, c_void_p(tensor.data_ptr()), int(tensor.shape[1]))

    def knn_last_fallback_rows(self, stream=0):
        return int(self.lib.cb_knn_last_fallback_rows(self.handle, c_void_p(stream)))

    TIMING_CLASSES = ("knn_candidates", "knn_rerank", "ransac_score", "ransac_fit", "chamfer_nn",
                      "rows_select", "kmeans", "ransac_waves")

    def timing(self, on):
        self.lib.cb_timing_enable(self.handle, int(bool(on)))

    def timing_reset(self):
        self.call("cb_timing_reset")

    def timing_read(self):
        """{class: (device ms, algorithmic work, launches)} -- synchronises the device."""
        out = {}
        for i, name in enumerate(self.TIMING_CLASSES):
            ms, work, n = c_double(), c_double(), ctypes.c_int64()
            self.call("cb_timing_read", i, ctypes.byref(ms), ctypes.byref(work), ctypes.byref(n))
            out[name] = (ms.value, work.value, n.value)
        return out

    # -- marshalling helpers --------------------------------------------------------------
    def dev(self, x, dtype):
        """numpy / torch (any device) -> contiguous torch tensor of `dtype` on this device."""
        if isinstance(x, torch.Tensor):
            t = x.detach()
        else:
            t = torch.from_numpy(np.ascontiguousarray(x))
        if not t.is_cuda:
            self.h2d_bytes += t.numel() * t.element_size()
        return t.to(device=self.device, dtype=dtype, non_blocking=False).contiguous()

    def host(self, t):
        """CUDA tensor -> NumPy (the device->host read of a result)."""
        if t.is_cuda:
            self.d2h_bytes += t.numel() * t.element_size()
        return t.cpu().numpy()

    def empty(self, shape, dtype):
        return torch.empty(shape, dtype=dtype, device=self.device)


def ptr(t):
    return c_void_p(t.data_ptr()) if t is not None else c_void_p(0)


_tls = threading.local()


def get_context(device=None):
    """Thread-local context for `device` (default: the current CUDA device)."""
    if device is None:
        device = torch.cuda.current_device() if torch.cuda.is_available() else 0
    if isinstance(device, torch.device):
        device = device.index or 0
    cache = getattr(_tls, "ctx", None)
    if cache is None:
        cache = _tls.ctx = {}
    if device not in cache:
        cache[device] = Context(device)
        with _lib_lock:
            _all_contexts.append(cache[device])
    return cache[device]


_all_contexts = []


def all_contexts():
    """Every thread-local context created so far (bench.py sums their launch counters)."""
    with _lib_lock:
        return list(_all_contexts)
