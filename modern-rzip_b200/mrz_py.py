"""ctypes view of the C-ABI in include/mrz_gpu.h (used by tests/, bench.py and __graft_entry__.py).

The product's host side is C (host/); this module only lets the Python harness call the same
entry points.  It never falls back to a CPU implementation: if libmrz_gpu.so is missing or no CUDA
device is present, calls raise.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libmrz_gpu.so")

DECLARED_SYMBOLS = [
    "mrz_device_count", "mrz_version", "mrz_last_error", "mrz_ctx_create", "mrz_ctx_destroy",
    "mrz_rzip_chunk", "mrz_upload", "mrz_run_resident", "mrz_download", "mrz_host_alloc", "mrz_host_free",
    "mrz_debug_tags", "mrz_debug_crc32", "mrz_debug_tokens",
    "mrz_ctx_create_variant", "mrz_run_automaton", "mrz_run_emit", "mrz_run_variants", "mrz_run_many",
    "mrz_config_default", "mrz_ctx_create_ex", "mrz_upload_part", "mrz_upload_wait", "mrz_upload_finish",
    "mrz_download_part", "mrz_get_profile", "mrz_debug_tagmap", "mrz_ctx_set_cluster", "mrz_crc32_noalloc", "mrz_noalloc_last_error", "mrz_runzip_chunk",
]

LEVELS = [(1, 4, 1), (2, 4, 2), (4, 4, 2), (8, 4, 2), (16, 4, 3), (32, 4, 4), (32, 2, 6), (64, 1, 16), (64, 1, 32), (64, 1, 128)]
CHAIN = [c for _, _, c in LEVELS]     # levels[].max_chain_len, rzip.c:65-73


def hash_index(skip_files=0):
    """st->hash_index as the host computes it (host/rzip_fd.c init_hash_indexes, rzip.c:669-673): 256 entries of
    (random() << 16) ^ random() from glibc's unseeded random() (== srandom(1)); file k of one command line uses draws
    512k .. 512k+511 (main.c:651).  Restated here so that the harness does not need the oracle for its parameters."""
    libc = C.CDLL(None)
    libc.random.restype = C.c_long
    libc.srandom(1)
    for _ in range(512 * skip_files):
        libc.random()
    out = np.empty(256, dtype=np.int64)
    for i in range(256):
        a = libc.random()
        b = libc.random()
        out[i] = (a << 16) ^ b
    return out


def chunk_bytes(n):
    """st->chunk_bytes (rzip.c:1006-1008): bytes needed for a distance inside a chunk of n bytes, at least 1."""
    bits = 8
    while n >> bits > 0:
        bits += 1
    return bits // 8 + (1 if bits % 8 else 0)


class MrzConfig(C.Structure):
    _fields_ = [("cluster_size", C.c_int32), ("threads", C.c_int32), ("filter_log2_words", C.c_int32),
                ("profile", C.c_int32), ("token_divisor", C.c_int32), ("poison", C.c_int32), ("l2_policy", C.c_int32), ("reserved", C.c_int32 * 1)]


class MrzUnzipResult(C.Structure):
    _fields_ = [("out_len", C.c_int64), ("literals", C.c_int64), ("matches", C.c_int64), ("waves", C.c_int64),
                ("crc32", C.c_uint32), ("crc32_stored", C.c_uint32), ("ms_kernels", C.c_float), ("gpu_launches", C.c_int32)]


class MrzParams(C.Structure):
    _fields_ = [("level", C.c_int32), ("chunk_bytes", C.c_int32), ("victim_round_in", C.c_int64),
                ("hash_index", C.c_int64 * 256)]


class MrzResult(C.Structure):
    _fields_ = [(n, C.c_int64) for n in (
        "s0_len", "s1_len", "matches", "match_bytes", "literals", "literal_bytes", "tag_hits", "tag_misses",
        "inserts", "victim_round_out", "victim_events", "final_min_mask", "n_ops")] + [
        ("crc32", C.c_uint32), ("reserved", C.c_uint32),
        ("ms_kernels", C.c_double), ("ms_tagscan", C.c_double), ("ms_automaton", C.c_double),
        ("ms_emit", C.c_double), ("gpu_launches", C.c_int64)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


class MrzError(RuntimeError):
    pass


_lib = None


def bind(path):
    """Load a library exporting include/mrz_gpu.h and set the ctypes prototypes.  The product library is
    LIB_PATH; tests/cusim builds an emulator library with the same ABI for CPU-only kernel tests."""
    L = C.CDLL(path)
    if True:
        L.mrz_version.restype = C.c_char_p
        L.mrz_last_error.restype = C.c_char_p
        L.mrz_last_error.argtypes = [C.c_void_p]
        L.mrz_ctx_create.argtypes = [C.c_int, C.c_int64, C.POINTER(C.c_void_p)]
        L.mrz_ctx_destroy.argtypes = [C.c_void_p]
        L.mrz_rzip_chunk.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.POINTER(MrzParams), C.c_void_p, C.c_int64,
                                     C.c_void_p, C.c_int64, C.POINTER(MrzResult)]
        L.mrz_upload.argtypes = [C.c_void_p, C.c_void_p, C.c_int64]
        L.mrz_run_resident.argtypes = [C.c_void_p, C.c_int64, C.POINTER(MrzParams), C.POINTER(MrzResult)]
        L.mrz_download.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64]
        L.mrz_host_alloc.restype = C.c_void_p
        L.mrz_host_alloc.argtypes = [C.c_int64]
        L.mrz_host_free.argtypes = [C.c_void_p]
        L.mrz_debug_tags.argtypes = [C.c_void_p, C.c_int64, C.POINTER(C.c_int64), C.c_void_p, C.c_int64]
        L.mrz_debug_crc32.argtypes = [C.c_void_p, C.c_int64, C.POINTER(C.c_uint32)]
        L.mrz_debug_tokens.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.POINTER(C.c_int64)]
        L.mrz_ctx_create_variant.argtypes = [C.c_void_p, C.POINTER(C.c_void_p)]
        L.mrz_run_automaton.argtypes = [C.c_void_p, C.c_int64, C.POINTER(MrzParams), C.POINTER(MrzResult)]
        L.mrz_run_emit.argtypes = [C.c_void_p, C.POINTER(MrzResult)]
        L.mrz_run_variants.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.c_int64, C.POINTER(MrzParams),
                                       C.POINTER(MrzResult)]
        L.mrz_run_many.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.POINTER(C.c_int64), C.POINTER(MrzParams),
                                   C.POINTER(MrzResult)]
        L.mrz_config_default.argtypes = [C.POINTER(MrzConfig)]
        L.mrz_ctx_create_ex.argtypes = [C.c_int, C.c_int64, C.POINTER(MrzConfig), C.POINTER(C.c_void_p)]
        L.mrz_ctx_set_cluster.argtypes = [C.c_void_p, C.c_int]
        L.mrz_crc32_noalloc.argtypes = [C.c_int, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]
        L.mrz_noalloc_last_error.restype = C.c_char_p
        L.mrz_upload_part.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_int64]
        L.mrz_upload_wait.argtypes = [C.c_void_p]
        L.mrz_upload_finish.argtypes = [C.c_void_p, C.c_int64]
        L.mrz_runzip_chunk.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_int, C.c_void_p, C.c_int64,
                                       C.POINTER(MrzUnzipResult)]
        L.mrz_download_part.argtypes = [C.c_void_p, C.c_int, C.c_int64, C.c_void_p, C.c_int64]
        L.mrz_get_profile.argtypes = [C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_int32)]
        L.mrz_debug_tagmap.argtypes = [C.c_void_p, C.c_int64, C.POINTER(C.c_int64), C.c_void_p, C.c_void_p]
    return L


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise MrzError("libmrz_gpu.so not built (run `make -C modern-rzip_b200`); there is no CPU fallback")
        _lib = bind(LIB_PATH)
    return _lib


def make_params(level, chunk_bytes, hash_index, victim_round_in=0):
    p = MrzParams()
    p.level = level
    p.chunk_bytes = chunk_bytes
    p.victim_round_in = victim_round_in
    for i in range(256):
        p.hash_index[i] = int(hash_index[i])
    return p


class PinnedBuffer:
    """cudaHostAlloc'd byte buffer exposed as a numpy array."""

    def __init__(self, nbytes):
        self.ptr = lib().mrz_host_alloc(max(1, nbytes))
        if not self.ptr:
            raise MrzError("mrz_host_alloc failed")
        self.nbytes = nbytes
        self.array = np.ctypeslib.as_array(C.cast(self.ptr, C.POINTER(C.c_uint8)), shape=(max(1, nbytes),))

    def free(self):
        if self.ptr:
            lib().mrz_host_free(self.ptr)
            self.ptr = None


class Context:
    def __init__(self, device=0, max_chunk=1 << 28, library=None, parent=None, **config):
        """parent: create a variant context (mrz_ctx_create_variant) sharing the parent's chunk bytes and tag map.
        config: fields of mrz_config (cluster_size, threads, filter_log2_words, profile, token_divisor, poison)."""
        self.h = C.c_void_p()
        if parent is not None:
            self.L = parent.L
            rc = self.L.mrz_ctx_create_variant(parent.h, C.byref(self.h))
            max_chunk = parent.max_chunk
        else:
            self.L = library if library is not None else lib()
            cfg = MrzConfig()
            self.L.mrz_config_default(C.byref(cfg))
            for k, v in config.items():
                setattr(cfg, k, v)
            rc = self.L.mrz_ctx_create_ex(device, max_chunk, C.byref(cfg), C.byref(self.h))
        if rc:
            raise MrzError("mrz_ctx_create rc=%d: %s" % (rc, self.L.mrz_last_error(None).decode()))
        self.max_chunk = max_chunk

    def close(self):
        if self.h:
            self.L.mrz_ctx_destroy(self.h)
            self.h = C.c_void_p()

    def _check(self, rc, what):
        if rc:
            raise MrzError("%s rc=%d: %s" % (what, rc, self.L.mrz_last_error(self.h).decode()))

    def upload(self, arr):
        arr = np.ascontiguousarray(arr, dtype=np.uint8)
        self._check(self.L.mrz_upload(self.h, arr.ctypes.data if arr.size else None, arr.size), "mrz_upload")

    def runzip_chunk(self, s0, s1, chunk_bytes, out_cap):
        """mrz_runzip_chunk: the two streams of a chunk -> (MrzUnzipResult, the chunk's bytes).  Raises MrzError on a corrupt
        stream or a checksum mismatch."""
        s0 = np.ascontiguousarray(s0, dtype=np.uint8)
        s1 = np.ascontiguousarray(s1, dtype=np.uint8)
        out = np.empty(max(1, out_cap), dtype=np.uint8)
        res = MrzUnzipResult()
        self._check(self.L.mrz_runzip_chunk(self.h, s0.ctypes.data, s0.size, s1.ctypes.data if s1.size else None, s1.size, chunk_bytes,
                                            out.ctypes.data, out_cap, C.byref(res)), "mrz_runzip_chunk")
        return res, out[:res.out_len]

    def invalidate_map(self, n):
        """The resident chunk counts as uploaded anew: the next run recomputes its tag map and CRC (mrz_upload_finish)."""
        self._check(self.L.mrz_upload_finish(self.h, n), "mrz_upload_finish")

    def run_resident(self, n, params):
        res = MrzResult()
        self._check(self.L.mrz_run_resident(self.h, n, C.byref(params), C.byref(res)), "mrz_run_resident")
        return res

    def run_automaton(self, n, params):
        res = MrzResult()
        self._check(self.L.mrz_run_automaton(self.h, n, C.byref(params), C.byref(res)), "mrz_run_automaton")
        return res

    def run_emit(self):
        res = MrzResult()
        self._check(self.L.mrz_run_emit(self.h, C.byref(res)), "mrz_run_emit")
        return res

    def download(self, res, s0=None, s1=None):
        s0 = np.empty(max(1, res.s0_len), dtype=np.uint8) if s0 is None else s0
        s1 = np.empty(max(1, res.s1_len), dtype=np.uint8) if s1 is None else s1
        self._check(self.L.mrz_download(self.h, s0.ctypes.data, s0.size, s1.ctypes.data, s1.size), "mrz_download")
        return s0[:res.s0_len], s1[:res.s1_len]

    def rzip_chunk(self, arr, params, s0=None, s1=None):
        """hash_search for one chunk through the host-buffer entry point (H2D + kernels + D2H)."""
        arr = np.ascontiguousarray(arr, dtype=np.uint8)
        n = arr.size
        s0 = np.empty(n // 2 + 64, dtype=np.uint8) if s0 is None else s0
        s1 = np.empty(n + 16, dtype=np.uint8) if s1 is None else s1
        res = MrzResult()
        rc = self.L.mrz_rzip_chunk(self.h, arr.ctypes.data if n else None, n, C.byref(params), s0.ctypes.data, s0.size,
                                  s1.ctypes.data, s1.size, C.byref(res))
        self._check(rc, "mrz_rzip_chunk")
        return res, s0[:res.s0_len], s1[:res.s1_len]

    def debug_tags(self, n, hash_index, count):
        hi = (C.c_int64 * 256)(*[int(x) for x in hash_index])
        out = np.empty(count, dtype=np.int64)
        self._check(self.L.mrz_debug_tags(self.h, n, hi, out.ctypes.data, count), "mrz_debug_tags")
        return out

    def debug_tagmap(self, n, hash_index):
        hi = (C.c_int64 * 256)(*[int(x) for x in hash_index])
        bmap = np.empty(max(1, n), dtype=np.uint8)
        bsum = np.empty(max(1, (n + 63) // 64), dtype=np.uint8)
        self._check(self.L.mrz_debug_tagmap(self.h, n, hi, bmap.ctypes.data, bsum.ctypes.data), "mrz_debug_tagmap")
        return bmap[:n], bsum[:(n + 63) // 64]

    def profile(self):
        """-> (64 phase counters of the last automaton run (profile=1), (CTAs per cluster, threads, automata))"""
        cnt = (C.c_int64 * 64)()
        geom = (C.c_int32 * 3)()
        self._check(self.L.mrz_get_profile(self.h, cnt, geom), "mrz_get_profile")
        return list(cnt), tuple(geom)

    def debug_tokens(self):
        cnt = C.c_int64()
        self.L.mrz_debug_tokens(self.h, None, 0, C.byref(cnt))
        out = np.empty((max(1, cnt.value), 3), dtype=np.int64)
        self._check(self.L.mrz_debug_tokens(self.h, out.ctypes.data, cnt.value, C.byref(cnt)), "mrz_debug_tokens")
        return out[:cnt.value]


class VariantSet:
    """One resident chunk, `nvar` automata side by side: variant v runs with victim_round_in = v (include/mrz_gpu.h,
    mrz_run_variants).  ctxs[0] is the parent (owns the chunk bytes), the others are variant contexts."""

    def __init__(self, device, max_chunk, nvar, library=None, **config):
        self.parent = Context(device, max_chunk, library=library, **config)
        self.ctxs = [self.parent] + [Context(parent=self.parent) for _ in range(nvar - 1)]
        self.nvar = nvar
        self.L = self.parent.L

    def close(self):
        for c in reversed(self.ctxs):
            c.close()
        self.ctxs = []

    def upload(self, arr):
        self.parent.upload(arr)

    def run(self, n, params, nvar=None):
        """-> list of MrzResult, one per victim_round_in in range(nvar)."""
        nvar = self.nvar if nvar is None else nvar
        hs = (C.c_void_p * nvar)(*[c.h for c in self.ctxs[:nvar]])
        res = (MrzResult * nvar)()
        rc = self.L.mrz_run_variants(hs, nvar, n, C.byref(params), res)
        if rc:
            raise MrzError("mrz_run_variants rc=%d: %s" % (rc, self.L.mrz_last_error(None).decode()))
        return list(res)

    def emit(self, v, s0=None, s1=None):
        """Streams of variant v (after run): -> (MrzResult, stream 0, stream 1)."""
        r = self.ctxs[v].run_emit()
        s0, s1 = self.ctxs[v].download(r, s0, s1)
        return r, s0, s1


def run_many(ctxs, ns, params):
    """mrz_run_many: the automata of several contexts (same device; chunks of different parents allowed) in ONE launch.
    -> list of MrzResult (as mrz_run_automaton); emit each with ctx.run_emit()."""
    m = len(ctxs)
    hs = (C.c_void_p * m)(*[c.h for c in ctxs])
    nn = (C.c_int64 * m)(*[int(x) for x in ns])
    pp = (MrzParams * m)(*params)
    res = (MrzResult * m)()
    rc = ctxs[0].L.mrz_run_many(hs, m, nn, pp, res)
    if rc:
        raise MrzError("mrz_run_many rc=%d: %s" % (rc, ctxs[0].L.mrz_last_error(None).decode()))
    return list(res)
