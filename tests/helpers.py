"""Shared test helpers: oracle bindings, reference-binary runner, archive parser, data generators.

TEST INFRASTRUCTURE ONLY.  Nothing in the product imports this file.
Archive layout parsed here follows /root/reference/src/stream.c:1199-1293 (chunk header,
initial stream headers, block headers with next_head links) and src/mrzip.c:127-188 (magic).
"""
import ctypes as C
import hashlib
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
ORACLE_SO = os.path.join(ORACLE_DIR, "librzip_oracle.so")
REF_BIN = os.path.join(ORACLE_DIR, "_ref", "mrzip")
GOLDEN = os.path.join(ROOT, "tests", "golden")


class OrcStats(C.Structure):
    _fields_ = [(n, C.c_int64) for n in (
        "matches", "match_bytes", "literals", "literal_bytes", "tag_hits", "tag_misses", "inserts",
        "victim_events", "displaced", "minbit_overwrites", "sweeps_wrapped", "final_min_mask",
        "lookups", "slots_scanned")]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


class SpecStats(C.Structure):
    _fields_ = [(n, C.c_int64) for n in (
        "batches", "planned_commits", "serial_ops", "static_conflicts", "extra_conflicts", "complex_ops",
        "emit_cuts", "count_cuts")]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


class OrcToken(C.Structure):
    _fields_ = [("pos", C.c_int64), ("len", C.c_int64), ("ofs", C.c_int64)]


_orc = None


def build_oracle():
    srcs = [os.path.join(ORACLE_DIR, f) for f in ("rzip_oracle.c", "spec_model.c", "rzip_oracle.h")]
    if not os.path.exists(ORACLE_SO) or os.path.getmtime(ORACLE_SO) < max(os.path.getmtime(f) for f in srcs):
        subprocess.check_call(["make", "-C", ORACLE_DIR, "librzip_oracle.so"], stdout=subprocess.DEVNULL)


def orc():
    global _orc
    if _orc is None:
        build_oracle()
        lib = C.CDLL(ORACLE_SO)
        lib.orc_hash_index.argtypes = [C.POINTER(C.c_int64), C.c_int64]
        lib.orc_chunk_bytes.argtypes = [C.c_int64]
        lib.orc_crc32.argtypes = [C.c_uint32, C.c_void_p, C.c_int64]
        lib.orc_crc32.restype = C.c_uint32
        lib.orc_rzip_chunk.argtypes = [
            C.c_void_p, C.c_int64, C.POINTER(C.c_int64), C.c_int, C.c_int, C.POINTER(C.c_int64),
            C.c_void_p, C.c_int64, C.POINTER(C.c_int64), C.c_void_p, C.c_int64, C.POINTER(C.c_int64),
            C.c_void_p, C.c_int64, C.POINTER(C.c_int64), C.POINTER(OrcStats)]
        lib.orc_rzip_chunk_spec.argtypes = [
            C.c_void_p, C.c_int64, C.POINTER(C.c_int64), C.c_int, C.c_int, C.POINTER(C.c_int64),
            C.c_void_p, C.c_int64, C.POINTER(C.c_int64), C.c_void_p, C.c_int64, C.POINTER(C.c_int64),
            C.POINTER(OrcStats), C.c_int, C.POINTER(SpecStats)]
        _orc = lib
    return _orc


def hash_index(skip=0):
    arr = (C.c_int64 * 256)()
    orc().orc_hash_index(arr, skip)
    return np.frombuffer(arr, dtype=np.int64).copy()


def chunk_bytes(n):
    return orc().orc_chunk_bytes(n)


def as_u8(data):
    if isinstance(data, np.ndarray):
        return np.ascontiguousarray(data, dtype=np.uint8)
    return np.frombuffer(bytes(data), dtype=np.uint8)


def oracle_chunk(data, level=7, cb=None, vr_in=0, hidx=None, want_tokens=False):
    """Run the CPU restatement on one chunk.  Returns dict(s0, s1, stats, vr_out[, tokens])."""
    a = as_u8(data)
    n = a.size
    if cb is None:
        cb = chunk_bytes(n)
    if hidx is None:
        hidx = hash_index()
    hi = (C.c_int64 * 256)(*[int(x) for x in hidx])
    s0 = np.empty(n // 2 + 64, dtype=np.uint8)
    s1 = np.empty(n + 16, dtype=np.uint8)
    s0l, s1l, ntok = C.c_int64(), C.c_int64(), C.c_int64()
    vr = C.c_int64(vr_in)
    st = OrcStats()
    tok_cap = (n // 16 + 16) if want_tokens else 0
    toks = (OrcToken * tok_cap)() if want_tokens else None
    buf = a if n else np.zeros(1, dtype=np.uint8)
    rc = orc().orc_rzip_chunk(buf.ctypes.data, n, hi, level, cb, C.byref(vr),
                              s0.ctypes.data, s0.size, C.byref(s0l), s1.ctypes.data, s1.size, C.byref(s1l),
                              C.cast(toks, C.c_void_p) if want_tokens else None, tok_cap, C.byref(ntok),
                              C.byref(st))
    assert rc == 0, "oracle buffer overflow"
    out = dict(s0=s0[:s0l.value].tobytes(), s1=s1[:s1l.value].tobytes(), stats=st.as_dict(), vr_out=vr.value)
    if want_tokens:
        out["tokens"] = [(toks[i].pos, toks[i].len, toks[i].ofs) for i in range(ntok.value)]
    return out


def spec_model_chunk(data, level=7, cb=None, vr_in=0, hidx=None, batch=1024):
    """CPU model of the GPU's batched speculation (oracle/spec_model.c)."""
    a = as_u8(data)
    n = a.size
    cb = cb or chunk_bytes(n)
    hidx = hash_index() if hidx is None else hidx
    hi = (C.c_int64 * 256)(*[int(x) for x in hidx])
    s0 = np.empty(n // 2 + 64, dtype=np.uint8)
    s1 = np.empty(n + 16, dtype=np.uint8)
    s0l, s1l = C.c_int64(), C.c_int64()
    vr = C.c_int64(vr_in)
    st, ss = OrcStats(), SpecStats()
    buf = a if n else np.zeros(1, dtype=np.uint8)
    rc = orc().orc_rzip_chunk_spec(buf.ctypes.data, n, hi, level, cb, C.byref(vr), s0.ctypes.data, s0.size,
                                   C.byref(s0l), s1.ctypes.data, s1.size, C.byref(s1l), C.byref(st), batch, C.byref(ss))
    assert rc == 0
    return dict(s0=s0[:s0l.value].tobytes(), s1=s1[:s1l.value].tobytes(), stats=st.as_dict(), vr_out=vr.value,
                spec=ss.as_dict())


# ---------------------------------------------------------------- reference binary
def have_ref():
    return os.path.exists(REF_BIN) and os.access(REF_BIN, os.X_OK)


def run_mrzip(binary, infile, outfile, extra=(), env_extra=None):
    env = dict(os.environ, MRZIP="NOCONFIG")
    if env_extra:
        env.update(env_extra)
    cmd = [binary, "-n", "-q", "-f", *extra, "-o", outfile, infile]
    return subprocess.run(cmd, env=env, capture_output=True, text=True, check=True)


def run_reference(infile, outfile, extra=()):
    return run_mrzip(REF_BIN, infile, outfile, extra)


# ---------------------------------------------------------------- archive parser
def _le(b):
    return int.from_bytes(b, "little")


def parse_archive(raw):
    """-> dict(magic=..., chunks=[dict(cb, eof, size, s0, s1, blocks=[(stream, hdr_pos, ctype, c_len, u_len)])], trailer=bytes)"""
    assert raw[:4] == b"MRZI", raw[:4]
    magic = dict(major=raw[4], minor=raw[5], st_size=_le(raw[6:14]), hash_code=raw[14], enc=raw[15],
                 props=raw[17], levels=raw[18], comment_len=raw[19])
    pos = 20 + raw[19]
    chunks = []
    while True:
        cb, eof = raw[pos], raw[pos + 1]
        size = _le(raw[pos + 2:pos + 2 + cb])
        init = pos + 2 + cb
        hl = 1 + 3 * cb
        streams, blocks, endpos = [], [], init + 2 * hl
        for j in range(2):
            h = init + j * hl
            assert raw[h] == 3
            nxt = _le(raw[h + 1 + 2 * cb:h + hl])
            data = bytearray()
            while nxt:
                b = init + nxt
                ctype, c_len, u_len = raw[b], _le(raw[b + 1:b + 1 + cb]), _le(raw[b + 1 + cb:b + 1 + 2 * cb])
                nxt = _le(raw[b + 1 + 2 * cb:b + hl])
                assert ctype == 3, "only CTYPE_NONE blocks are parsed here"
                data += raw[b + hl:b + hl + c_len]
                blocks.append((j, b - init, ctype, c_len, u_len))
                endpos = max(endpos, b + hl + c_len)
            streams.append(bytes(data))
        chunks.append(dict(cb=cb, eof=eof, size=size, s0=streams[0], s1=streams[1], blocks=blocks))
        pos = endpos
        if eof:
            break
    return dict(magic=magic, chunks=chunks, trailer=bytes(raw[pos:]))


def parse_archive_hashed(path):
    """parse_archive for archives too large to hold as bytes: walks the file through np.memmap and returns per-chunk
    SHA-256 of the two streams instead of the streams themselves (same layout, stream.c:1199-1293)."""
    raw = np.memmap(path, dtype=np.uint8, mode="r")
    total = raw.size

    def le(a, b):
        return int.from_bytes(bytes(raw[a:b]), "little")
    assert bytes(raw[:4]) == b"MRZI"
    magic = dict(major=int(raw[4]), minor=int(raw[5]), st_size=le(6, 14), hash_code=int(raw[14]), levels=int(raw[18]))
    pos = 20 + int(raw[19])
    chunks = []
    while True:
        cb, eof = int(raw[pos]), int(raw[pos + 1])
        size = le(pos + 2, pos + 2 + cb)
        init = pos + 2 + cb
        hl = 1 + 3 * cb
        endpos = init + 2 * hl
        rec = dict(cb=cb, eof=eof, size=size, n_blocks=0)
        for j in range(2):
            h = init + j * hl
            assert int(raw[h]) == 3
            nxt = le(h + 1 + 2 * cb, h + hl)
            sha, ln = hashlib.sha256(), 0
            while nxt:
                b = init + nxt
                ctype, c_len, u_len = int(raw[b]), le(b + 1, b + 1 + cb), le(b + 1 + cb, b + 1 + 2 * cb)
                nxt = le(b + 1 + 2 * cb, b + hl)
                assert ctype == 3 and c_len == u_len, "only CTYPE_NONE blocks are parsed here"
                for o in range(b + hl, b + hl + c_len, 1 << 26):
                    sha.update(raw[o:min(o + (1 << 26), b + hl + c_len)])
                ln += c_len
                rec["n_blocks"] += 1
                endpos = max(endpos, b + hl + c_len)
            rec["s%d_len" % j] = ln
            rec["s%d_sha256" % j] = sha.hexdigest()
        chunks.append(rec)
        pos = endpos
        if eof:
            break
    trailer = bytes(raw[pos:total])
    del raw
    return dict(magic=magic, chunks=chunks, trailer=trailer)


def sha256_file(path):
    h = hashlib.sha256()
    with open(path, "rb") as f:
        while True:
            b = f.read(1 << 26)
            if not b:
                break
            h.update(b)
    return h.hexdigest()


def decode_tokens(s0, s1, cb, history=b""):
    """Replay a chunk's streams (runzip.c:120-207 semantics) -> bytes; checks the CRC."""
    out = bytearray()
    i = j = 0
    while True:
        head, ln = s0[i], _le(s0[i + 1:i + 3])
        i += 3
        if head == 0:
            if ln == 0:
                break
            out += s1[j:j + ln]
            j += ln
        else:
            dist = _le(s0[i:i + cb])
            i += cb
            for _ in range(ln):  # byte-wise: overlapping copies allowed
                out.append(out[-dist])
    crc = int.from_bytes(s0[i:i + 4], "big")
    import zlib
    assert crc == zlib.crc32(bytes(out)), "chunk CRC mismatch"
    assert i + 4 == len(s0) and j == len(s1)
    return bytes(out)


def blake2b(data):
    return hashlib.blake2b(bytes(data)).hexdigest()


# ---------------------------------------------------------------- synthetic data (scaled)
def gen_random_repeats(n, n_inject, seed=12345, lo=4096, hi=262144):
    """config 1 generator: uniform bytes + sequential injected copies."""
    rng = np.random.default_rng(seed)
    buf = rng.integers(0, 256, n, dtype=np.uint8)
    hi = min(hi, max(lo + 1, n // 8))
    for _ in range(n_inject):
        L = int(rng.integers(lo, hi))
        dst = int(rng.integers(L + 1, n - L))
        src = int(rng.integers(0, dst - L))
        buf[dst:dst + L] = buf[src:src + L]
    return buf


def gen_random(n, seed=4):
    return np.random.default_rng(seed).integers(0, 256, n, dtype=np.uint8)


def gen_vm_like(n, seed=3, block=4096):
    """config 3 generator: 4 KiB blocks: 35% zero, 25% copy of an earlier block, 20% random, 20% text-like."""
    rng = np.random.default_rng(seed)
    nb = n // block
    buf = np.zeros(nb * block, dtype=np.uint8)
    kinds = rng.random(nb)
    alphabet = np.frombuffer(b"abcdefghijklmnopqrstuvwxyzABCDEFGHIJKLMNOPQRSTUVWXYZ0123456789 \n", dtype=np.uint8)
    for b in range(nb):
        k = kinds[b]
        s = b * block
        if k < 0.35:
            continue
        if k < 0.60 and b > 0:
            src = int(rng.integers(0, b)) * block
            buf[s:s + block] = buf[src:src + block]
        elif k < 0.80:
            buf[s:s + block] = rng.integers(0, 256, block, dtype=np.uint8)
        else:
            buf[s:s + block] = alphabet[rng.integers(0, 64, block)]
    if buf.size < n:
        buf = np.concatenate([buf, np.zeros(n - buf.size, dtype=np.uint8)])
    return buf


def gen_dense(n, seed=5, base=1 << 20, mutations=100):
    """config 5 generator: one random base block repeated with point mutations per copy."""
    rng = np.random.default_rng(seed)
    base = min(base, max(4096, n // 8))
    blk = rng.integers(0, 256, base, dtype=np.uint8)
    reps = (n + base - 1) // base
    buf = np.tile(blk, reps)[:n].copy()
    for r in range(reps):
        m = max(1, mutations * min(base, n - r * base) // base)
        pos = rng.integers(r * base, min(n, (r + 1) * base), m)
        buf[pos] = rng.integers(0, 256, m, dtype=np.uint8)
    return buf


def gen_tar_like(n, seed=2, max_body=1 << 20):
    """config 2 generator (scaled): 512-byte headers + bodies; 35% verbatim copies of earlier bodies."""
    rng = np.random.default_rng(seed)
    out = np.zeros(n, dtype=np.uint8)
    bodies = []
    pos = 0
    idx = 0
    alphabet = np.frombuffer(b"etaoinshrdlucmfwypvbgkqjxz ETAOINSHRDLUCMFWYPVBGKQJXZ.,;\n0123456789", dtype=np.uint8)[:64]
    while pos + 1024 <= n:
        size = int(np.exp(rng.uniform(np.log(1024), np.log(max_body))))
        size = min(size, n - pos - 512)
        hdr = np.zeros(512, dtype=np.uint8)
        name = ("member%08d.dat" % idx).encode()
        hdr[:len(name)] = np.frombuffer(name, dtype=np.uint8)
        sz = ("%011o" % size).encode()
        hdr[124:124 + len(sz)] = np.frombuffer(sz, dtype=np.uint8)
        out[pos:pos + 512] = hdr
        pos += 512
        r = rng.random()
        if r < 0.35 and bodies:
            s, l = bodies[int(rng.integers(0, len(bodies)))]
            l = min(l, size)
            out[pos:pos + l] = out[s:s + l]
            size = l
        elif r < 0.675:
            out[pos:pos + size] = rng.integers(0, 256, size, dtype=np.uint8)
        else:
            out[pos:pos + size] = alphabet[rng.integers(0, 64, size)]
        bodies.append((pos, size))
        pos += (size + 511) // 512 * 512
        idx += 1
    return out


def gen_motif(n, copies, motif_len=48, seed=7):
    """victim/displacement stressor: random bytes with many copies of one short motif (SURVEY §8c vector 3)."""
    rng = np.random.default_rng(seed)
    buf = rng.integers(0, 256, n, dtype=np.uint8)
    motif = rng.integers(0, 256, motif_len, dtype=np.uint8)
    for p in rng.integers(0, n - motif_len, copies):
        buf[p:p + motif_len] = motif
    return buf


# ---------------------------------------------------------------- golden cases (tests/golden/make_golden.py)
def survey_650():
    """SURVEY.md §8(a) verified example: A(300) B(100) A[:200] C(50), Python random.seed(1)."""
    import random
    random.seed(1)
    A = bytes(random.getrandbits(8) for _ in range(300))
    B = bytes(random.getrandbits(8) for _ in range(100))
    Cc = bytes(random.getrandbits(8) for _ in range(50))
    return A + B + A[:200] + Cc


def golden_cases():
    """name -> (input bytes, extra mrzip flags).  All single-process, `-n -q -f`, MRZIP=NOCONFIG."""
    pat = b"abcdefghijklmnopqrstuvwxyz012345"
    cases = {"survey650": (survey_650(), ())}
    for n in (0, 1, 30, 31, 32, 63, 64, 100):
        cases["pattern%d" % n] = ((pat * 4)[:n], ())
    cases["random100003"] = (gen_random(100003, seed=5), ())
    cases["repeats1m"] = (gen_random_repeats(1 << 20, 12, lo=1024, hi=32768), ())
    cases["motif_R1"] = (gen_motif(3 << 20, 3000), ("-R", "1"))
    cases["vm_like_L9"] = (gen_vm_like(2 << 20, seed=19), ("-L", "9"))
    cases["dense_R3"] = (gen_dense(2 << 20, seed=6), ("-R", "3"))
    cases["tar_like"] = (gen_tar_like(4 << 20, seed=8), ())
    return cases


def load_golden():
    import json
    with open(os.path.join(GOLDEN, "reference_vectors.json")) as f:
        return json.load(f)
