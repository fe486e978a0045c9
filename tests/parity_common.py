"""Shared parity checker: a context (CUDA library on the GPU, or the cusim build of the same kernel sources on
the CPU) against the CPU oracle, bit-exact (streams, counters, victim carry).  TEST INFRASTRUCTURE."""
import pytest

import helpers as H

STAT_KEYS = ("matches", "match_bytes", "literals", "literal_bytes", "tag_hits", "tag_misses", "inserts")


def first_token_diff(ctx, ref):
    got = [tuple(int(x) for x in r) for r in ctx.debug_tokens()]
    want = [t for t in ref["tokens"] if t[2] >= 0]
    for i, (g, w) in enumerate(zip(got, want)):
        if g != w:
            return "token %d: got %r want %r" % (i, g, w)
    return "token counts: got %d want %d" % (len(got), len(want))


def check(ctx, data, level=7, vr_in=0, cb=None):
    import mrz_py
    data = H.as_u8(data)
    cb = cb or H.chunk_bytes(data.size)
    hidx = H.hash_index()
    ref = H.oracle_chunk(data, level=level, cb=cb, vr_in=vr_in, hidx=hidx, want_tokens=True)
    prm = mrz_py.make_params(level, cb, hidx, vr_in)
    res, s0, s1 = ctx.rzip_chunk(data, prm)
    if s0.tobytes() != ref["s0"] or s1.tobytes() != ref["s1"]:
        pytest.fail("stream mismatch (n=%d level=%d): %s" % (data.size, level, first_token_diff(ctx, ref)))
    for k in STAT_KEYS:
        assert getattr(res, k) == ref["stats"][k], k
    assert res.victim_round_out == ref["vr_out"]
    assert res.victim_events == ref["stats"]["victim_events"]
    assert res.final_min_mask == ref["stats"]["final_min_mask"]
    assert res.n_ops == ref["stats"]["lookups"]
    assert res.gpu_launches > 0
    # and back: the device token replay (mrz_runzip_chunk, runzip.c:226-327) rebuilds the chunk from the two streams and
    # finds the checksum the encoder stored
    ures, back = ctx.runzip_chunk(s0, s1, cb, data.size)
    assert ures.out_len == data.size and ures.crc32 == ures.crc32_stored == res.crc32
    assert back.tobytes() == data.tobytes()
    assert ures.matches >= res.matches and ures.literals >= (1 if res.literal_bytes else 0)
    if hasattr(ctx.L, "mrz_sim_selfcheck_failures"):   # emulator build: its per-batch self-checks (table vs hash_count, ring) stayed quiet
        assert ctx.L.mrz_sim_selfcheck_failures() == 0
    return res, ref


def check_tagmap(ctx, data, hidx):
    """tagmap.cu against an independent restatement: tag[p] = X[p+31] ^ X[p] over a prefix-XOR of hash_index[buf[i]]."""
    import numpy as np
    data = H.as_u8(data)
    ctx.upload(data)
    bmap, bsum = ctx.debug_tagmap(data.size, hidx)
    h = hidx[data]
    x = np.concatenate([[0], np.bitwise_xor.accumulate(h)])
    tags = x[31:] ^ x[:-31]
    tz = np.zeros(tags.size, dtype=np.int64)
    t = tags.astype(np.uint64)
    for i in range(64):
        tz += ((t & np.uint64(1)) == 1) & (tz == i)
        t >>= np.uint64(1)
    want = np.zeros(data.size, dtype=np.uint8)
    want[:tags.size] = tz
    assert np.array_equal(bmap, want)
    ws = np.zeros((data.size + 63) // 64 * 64, dtype=np.uint8)
    ws[:data.size] = want
    assert np.array_equal(bsum, ws.reshape(-1, 64).max(axis=1))
    return want


def wide_hash_index(seed=5):
    """A hash_index whose every tag has >= 32 trailing one bits (31 values XORed, low words all ones): the map's rare
    64-bit path on every position."""
    import numpy as np
    rng = np.random.default_rng(seed)
    hi = rng.integers(0, 1 << 31, size=256, dtype=np.int64)
    return ((hi << 32) | 0xffffffff).astype(np.int64)


def check_runzip_rejects_corruption(ctx):
    """mrz_runzip_chunk refuses what runzip_chunk would refuse (runzip.c:178 corrupt history, :317-320 bad checksum)."""
    import numpy as np
    import mrz_py
    data = H.as_u8(H.gen_random_repeats(1 << 20, 12, lo=2048, hi=32768))
    cb = H.chunk_bytes(data.size)
    res, s0, s1 = ctx.rzip_chunk(data, mrz_py.make_params(7, cb, H.hash_index(), 0))
    assert res.matches > 0
    _, back = ctx.runzip_chunk(s0, s1, cb, data.size)
    assert back.tobytes() == data.tobytes()
    bad1 = s1.copy()
    bad1[bad1.size // 2] ^= 0x40                      # one literal byte differs: the chunk's CRC32 no longer matches
    with pytest.raises(mrz_py.MrzError, match="Bad checksum"):
        ctx.runzip_chunk(s0, bad1, cb, data.size)
    with pytest.raises(mrz_py.MrzError):              # the terminator and checksum are cut off
        ctx.runzip_chunk(s0[:-5], s1, cb, data.size)
    with pytest.raises(mrz_py.MrzError):              # stream 1 shorter than the headers say
        ctx.runzip_chunk(s0, s1[:-1], cb, data.size)
    first_match = np.array([1, 40, 0] + [0xff] * cb, dtype=np.uint8)   # a match as the first token: nothing to copy from
    with pytest.raises(mrz_py.MrzError, match="corrupt"):
        ctx.runzip_chunk(np.concatenate([first_match, s0]), s1, cb, data.size + 40)
    with pytest.raises(mrz_py.MrzError):              # caller's buffer too small
        ctx.runzip_chunk(s0, s1, cb, data.size - 1)
    _, back = ctx.runzip_chunk(s0, s1, cb, data.size)   # the context is still good
    assert back.tobytes() == data.tobytes()


def synthetic_streams(seed, n_tokens, cb=3, max_lit=300, max_len=500, max_ofs=600):
    """Stream 0 / stream 1 of a made-up chunk with many short, overlapping and chained matches (what the encoder rarely
    emits), and the chunk they describe (byte-wise replay, runzip.c:159-207)."""
    import zlib
    import numpy as np
    rng = np.random.default_rng(seed)
    out = bytearray()
    s0 = bytearray()
    s1 = bytearray()
    for _ in range(n_tokens):
        if not out or rng.random() < 0.4:
            ln = int(rng.integers(1, max_lit))
            lit = rng.integers(0, 256, size=ln, dtype=np.uint8).tobytes()
            s0 += bytes([0, ln & 255, ln >> 8])
            s1 += lit
            out += lit
        else:
            ln = int(rng.integers(1, max_len))
            ofs = int(rng.integers(1, min(len(out), max_ofs) + 1))
            s0 += bytes([1, ln & 255, ln >> 8]) + ofs.to_bytes(cb, "little")
            start = len(out) - ofs
            for i in range(ln):                      # byte-wise forward copy: overlaps repeat
                out.append(out[start + i])
    crc = zlib.crc32(bytes(out)) & 0xffffffff
    s0 += bytes([0, 0, 0]) + crc.to_bytes(4, "big")
    return np.frombuffer(bytes(s0), dtype=np.uint8), np.frombuffer(bytes(s1), dtype=np.uint8), bytes(out)


def check_runzip_synthetic(ctx, seeds=(1, 2, 3), n_tokens=3000):
    for seed in seeds:
        s0, s1, want = synthetic_streams(seed, n_tokens)
        res, back = ctx.runzip_chunk(s0, s1, 3, len(want))
        assert back.tobytes() == want, "seed %d" % seed
        assert res.waves > 10 and res.matches > n_tokens // 3     # chained matches really split into waves
