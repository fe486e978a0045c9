"""`-m gpu` parity at BASELINE.json's sizes, through the reference binary where it finishes in seconds and through
size-independent properties (replay of the token streams == input, CRC, counters) at the full 4 GiB of configs[1]."""
import os
import sys
import zlib

import numpy as np
import pytest

import helpers as H
from test_archive_parity import gpu_vs_reference

ROOT = H.ROOT
sys.path.insert(0, ROOT)


def replay(s0, s1, cb, n):
    """Vectorised replay of a chunk's streams (runzip.c:120-207): literal runs are slice copies, matches are slice
    copies repeated with period `dist` when they overlap their own output."""
    out = np.empty(n, dtype=np.uint8)
    s0 = np.asarray(s0)
    i = j = p = 0
    nmatch = 0
    while True:
        head, ln = int(s0[i]), int(s0[i + 1]) | (int(s0[i + 2]) << 8)
        i += 3
        if head == 0:
            if ln == 0:
                break
            out[p:p + ln] = s1[j:j + ln]
            j += ln
            p += ln
        else:
            dist = int.from_bytes(bytes(s0[i:i + cb]), "little")
            i += cb
            assert 0 < dist <= p, "match reaches before the start of the chunk"
            done = 0
            while done < ln:   # period-dist copy
                step = min(dist, ln - done)
                out[p + done:p + done + step] = out[p + done - dist:p + done - dist + step]
                done += step
            p += ln
            nmatch += 1
    assert p == n and j == len(s1) and i + 4 == len(s0)
    return out, int.from_bytes(bytes(s0[i:i + 4]), "big"), nmatch


@pytest.mark.gpu
def test_config0_256m_random_repeats_vs_reference():
    """configs[0]: 256 MiB uniform bytes with 2000 injected repeats, `mrzip -n`: archive byte-identical to the
    reference's, and the reference decodes ours back to the input (BLAKE2b)."""
    gpu_vs_reference(H.gen_random_repeats(256 << 20, 2000), ("-m", "40"))


@pytest.mark.gpu
def test_1gib_tar_like_single_chunk_vs_reference():
    import bench
    gpu_vs_reference(bench.gen_tar_like(1 << 30, seed=2), ("-m", "20"), expect_chunks=1)


@pytest.mark.gpu
@pytest.mark.parametrize("kind", ["tar", "vm"])
def test_config1_4gib_round_trip_properties(kind):
    """configs[1] at full size (4 GiB, one chunk, 5-byte distances): the streams replay to the input, the CRC in
    stream 0 is the input's, the counters add up.  configs[2]'s data kind at the same size rides along."""
    import bench
    import mrz_py
    import psutil
    n = 4 << 30
    if psutil.virtual_memory().available < 4 * n:
        n = 1 << 30
    pin = mrz_py.PinnedBuffer(n)
    data = pin.array
    (bench.gen_vm_like if kind == "vm" else bench.gen_tar_like)(n, seed=2, out=data)
    ctx = mrz_py.Context(0, n)
    try:
        cb = H.chunk_bytes(n)
        res, s0, s1 = ctx.rzip_chunk(data, mrz_py.make_params(7, cb, H.hash_index(), 0))
        assert res.match_bytes + res.literal_bytes == n
        assert res.literal_bytes == len(s1)
        out, crc, nmatch = replay(s0, s1, cb, n)
        assert crc == res.crc32 == zlib.crc32(data)
        assert nmatch >= res.matches          # matches longer than 65535 bytes are split into pieces (rzip.c:207-219)
        assert np.array_equal(out, data)
    finally:
        ctx.close()
