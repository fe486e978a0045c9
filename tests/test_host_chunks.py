"""The host's chunk scheduler (modern-rzip_b200/host/rzip_fd.c: slots, variant-parallel chunks, ordered writer) on the
CPU: the product's host code linked against the emulator build of the kernels, built with 256 KiB window units
(tests/cusim/mrzip_sim_small) so that a few MiB make a multi-chunk file.  Every chunk's two streams must equal the
oracle's with the victim_round chain carried exactly (rzip.c:259), whatever the number of chunks in flight."""
import os
import subprocess
import tempfile

import numpy as np
import pytest

import helpers as H

ROOT = H.ROOT
BIN = os.path.join(ROOT, "tests", "cusim", "mrzip_sim_small")
UNIT = 262144


@pytest.fixture(scope="module")
def small_bin():
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "tests", "cusim"), "mrzip_sim_small"], stdout=subprocess.DEVNULL)
    return BIN


def oracle_chain(data, chunk, level, hidx):
    out, vr = [], 0
    for lo in range(0, max(1, data.size), chunk):
        piece = data[lo:lo + chunk]
        r = H.oracle_chunk(piece, level=level, cb=H.chunk_bytes(piece.size), vr_in=vr, hidx=hidx)
        out.append(r)
        vr = r["vr_out"]
    return out


@pytest.mark.parametrize("slots", [1, 3])
def test_multi_chunk_variants_vs_oracle_chain(small_bin, slots):
    level, w = 1, 1                               # chain length 2; chunks of 256 KiB
    data = np.concatenate([H.gen_motif(UNIT, 600, seed=s) for s in (3, 4, 5)] + [H.gen_motif(70000, 200, seed=6)])
    hidx = H.hash_index()
    want = oracle_chain(data, w * UNIT, level, hidx)
    assert sum(r["stats"]["victim_events"] > 0 for r in want) >= 3, "the chunks must depend on the carried victim_round"
    assert len({r["vr_out"] for r in want}) > 1 or want[0]["vr_out"] != 0
    with tempfile.TemporaryDirectory() as d:
        fin, fout = os.path.join(d, "in.bin"), os.path.join(d, "out.mar")
        data.tofile(fin)
        p = subprocess.run([small_bin, "-n", "-vv", "-f", "-R", str(level), "-w", str(w), "-G", str(slots), "-o", fout, fin],
                           env=dict(os.environ, MRZIP="NOCONFIG"), capture_output=True, text=True)
        assert p.returncode == 0, p.stderr[-2000:]
        ar = H.parse_archive(open(fout, "rb").read())
        # the way back, chunk after chunk (host/runzip_fd.c)
        fback = os.path.join(d, "back.bin")
        q = subprocess.run([small_bin, "-d", "-vv", "-f", "-o", fback, fout], env=dict(os.environ, MRZIP="NOCONFIG"), capture_output=True, text=True)
        assert q.returncode == 0, q.stderr[-2000:]
        assert open(fback, "rb").read() == data.tobytes() and "gpu: 4 chunk(s)" in q.stderr
    assert len(ar["chunks"]) == len(want) == 4
    for k, (c, r) in enumerate(zip(ar["chunks"], want)):
        assert c["s0"] == r["s0"], "stream 0 of chunk %d" % k
        assert c["s1"] == r["s1"], "stream 1 of chunk %d" % k
        assert c["eof"] == (k == 3)
    import hashlib
    assert ar["trailer"] == hashlib.md5(data.tobytes()).digest()
    if slots > 1:
        assert "chunk(s) searched for every victim_round" in p.stderr and " 0 chunk(s) searched" not in p.stderr


def test_empty_and_single(small_bin):
    for data in (np.zeros(0, dtype=np.uint8), H.gen_random(1000, seed=3)):
        want = oracle_chain(data, 4 * UNIT, 7, H.hash_index())
        with tempfile.TemporaryDirectory() as d:
            fin, fout = os.path.join(d, "in.bin"), os.path.join(d, "out.mar")
            data.tofile(fin)
            subprocess.run([small_bin, "-n", "-q", "-f", "-w", "4", "-o", fout, fin], env=dict(os.environ, MRZIP="NOCONFIG"), check=True,
                           capture_output=True)
            ar = H.parse_archive(open(fout, "rb").read())
            fback = os.path.join(d, "back.bin")
            subprocess.run([small_bin, "-d", "-q", "-f", "-o", fback, fout], env=dict(os.environ, MRZIP="NOCONFIG"), check=True, capture_output=True)
            assert open(fback, "rb").read() == data.tobytes()
        assert len(ar["chunks"]) == 1 and ar["chunks"][0]["s0"] == want[0]["s0"] and ar["chunks"][0]["s1"] == want[0]["s1"]
