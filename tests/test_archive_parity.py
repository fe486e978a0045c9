"""Whole-archive parity: the product's `mrzip -n` host code (modern-rzip_b200/host) around the device call writes the
same bytes as the reference binary / the golden archives.  CPU variant: host code linked against the emulator build
of the kernels (tests/cusim/mrzip_sim).  GPU variant: the real binary, larger inputs, multi-chunk + multi-block."""
import hashlib
import os
import subprocess
import tempfile

import pytest

import helpers as H

ROOT = H.ROOT
SIM_BIN = os.path.join(ROOT, "tests", "cusim", "mrzip_sim")
GPU_BIN = os.path.join(ROOT, "modern-rzip_b200", "mrzip")
GOLD = H.load_golden()
CASES = H.golden_cases()


@pytest.fixture(scope="module")
def sim_bin():
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "tests", "cusim"), "mrzip_sim"], stdout=subprocess.DEVNULL)
    return SIM_BIN


def run(binary, fin, fout, flags=()):
    env = dict(os.environ, MRZIP="NOCONFIG", CUSIM_SEED="7")
    subprocess.run([binary, "-n", "-q", "-f", *flags, "-o", fout, fin], env=env, check=True, capture_output=True)
    return open(fout, "rb").read()


@pytest.mark.parametrize("name", ["survey650", "pattern0", "pattern1", "pattern31", "pattern64", "pattern100",
                                  "random100003", "repeats1m", "dense_R3"])
def test_sim_archive_equals_golden(sim_bin, name):
    data, flags = CASES[name]
    with tempfile.TemporaryDirectory() as d:
        fin = os.path.join(d, "in.bin")
        H.as_u8(data).tofile(fin)
        raw = run(sim_bin, fin, os.path.join(d, "o.mar"), flags)
    g = GOLD[name]
    assert len(raw) == g["archive_len"]
    assert hashlib.sha256(raw).hexdigest() == g["archive_sha256"]
    if "archive_hex" in g:
        assert raw.hex() == g["archive_hex"]


@pytest.mark.skipif(not H.have_ref(), reason="reference binary not built")
def test_sim_two_files_one_invocation(sim_bin):
    """random() state and victim_round continue across the files of one command line (main.c:651, rzip.c:259)."""
    a, b = H.gen_motif(1 << 20, 2000, seed=3), H.gen_motif(1 << 20, 2000, seed=4)
    with tempfile.TemporaryDirectory() as d:
        outs = {}
        for tag, binary in (("ref", H.REF_BIN), ("sim", sim_bin)):
            od = os.path.join(d, tag) + "/"
            os.mkdir(od)
            fa, fb = os.path.join(od, "a.bin"), os.path.join(od, "b.bin")
            a.tofile(fa); b.tofile(fb)
            subprocess.run([binary, "-n", "-q", "-f", "-R", "1", fa, fb], env=dict(os.environ, MRZIP="NOCONFIG"), check=True,
                           capture_output=True)
            outs[tag] = (open(fa + ".lrz", "rb").read(), open(fb + ".lrz", "rb").read())
        assert outs["ref"][0] == outs["sim"][0]
        assert outs["ref"][1] == outs["sim"][1]


# ------------------------------------------------------------------------------------------------ GPU
def gpu_vs_reference(data, flags, expect_chunks=None):
    assert H.have_ref(), "oracle/_ref/mrzip must travel to the GPU box (built by __graft_entry__.build())"
    tmp = "/dev/shm" if os.path.isdir("/dev/shm") else None
    with tempfile.TemporaryDirectory(dir=tmp) as d:
        fin = os.path.join(d, "in.bin")
        H.as_u8(data).tofile(fin)
        got = run(GPU_BIN, fin, os.path.join(d, "gpu.mar"), flags)
        H.run_reference(fin, os.path.join(d, "ref.mar"), flags)
        want = open(os.path.join(d, "ref.mar"), "rb").read()
        assert len(got) == len(want)
        assert got == want, "archive differs from the reference's"
        if expect_chunks:
            assert len(H.parse_archive(got)["chunks"]) == expect_chunks
        # the reference decodes our archive back to the input (BLAKE2b of the decoded output)
        subprocess.run([H.REF_BIN, "-d", "-q", "-f", "-o", os.path.join(d, "back.bin"), os.path.join(d, "gpu.mar")],
                       env=dict(os.environ, MRZIP="NOCONFIG"), check=True, capture_output=True)
        assert H.blake2b(open(os.path.join(d, "back.bin"), "rb").read()) == H.blake2b(H.as_u8(data).tobytes())
        # and the product's own way back (host/runzip_fd.c + mrz_runzip_chunk) decodes its archive and the reference's
        for arc in ("gpu.mar", "ref.mar"):
            assert decode(GPU_BIN, os.path.join(d, arc), os.path.join(d, "back2.bin")) == H.as_u8(data).tobytes()


def decode(binary, farc, fout, expect_ok=True):
    env = dict(os.environ, MRZIP="NOCONFIG", CUSIM_SEED="7")
    p = subprocess.run([binary, "-d", "-q", "-f", "-o", fout, farc], env=env, capture_output=True, text=True)
    if not expect_ok:
        return p
    assert p.returncode == 0, p.stderr[-2000:]
    return open(fout, "rb").read()


@pytest.mark.parametrize("name", ["survey650", "pattern0", "pattern31", "random100003", "repeats1m", "dense_R3"])
def test_sim_decode_roundtrip(sim_bin, name):
    """`mrzip -d` of this build (stream read side + device token replay, runzip.c:226-327) gives the input back, from its
    own archive and - when the reference binary is here - from the reference's."""
    data, flags = CASES[name]
    with tempfile.TemporaryDirectory() as d:
        fin = os.path.join(d, "in.bin")
        H.as_u8(data).tofile(fin)
        run(sim_bin, fin, os.path.join(d, "o.mar"), flags)
        assert decode(sim_bin, os.path.join(d, "o.mar"), os.path.join(d, "back.bin")) == H.as_u8(data).tobytes()
        if H.have_ref():
            H.run_reference(fin, os.path.join(d, "ref.mar"), flags)
            assert decode(sim_bin, os.path.join(d, "ref.mar"), os.path.join(d, "back2.bin")) == H.as_u8(data).tobytes()


def test_sim_decode_refuses_a_damaged_archive(sim_bin):
    data = H.gen_random_repeats(1 << 20, 12, lo=2048, hi=32768)
    with tempfile.TemporaryDirectory() as d:
        fin = os.path.join(d, "in.bin")
        H.as_u8(data).tofile(fin)
        raw = bytearray(run(sim_bin, fin, os.path.join(d, "o.mar")))
        raw[len(raw) // 2] ^= 0x10                                    # a literal byte
        open(os.path.join(d, "bad.mar"), "wb").write(raw)
        p = decode(sim_bin, os.path.join(d, "bad.mar"), os.path.join(d, "x.bin"), expect_ok=False)
        assert p.returncode != 0 and "Bad checksum" in p.stderr
        raw[len(raw) // 2] ^= 0x10
        raw[-1] ^= 1                                                  # the stored MD5
        open(os.path.join(d, "bad2.mar"), "wb").write(raw)
        p = decode(sim_bin, os.path.join(d, "bad2.mar"), os.path.join(d, "y.bin"), expect_ok=False)
        assert p.returncode != 0 and "MD5 CHECK FAILED" in p.stderr


def test_sim_decode_survives_damaged_headers(sim_bin):
    """One damaged byte anywhere in the container / chunk / block headers: `mrzip -d` either still gives the input back (a
    byte that does not matter, e.g. the stored levels) or stops with an error - it never hangs, crashes or writes wrong data
    silently."""
    data = H.gen_random_repeats(200000, 6, lo=2048, hi=8192)
    want = H.as_u8(data).tobytes()
    with tempfile.TemporaryDirectory() as d:
        fin = os.path.join(d, "in.bin")
        H.as_u8(data).tofile(fin)
        raw = run(sim_bin, fin, os.path.join(d, "o.mar"))
        env = dict(os.environ, MRZIP="NOCONFIG")
        refused = 0
        for pos in list(range(0, 70)) + [len(raw) - 17, len(raw) - 16]:
            bad = bytearray(raw)
            bad[pos] ^= 0x5a
            fbad, fout = os.path.join(d, "bad.mar"), os.path.join(d, "bad.out")
            open(fbad, "wb").write(bad)
            p = subprocess.run([sim_bin, "-d", "-q", "-f", "-o", fout, fbad], env=env, capture_output=True, text=True, timeout=60)
            assert p.returncode in (0, 1), "byte %d: killed by a signal (%d)\n%s" % (pos, p.returncode, p.stderr[-500:])
            if p.returncode == 0:
                assert open(fout, "rb").read() == want, "byte %d: wrong output accepted" % pos
            else:
                refused += 1
        assert refused > 20


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["survey650", "pattern0", "motif_R1", "vm_like_L9", "tar_like"])
def test_gpu_archive_equals_golden(name):
    data, flags = CASES[name]
    tmp = "/dev/shm" if os.path.isdir("/dev/shm") else None
    with tempfile.TemporaryDirectory(dir=tmp) as d:
        fin = os.path.join(d, "in.bin")
        H.as_u8(data).tofile(fin)
        raw = run(GPU_BIN, fin, os.path.join(d, "o.mar"), flags)
    assert hashlib.sha256(raw).hexdigest() == GOLD[name]["archive_sha256"]


@pytest.mark.gpu
def test_gpu_multi_chunk_multi_block_vs_reference():
    # -m 1: ramsize 100 MB => 66 MB chunks, 33 MB stream buffers => 3 chunks, several blocks per stream, victim carry
    gpu_vs_reference(H.gen_vm_like(150 << 20, seed=31), ("-m", "1"), expect_chunks=3)


@pytest.mark.gpu
def test_gpu_config1_like_vs_reference():
    gpu_vs_reference(H.gen_random_repeats(64 << 20, 500), ("-m", "40"))


@pytest.mark.gpu
def test_gpu_window_chunks_level9_vs_reference():
    gpu_vs_reference(H.gen_dense(230 << 20, seed=15), ("-w", "1", "-L", "9"), expect_chunks=3)
