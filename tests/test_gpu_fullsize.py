"""`-m gpu` parity at BASELINE.json's sizes against golden vectors minted from the UNMODIFIED reference binary in the
build container (tests/golden/fullsize_vectors.json, tests/golden/make_fullsize_golden.py): the product's `mrzip -n`
archive must have the reference archive's SHA-256, chunk by chunk the reference's two streams, and its MD5 trailer.
Inputs are regenerated from the seeds of tests/fullsize_cases.py with bench.py's generators.  Where the box allows it
the reference binary also decodes our archive back (BLAKE2b of the decoded output == input's)."""
import hashlib
import json
import os
import subprocess
import sys
import tempfile

import numpy as np
import pytest

import helpers as H
from fullsize_cases import CASES

ROOT = H.ROOT
sys.path.insert(0, ROOT)
GPU_BIN = os.path.join(ROOT, "modern-rzip_b200", "mrzip")
GOLD = json.load(open(os.path.join(H.GOLDEN, "fullsize_vectors.json")))
pytestmark = pytest.mark.gpu
# The whole GPU suite has to fit the 20-minute window of the round-end run.  Two of the three 32 GiB cases (3-4 minutes
# each, mostly generating and hashing 32 GiB) therefore run only on request; they passed on the B200 in round 2
# (profiles/r2_gpu_tests_full_earlier_build.log), and `bench.py --workload vm32 --window 41` / `bench.py --gpus N` check
# the same archives / streams against the same golden vectors on every run (profiles/r2_bench_vm32_w41_*.json,
# r2_bench_tar_*gpu.json).  The 33-chunk case stays in the default suite.
all_32g = pytest.mark.skipif(not os.environ.get("MRZ_TEST_32G_ALL"), reason="set MRZ_TEST_32G_ALL=1 for every 32 GiB case (adds ~8 minutes)")


def blake2b_file(path):
    h = hashlib.blake2b()
    with open(path, "rb") as f:
        while True:
            b = f.read(1 << 26)
            if not b:
                break
            h.update(b)
    return h.hexdigest()


def run_case(name, extra=(), decode=False):
    import bench
    import psutil
    c, g = CASES[name], GOLD[name]
    if psutil.virtual_memory().available < 2.6 * c["n"] + (8 << 30):
        pytest.skip("not enough host memory for %s on this box" % name)
    with tempfile.TemporaryDirectory(dir="/dev/shm") as d:
        fin, fout = os.path.join(d, "in.bin"), os.path.join(d, "out.mar")
        mm = np.memmap(fin, dtype=np.uint8, mode="w+", shape=(c["n"],))
        bench.GENS[c["kind"]](c["n"], seed=c["seed"], out=mm)
        mm.flush()
        del mm
        p = subprocess.run([GPU_BIN, "-n", "-vv", "-f", *c["flags"], *extra, "-o", fout, fin], env=dict(os.environ, MRZIP="NOCONFIG"),
                           capture_output=True, text=True)
        assert p.returncode == 0, p.stderr[-3000:]
        assert os.path.getsize(fout) == g["archive_len"]
        ar = H.parse_archive_hashed(fout)
        assert len(ar["chunks"]) == len(g["chunks"])
        for k, (a, b) in enumerate(zip(ar["chunks"], g["chunks"])):
            for key in ("cb", "eof", "size", "s0_len", "s1_len", "s0_sha256", "s1_sha256", "n_blocks"):
                assert a[key] == b[key], "chunk %d: %s differs from the reference's" % (k, key)
        assert ar["trailer"].hex() == g["md5"]
        assert H.sha256_file(fout) == g["archive_sha256"], "archive differs from the reference's"
        err = p.stderr
        for key, want in g["counters_u32"].items():        # the seven -vv counters (rzip.c:1108-1115, printed through (unsigned) casts)
            assert ("%s=%d" % (key, want)) in err, key
        if decode and H.have_ref():
            back = os.path.join(d, "back.bin")
            want = blake2b_file(fin)
            os.unlink(fin)
            subprocess.run([H.REF_BIN, "-d", "-q", "-f", "-o", back, fout], env=dict(os.environ, MRZIP="NOCONFIG"), check=True,
                           capture_output=True)
            assert blake2b_file(back) == want
            os.unlink(back)
            # the product's own way back at full size (host/runzip_fd.c + mrz_runzip_chunk: literal scatter, match waves, CRC32, MD5)
            q = subprocess.run([GPU_BIN, "-d", "-vv", "-f", "-o", back, fout], env=dict(os.environ, MRZIP="NOCONFIG"), capture_output=True, text=True)
            assert q.returncode == 0, q.stderr[-3000:]
            assert blake2b_file(back) == want
            print("mrzip -d:", [l for l in q.stderr.splitlines() if l.startswith("gpu:") or "DeCompression" in l])
    return p.stderr


def test_config1_tar_4gib_archive_equals_reference():
    """configs[1]: 4 GiB tar-like, one chunk, 5-byte distances; the timed configuration of bench.py."""
    run_case("tar4g", decode=True)


def test_vm_4gib_archive_equals_reference():
    run_case("vm4g", decode=True)


def test_config3_random_8gib_archive_equals_reference():
    """configs[3]: 8 GiB incompressible; the sweeps reach mask 2047."""
    assert GOLD["random8g"]["max_sweep_mask"] >= 1023
    run_case("random8g")


def test_config4_dense_4gib_R9_archive_equals_reference():
    """configs[4]'s data at rzip level 9 (max_chain_len 128), 4 GiB."""
    run_case("dense4g_R9")


def test_config4_dense_16gib_R9_archive_equals_reference():
    """configs[4] at its stated size: 16 GiB dense repeats, rzip level 9 (the rzip stage of `-L9`)."""
    run_case("dense16g_R9")


@all_32g
def test_config2_vm_32gib_w41_one_gpu_archive_equals_reference():
    """configs[2]: 32 GiB VM-image-like, `-w 41` -> 8 chunks; chunks in flight side by side, each searched for all 16
    values of victim_round_in, the chain resolved on the host (rzip_fd.c)."""
    err = run_case("vm32g_w41", extra=("-G", "2"))
    assert " 0 chunk(s) searched for every victim_round" not in err


def test_config2_vm_32gib_w10_all_gpus_archive_equals_reference():
    """configs[2] with `-w 10` -> 33 chunks, spread over every GPU of the box (`-g 0,1,..`)."""
    import mrz_py
    n = mrz_py.lib().mrz_device_count()
    run_case("vm32g_w10", extra=("-g", ",".join(str(i) for i in range(max(1, n))), "-G", "4"))


@all_32g
def test_tar_32gib_w41_archive_equals_reference():
    """bench.py's multi-GPU weak-scaling input: 8 tar-like pieces of one `-w 41` chunk each."""
    if "tar32g_w41" not in GOLD:
        pytest.skip("golden vector not minted")
    import mrz_py
    n = mrz_py.lib().mrz_device_count()
    run_case("tar32g_w41", extra=("-g", ",".join(str(i) for i in range(max(1, n))), "-G", "2"))
