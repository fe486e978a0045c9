#!/usr/bin/env python
"""Mint golden vectors for BASELINE.json's configurations AT THEIR STATED SIZES from the UNMODIFIED reference binary
(oracle/_ref/mrzip), in the build container (62 GB RAM, no GPU):

    python tests/golden/make_fullsize_golden.py [case ...]

Inputs come from bench.py's seeded generators (so the GPU box regenerates them bit-identically); per chunk the SHA-256
of each of the two rzip streams is kept, plus the archive's SHA-256, the MD5 trailer, the reference's -vv counters and
its wall time here.  Results are merged into tests/golden/fullsize_vectors.json.
"""
import json
import os
import re
import subprocess
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, ROOT)
import helpers as H  # noqa: E402
import bench  # noqa: E402
from fullsize_cases import CASES  # noqa: E402

OUT = os.path.join(HERE, "fullsize_vectors.json")
SHM = "/dev/shm"
ARC_DIR = os.environ.get("MRZ_GOLD_TMP", "/tmp/mrz_gold")


def input_path(case):
    c = CASES[case]
    return os.path.join(SHM, "mrz_in_%s_%d_%d.bin" % (c["kind"], c["n"] >> 20, c["seed"]))


def make_input(case):
    c = CASES[case]
    path = input_path(case)
    if os.path.exists(path) and os.path.getsize(path) == c["n"]:
        return path
    mm = np.memmap(path + ".tmp", dtype=np.uint8, mode="w+", shape=(c["n"],))
    bench.GENS[c["kind"]](c["n"], seed=c["seed"], out=mm)
    mm.flush()
    del mm
    os.rename(path + ".tmp", path)
    return path


def run_case(case):
    c = CASES[case]
    os.makedirs(ARC_DIR, exist_ok=True)
    fin = make_input(case)
    fout = os.path.join(ARC_DIR, case + ".mar")
    env = dict(os.environ, MRZIP="NOCONFIG")
    t0 = time.perf_counter()
    p = subprocess.run([H.REF_BIN, "-n", "-vv", "-f", *c["flags"], "-o", fout, fin], env=env, capture_output=True, text=True)
    wall = time.perf_counter() - t0
    if p.returncode:
        raise SystemExit("reference failed on %s: %s" % (case, p.stderr[-2000:]))
    err = p.stderr.replace("\r", "\n")
    counters = {}
    for k in ("matches", "match_bytes", "literals", "literal_bytes", "true_tag_positives", "false_tag_positives", "inserts"):
        m = re.search(r"\b%s=(\d+)" % k, err)
        if m:
            counters[k] = int(m.group(1))
    sweeps = [int(x) for x in re.findall(r"Starting sweep for mask (\d+)", err)]
    ar = H.parse_archive_hashed(fout)
    rec = dict(kind=c["kind"], n=c["n"], seed=c["seed"], flags=list(c["flags"]), archive_len=os.path.getsize(fout),
               archive_sha256=H.sha256_file(fout), md5=ar["trailer"].hex(), magic=ar["magic"], chunks=ar["chunks"],
               counters_u32=counters, max_sweep_mask=max(sweeps) if sweeps else 0,
               sliding_mmap="Enabling sliding mmap" in err,
               reference_wall_s_build_container=round(wall, 2), reference_nproc=os.cpu_count())
    os.unlink(fout)
    return rec


def main():
    assert H.have_ref(), "build the reference first: make -C oracle ref"
    names = sys.argv[1:] or list(CASES)
    for case in names:
        rec = run_case(case)
        import fcntl
        with open(OUT + ".lock", "w") as lk:      # several cases may be minted side by side
            fcntl.flock(lk, fcntl.LOCK_EX)
            data = {}
            if os.path.exists(OUT):
                with open(OUT) as f:
                    data = json.load(f)
            data[case] = rec
            with open(OUT + ".tmp", "w") as f:
                json.dump(data, f, indent=1, sort_keys=True)
            os.rename(OUT + ".tmp", OUT)
        print(case, "n=%d" % rec["n"], "chunks=%d" % len(rec["chunks"]), "wall=%.1fs" % rec["reference_wall_s_build_container"],
              rec["counters_u32"], flush=True)


if __name__ == "__main__":
    main()
