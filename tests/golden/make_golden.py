#!/usr/bin/env python
"""Mint golden vectors from the UNMODIFIED reference binary (oracle/_ref/mrzip, built from /root/reference by
oracle/Makefile.ref).  The reference ships no tests or fixtures of its own, so these are the
pins for the oracle restatement and for the CUDA path.  Run in the build container:

    python tests/golden/make_golden.py

Small cases keep the whole archive (hex); larger ones keep SHA-256 of each chunk's two streams plus the archive's.
Inputs are regenerated from seeds by tests/helpers.py, so only the expected outputs are stored.
"""
import hashlib
import json
import os
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import helpers as H  # noqa: E402

CASES = H.golden_cases()


def main():
    assert H.have_ref(), "build the reference first: make -C oracle ref"
    out = {}
    with tempfile.TemporaryDirectory() as d:
        for name, (data, flags) in CASES.items():
            fin, fout = os.path.join(d, "in.bin"), os.path.join(d, "out.mar")
            H.as_u8(data).tofile(fin)
            H.run_reference(fin, fout, flags)
            raw = open(fout, "rb").read()
            ar = H.parse_archive(raw)
            rec = {"flags": list(flags), "n": int(H.as_u8(data).size), "archive_sha256": hashlib.sha256(raw).hexdigest(),
                   "archive_len": len(raw), "md5": ar["trailer"].hex(),
                   "chunks": [{"cb": c["cb"], "eof": c["eof"], "size": c["size"], "s0_len": len(c["s0"]), "s1_len": len(c["s1"]),
                               "s0_sha256": hashlib.sha256(c["s0"]).hexdigest(), "s1_sha256": hashlib.sha256(c["s1"]).hexdigest(),
                               "blocks": [list(b) for b in c["blocks"]]} for c in ar["chunks"]]}
            if len(raw) <= 4096:
                rec["archive_hex"] = raw.hex()
            out[name] = rec
            print(name, rec["n"], "->", len(raw), "bytes,", len(ar["chunks"]), "chunk(s)")
    with open(os.path.join(HERE, "reference_vectors.json"), "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
