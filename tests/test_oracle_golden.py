"""The CPU oracle (oracle/rzip_oracle.c) against the golden vectors minted from the unmodified reference binary
(tests/golden/make_golden.py) — no reference needed at test time — and, when oracle/_ref/mrzip is present, against
the reference run live.  This is what pins the oracle (the reference has no tests of its own)."""
import hashlib
import os
import tempfile

import pytest

import helpers as H

GOLD = H.load_golden()
CASES = H.golden_cases()


def level_of(flags):
    lvl = 7
    f = list(flags)
    for k in ("-L", "-R"):   # -R wins over -L for the rzip level (main.c:575)
        if k in f:
            lvl = int(f[f.index(k) + 1])
    return lvl


@pytest.mark.parametrize("name", sorted(CASES))
def test_oracle_matches_golden(name):
    data, flags = CASES[name]
    g = GOLD[name]
    assert len(g["chunks"]) == 1 and g["n"] == len(H.as_u8(data))
    ch = g["chunks"][0]
    r = H.oracle_chunk(data, level=level_of(flags), cb=ch["cb"])
    assert len(r["s0"]) == ch["s0_len"] and len(r["s1"]) == ch["s1_len"]
    assert hashlib.sha256(r["s0"]).hexdigest() == ch["s0_sha256"]
    assert hashlib.sha256(r["s1"]).hexdigest() == ch["s1_sha256"]
    assert hashlib.md5(H.as_u8(data).tobytes()).hexdigest() == g["md5"]
    assert H.decode_tokens(r["s0"], r["s1"], ch["cb"]) == H.as_u8(data).tobytes()


def test_golden_archive_layout():
    """The stored small archives parse with the layout of SURVEY.md §8(a19/a20) and carry the survey's bytes."""
    raw = bytes.fromhex(GOLD["survey650"]["archive_hex"])
    assert raw[:6] == b"MRZI\x00\x09" and len(raw) == 536
    ar = H.parse_archive(raw)
    assert ar["magic"]["st_size"] == 650 and ar["magic"]["hash_code"] == 1 and ar["magic"]["levels"] == 0x77
    c = ar["chunks"][0]
    assert c["cb"] == 2 and c["eof"] == 1 and c["size"] == 4096
    assert c["s0"].hex() == "009001" "01c8009001" "003200" "000000" "d42c7be1"
    assert bytes.fromhex(GOLD["pattern0"]["archive_hex"])[-16:].hex() == hashlib.md5(b"").hexdigest()


def test_hash_index_first_entry():
    assert int(H.hash_index()[0]) == 0x6b8b771c23c6   # SURVEY.md TL;DR 2: unseeded glibc random()


@pytest.mark.skipif(not H.have_ref(), reason="oracle/_ref/mrzip not built (no /root/reference here)")
@pytest.mark.parametrize("case", ["two_chunks_victim_carry", "second_file_hash_index"])
def test_oracle_vs_live_reference(case):
    with tempfile.TemporaryDirectory() as d:
        if case == "two_chunks_victim_carry":
            # -m 1 => ramsize 100 MB => chunks of 2/3 of that; level 1 motif data: victim_round crosses the chunk boundary
            data = H.gen_motif(72 << 20, 60000, seed=11)
            fin, fout = os.path.join(d, "in.bin"), os.path.join(d, "o.mar")
            data.tofile(fin)
            H.run_reference(fin, fout, ("-m", "1", "-R", "1"))
            ar = H.parse_archive(open(fout, "rb").read())
            assert len(ar["chunks"]) == 2
            vr, off = 0, 0
            for c in ar["chunks"]:
                n = c["size"]
                r = H.oracle_chunk(data[off:off + n], level=1, cb=c["cb"], vr_in=vr)
                assert r["s0"] == c["s0"] and r["s1"] == c["s1"]
                vr, off = r["vr_out"], off + n
            assert off == data.size
        else:
            # two files on one command line: the second uses random() draws 512..1023 (main.c:651 loop)
            a, b = H.gen_random_repeats(1 << 20, 10, lo=1024, hi=8192), H.gen_motif(1 << 20, 1000, seed=3)
            fa, fb = os.path.join(d, "a.bin"), os.path.join(d, "b.bin")
            a.tofile(fa); b.tofile(fb)
            import subprocess
            subprocess.run([H.REF_BIN, "-n", "-q", "-f", "-O", d, fa, fb], env=dict(os.environ, MRZIP="NOCONFIG"),
                           check=True, capture_output=True)
            arb = H.parse_archive(open(fb + ".lrz", "rb").read())
            ra = H.oracle_chunk(a)
            rb = H.oracle_chunk(b, hidx=H.hash_index(512), vr_in=ra["vr_out"])
            assert rb["s0"] == arb["chunks"][0]["s0"] and rb["s1"] == arb["chunks"][0]["s1"]
