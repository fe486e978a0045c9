"""The host's positional stream writer (several chunks of one archive written at the same time) against its sequential
form: same bytes, predicted chunk size == written size, streams several blocks long.  CPU only."""
import filecmp
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "modern-rzip_b200", "host")


def test_positional_writer_matches_sequential(tmp_path):
    exe = str(tmp_path / "stream_writer_test")
    subprocess.check_call(["gcc", "-O2", "-o", exe, os.path.join(ROOT, "tests", "host", "stream_writer_test.c"),
                           os.path.join(HOST, "stream_out.c"), os.path.join(HOST, "util.c"), os.path.join(HOST, "md5.c"),
                           "-I" + os.path.join(ROOT, "include")])
    # literal runs per chunk: 1 (both streams in one block), 170 (stream 1: two blocks of 10 MiB), 330 (three), 0 (empty)
    out = subprocess.check_output([exe, str(tmp_path), "1", "170", "330", "0", "161"], text=True)
    assert out.startswith("ok ")
    assert int(out.split()[1]) == os.path.getsize(tmp_path / "seq.bin")
    assert filecmp.cmp(tmp_path / "seq.bin", tmp_path / "pos.bin", shallow=False)
