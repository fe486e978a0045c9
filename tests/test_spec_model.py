"""The speculative-batch algorithm the GPU automaton uses (oracle/spec_model.c) must be exactly
equivalent to the sequential restatement for every batch size: streams, counters, victim carry."""
import numpy as np
import pytest

import helpers as H

KEYS = ("matches", "match_bytes", "literals", "literal_bytes", "tag_hits", "tag_misses", "inserts", "lookups",
        "final_min_mask", "displaced")


def same(data, level=7, vr=0, batches=(1, 33, 1024)):
    ref = H.oracle_chunk(data, level=level, vr_in=vr)
    for b in batches:
        r = H.spec_model_chunk(data, level=level, vr_in=vr, batch=b)
        assert r["s0"] == ref["s0"] and r["s1"] == ref["s1"], (b, level)
        assert r["vr_out"] == ref["vr_out"]
        for k in KEYS:
            assert r["stats"][k] == ref["stats"][k], (k, b)
    return r


@pytest.mark.parametrize("n", [0, 31, 32, 64, 650, 8193, 100003])
def test_small(n):
    same((b"abcdefghijklmnopqrstuvwxyz012345" * (n // 32 + 1))[:n])
    same(H.gen_random(n, seed=n + 3) if n else b"")


def test_random_repeats():
    r = same(H.gen_random_repeats(6 << 20, 50), batches=(64, 1024, 4096))
    assert r["spec"]["planned_commits"] > 50 * r["spec"]["serial_ops"]


def test_motif_victims_level1():
    for vr in (0, 1):
        same(H.gen_motif(3 << 20, 3000), level=1, vr=vr, batches=(256, 1024))


def test_vm_like_and_tar_like():
    same(H.gen_vm_like(8 << 20), batches=(1024,))
    same(H.gen_tar_like(8 << 20), batches=(1024,))


def test_dense_level9():
    same(H.gen_dense(4 << 20), level=9, batches=(128, 1024))


def test_sweeps_small_table():
    # level 2 = 4 MiB table: several sweep wraps on 8 MiB of random data
    r = same(H.gen_random(8 << 20, seed=21), level=2, batches=(512, 2048))
    assert r["stats"]["final_min_mask"] >= 31


@pytest.mark.parametrize("level", [0, 1, 3, 5, 6, 8])
def test_levels(level):
    same(H.gen_vm_like(1 << 20, seed=level + 10), level=level, batches=(1024,))
    same(H.gen_motif(1 << 20, 2000, seed=level), level=level, batches=(1024,))


def test_degenerate():
    same(np.zeros(1 << 18, dtype=np.uint8), batches=(1024,))
    same(np.full(1 << 18, 255, dtype=np.uint8), batches=(1024,))
    same(np.tile(np.arange(7, dtype=np.uint8), (1 << 18) // 7), batches=(1024,))
    same(np.tile(H.gen_random(4099, seed=9), 100), batches=(1024,))
