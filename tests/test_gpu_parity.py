"""GPU parity: the CUDA path through the C-ABI vs the CPU oracle, bit-exact (streams, counters, victim carry)."""
import numpy as np
import pytest

import helpers as H

pytestmark = pytest.mark.gpu

from parity_common import check


@pytest.mark.parametrize("n", [0, 1, 30, 31, 32, 33, 62, 63, 64, 650, 8191, 8192, 8193, 8222, 8223, 8224, 100003])
def test_tiny_and_ragged(gpu_ctx, n):
    seq = (b"abcdefghijklmnopqrstuvwxyz012345" * (n // 32 + 1))[:n]
    check(gpu_ctx, seq)
    check(gpu_ctx, H.gen_random(n, seed=n + 1) if n else b"")


def test_survey_650_byte_example(gpu_ctx):
    import random
    random.seed(1)
    A = bytes(random.getrandbits(8) for _ in range(300))
    B = bytes(random.getrandbits(8) for _ in range(100))
    Cc = bytes(random.getrandbits(8) for _ in range(50))
    res, ref = check(gpu_ctx, A + B + A[:200] + Cc)
    assert res.matches == 1 and res.s0_len == 18 and res.s1_len == 450


def test_random_with_repeats(gpu_ctx):
    res, _ = check(gpu_ctx, H.gen_random_repeats(8 << 20, 60))
    assert res.matches > 50


def test_motif_level1_victims_and_sweeps(gpu_ctx):
    res, ref = check(gpu_ctx, H.gen_motif(3 << 20, 3000), level=1)
    assert ref["stats"]["victim_events"] > 0 and ref["stats"]["sweeps_wrapped"] > 0


@pytest.mark.parametrize("vr", [0, 1])
def test_victim_round_carry(gpu_ctx, vr):
    check(gpu_ctx, H.gen_motif(3 << 20, 3000), level=1, vr_in=vr)


def test_vm_like(gpu_ctx):
    check(gpu_ctx, H.gen_vm_like(16 << 20))


def test_dense(gpu_ctx):
    check(gpu_ctx, H.gen_dense(16 << 20))


def test_dense_level9(gpu_ctx):
    check(gpu_ctx, H.gen_dense(8 << 20), level=9)


def test_random_sweeps(gpu_ctx):
    res, ref = check(gpu_ctx, H.gen_random(32 << 20))
    assert ref["stats"]["sweeps_wrapped"] >= 2


def test_tar_like(gpu_ctx):
    check(gpu_ctx, H.gen_tar_like(16 << 20))


@pytest.mark.parametrize("level", [0, 1, 2, 3, 4, 5, 6, 8, 9])
def test_levels(gpu_ctx, level):
    check(gpu_ctx, H.gen_vm_like(2 << 20, seed=level + 10), level=level)
    check(gpu_ctx, H.gen_motif(1 << 20, 2000, seed=level), level=level)


def test_constant_and_periodic(gpu_ctx):
    check(gpu_ctx, np.zeros(1 << 20, dtype=np.uint8))
    check(gpu_ctx, np.full(1 << 20, 255, dtype=np.uint8))
    check(gpu_ctx, np.tile(np.arange(7, dtype=np.uint8), (1 << 20) // 7))
    check(gpu_ctx, np.tile(H.gen_random(4099, seed=9), 300))


def test_wide_chunk_bytes(gpu_ctx):
    # distances written with 5 bytes, as for chunks >= 4 GiB (rzip.c:1006-1008)
    check(gpu_ctx, H.gen_random_repeats(2 << 20, 40, lo=1024, hi=65536), cb=5)


def test_tags_match_full_tag(gpu_ctx):
    data = H.gen_random(100000, seed=77)
    hidx = H.hash_index()
    gpu_ctx.upload(data)
    tags = gpu_ctx.debug_tags(data.size, hidx, data.size - 30)
    h = hidx[data]
    x = np.concatenate([[0], np.bitwise_xor.accumulate(h)])
    want = x[31:] ^ x[:-31]
    assert np.array_equal(tags, want)


def test_crc_matches_zlib(gpu_ctx):
    import zlib
    import mrz_py
    hidx = H.hash_index()
    for n in (0, 5, 16384, 16385, 524288, 524288 + 16384 * 3 + 77, 3 << 20):
        data = H.gen_random(n, seed=n % 97 + 1)
        res, _, _ = gpu_ctx.rzip_chunk(data, mrz_py.make_params(7, 4, hidx))
        assert res.crc32 == zlib.crc32(data.tobytes()), n


def _variant_case(level, size, reps=2, **config):
    """Every victim_round_in of one resident chunk through mrz_run_variants (ONE launch, a cluster per variant), checked
    against the plain entry point (all variants) and the oracle (a few), then emitted in arbitrary order and compared
    byte for byte."""
    import mrz_py
    chain = mrz_py.CHAIN[level]
    data = H.as_u8(H.gen_motif(size, 3000 * level) if level != 7 else H.gen_vm_like(size, seed=47))
    hidx, cb = H.hash_index(), H.chunk_bytes(data.size)
    vs = mrz_py.VariantSet(0, data.size, chain, **config)
    try:
        vs.upload(data)
        for _ in range(reps):                       # contexts are reusable
            res = vs.run(data.size, mrz_py.make_params(level, cb, hidx, 0))
        check_vs = sorted(set([0, 1, chain // 2, chain - 1]))
        refs = {v: H.oracle_chunk(data, level=level, cb=cb, vr_in=v, hidx=hidx) for v in check_vs}
        assert len({r["s0"] for r in refs.values()}) > 1, "the case must depend on victim_round_in"
        seq = mrz_py.Context(0, data.size)          # the same variants one after the other through the plain entry point
        try:
            seq.upload(data)
            for v in range(chain):
                r1 = seq.run_resident(data.size, mrz_py.make_params(level, cb, hidx, v))
                for k in ("victim_round_out", "victim_events", "inserts", "tag_hits", "tag_misses", "n_ops", "final_min_mask"):
                    assert getattr(res[v], k) == getattr(r1, k), (v, k)
        finally:
            seq.close()
        for v in reversed(check_vs):
            assert res[v].victim_round_out == refs[v]["vr_out"]
            assert res[v].victim_events == refs[v]["stats"]["victim_events"] > 0
            r, s0, s1 = vs.emit(v)
            assert s0.tobytes() == refs[v]["s0"] and s1.tobytes() == refs[v]["s1"], v
            assert r.matches == refs[v]["stats"]["matches"] and r.literal_bytes == refs[v]["stats"]["literal_bytes"]
    finally:
        vs.close()


CASES = [(7, 24 << 20), (1, 3 << 20), (8, 2 << 20)]
CASE_IDS = ["L7-vm-16var", "L1-motif-2var", "L8-motif-32var"]


@pytest.mark.parametrize("level,size", CASES, ids=CASE_IDS)
def test_variant_chunk(level, size):
    """mrz_run_variants (include/mrz_gpu.h): all variants of one resident chunk side by side in one launch, in process."""
    _variant_case(level, size, reps=2, poison=1)


@pytest.mark.parametrize("config", [dict(cluster_size=1, threads=512), dict(cluster_size=2, threads=256), dict(cluster_size=8, threads=128),
                                    dict(cluster_size=16, threads=64), dict(cluster_size=16, threads=512),
                                    dict(cluster_size=4, threads=256, filter_log2_words=12)],
                         ids=["1x512", "2x256", "8x128", "16x64", "16x512", "4x256-tiny-filter"])
def test_cluster_geometries(config):
    """The same automaton under other launch geometries (mrz_config): one CTA, small and the largest clusters; a tiny
    tag filter makes almost every op take the exact executor."""
    import mrz_py
    ctx = mrz_py.Context(0, 1 << 26, poison=1, **config)
    try:
        check(ctx, H.gen_random_repeats(8 << 20, 60))
        check(ctx, H.gen_motif(3 << 20, 3000), level=1)
        check(ctx, H.gen_vm_like(16 << 20))
        check(ctx, H.gen_random(32 << 20), level=4)
        check(ctx, H.gen_dense(8 << 20), level=9)
    finally:
        ctx.close()


@pytest.mark.parametrize("kind", ["glibc", "wide"])
def test_tagmap_matches_prefix_xor(gpu_ctx, kind):
    """tagmap.cu against an independent restatement; "wide": every tag has >= 32 trailing ones (the 64-bit path)."""
    from parity_common import check_tagmap, wide_hash_index
    data = H.gen_random((3 << 20) + 12345, seed=78)
    want = check_tagmap(gpu_ctx, data, H.hash_index() if kind == "glibc" else wide_hash_index())
    if kind == "wide":
        assert want[: data.size - 30].min() >= 32


def test_runzip_rejects_corruption(gpu_ctx):
    """The device token replay refuses corrupt streams and a wrong checksum (every parity case above also round-trips
    through it: parity_common.check)."""
    from parity_common import check_runzip_rejects_corruption
    check_runzip_rejects_corruption(gpu_ctx)


def test_runzip_synthetic_streams(gpu_ctx):
    """Made-up token streams with chained and self-overlapping matches: the wave rule of mrz_runzip_chunk."""
    from parity_common import check_runzip_synthetic
    check_runzip_synthetic(gpu_ctx, seeds=(1, 2, 3, 4, 5), n_tokens=20000)
