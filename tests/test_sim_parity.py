"""CPU-only parity: the product's kernel sources (modern-rzip_b200/csrc/*.cu) compiled for the fiber emulator in
tests/cusim and driven through the same C-ABI, against the oracle.  Catches logic errors in the kernels without a
GPU; the `-m gpu` tests in test_gpu_parity.py remain the parity tests proper."""
import numpy as np
import pytest

import helpers as H
from parity_common import check


@pytest.mark.parametrize("n", [0, 1, 31, 32, 64, 650, 8193, 8223, 100003])
def test_tiny_and_ragged(sim_ctx, n):
    seq = (b"abcdefghijklmnopqrstuvwxyz012345" * (n // 32 + 1))[:n]
    check(sim_ctx, seq)
    check(sim_ctx, H.gen_random(n, seed=n + 1) if n else b"")


def test_random_with_repeats(sim_ctx):
    res, _ = check(sim_ctx, H.gen_random_repeats(4 << 20, 40, lo=2048, hi=65536))
    assert res.matches > 30


@pytest.mark.parametrize("vr", [0, 1])
def test_motif_level1_victims_and_sweeps(sim_ctx, vr):
    res, ref = check(sim_ctx, H.gen_motif(3 << 20, 3000), level=1, vr_in=vr)
    assert ref["stats"]["victim_events"] > 0 and ref["stats"]["sweeps_wrapped"] > 0


def test_vm_like(sim_ctx):
    check(sim_ctx, H.gen_vm_like(4 << 20))


def test_dense_level9(sim_ctx):
    check(sim_ctx, H.gen_dense(4 << 20), level=9)


def test_tar_like(sim_ctx):
    check(sim_ctx, H.gen_tar_like(4 << 20))


def test_sweeps_small_table(sim_ctx):
    res, ref = check(sim_ctx, H.gen_random(8 << 20, seed=21), level=2)
    assert ref["stats"]["sweeps_wrapped"] >= 1


@pytest.mark.parametrize("level", [0, 1, 3, 5, 6, 8])
def test_levels(sim_ctx, level):
    check(sim_ctx, H.gen_vm_like(1 << 20, seed=level + 10), level=level)
    check(sim_ctx, H.gen_motif(1 << 20, 2000, seed=level), level=level)


def test_degenerate(sim_ctx):
    check(sim_ctx, np.zeros(1 << 18, dtype=np.uint8))
    check(sim_ctx, np.full(1 << 18, 255, dtype=np.uint8))
    check(sim_ctx, np.tile(np.arange(7, dtype=np.uint8), (1 << 18) // 7))
    check(sim_ctx, np.tile(H.gen_random(4099, seed=9), 100))


@pytest.mark.parametrize("kind", ["glibc", "wide"])
def test_tagmap_matches_prefix_xor(sim_ctx, kind):
    """tagmap.cu on the emulator against an independent restatement; "wide": every tag has >= 32 trailing ones."""
    from parity_common import check_tagmap, wide_hash_index
    data = H.gen_random((1 << 18) + 345, seed=78)
    want = check_tagmap(sim_ctx, data, H.hash_index() if kind == "glibc" else wide_hash_index())
    if kind == "wide":
        assert want[: data.size - 30].min() >= 32


def test_runzip_rejects_corruption(sim_ctx):
    from parity_common import check_runzip_rejects_corruption
    check_runzip_rejects_corruption(sim_ctx)


def test_runzip_synthetic_streams(sim_ctx):
    """Made-up token streams with chained and self-overlapping matches: the wave rule of mrz_runzip_chunk."""
    from parity_common import check_runzip_synthetic
    check_runzip_synthetic(sim_ctx)


def _sim_ctx(**config):
    """A separate emulator context with its own launch geometry (mrz_config)."""
    import os
    import mrz_py
    lib = mrz_py.bind(os.path.join(os.path.dirname(os.path.abspath(__file__)), "cusim", "libmrz_sim.so"))
    return mrz_py.Context(0, 1 << 24, library=lib, poison=1, **config)


@pytest.mark.parametrize("config", [dict(cluster_size=1, threads=128), dict(cluster_size=4, threads=32), dict(cluster_size=8, threads=32),
                                    dict(cluster_size=2, threads=64, filter_log2_words=10)],
                         ids=["1x128", "4x32", "8x32", "2x64-tiny-filter"])
def test_other_cluster_geometries(sim_ctx, config):
    """One CTA, and clusters of 4 and 8 CTAs, on inputs with matches, sweeps, victims and twins; a tiny tag filter
    (many false positives: every such op takes the exact executor)."""
    ctx = _sim_ctx(**config)
    k = 4 if "filter_log2_words" in config else 1      # with the tiny filter every op is serial: smaller inputs
    try:
        check(ctx, H.gen_random_repeats((2 << 20) // k, 30, lo=2048, hi=65536))
        check(ctx, H.gen_motif((1 << 20) // k, 2000), level=1)
        check(ctx, H.gen_random((4 << 20) // k, seed=21), level=2)
        check(ctx, H.gen_vm_like((1 << 20) // k, seed=13), level=6)
    finally:
        ctx.close()


@pytest.mark.parametrize("level,gen", [(1, lambda: H.gen_motif(1 << 20, 2000)), (6, lambda: H.gen_vm_like(2 << 20, seed=16))],
                         ids=["level1-motif", "level6-vm"])
def test_variant_parallel_chunk(sim_ctx, level, gen):
    """mrz_run_variants: every victim_round_in of a resident chunk at once (shared chunk bytes, own tables); each
    variant's counters, carry and — once chosen with mrz_run_emit — streams equal the oracle run with that input."""
    import mrz_py
    chain = [1, 2, 2, 2, 3, 4, 6, 16, 32, 128][level]
    data = H.as_u8(gen())
    hidx, cb = H.hash_index(), H.chunk_bytes(data.size)
    vs = mrz_py.VariantSet(0, data.size, chain, library=sim_ctx.L)
    try:
        vs.upload(data)
        res = vs.run(data.size, mrz_py.make_params(level, cb, hidx, 0))
        refs = [H.oracle_chunk(data, level=level, cb=cb, vr_in=v, hidx=hidx) for v in range(chain)]
        assert len({r["s0"] for r in refs}) > 1, "the case must depend on victim_round_in"
        for v in reversed(range(chain)):          # emission order is free: tokens of every variant stay resident
            assert res[v].victim_round_out == refs[v]["vr_out"]
            assert res[v].victim_events == refs[v]["stats"]["victim_events"]
            assert res[v].inserts == refs[v]["stats"]["inserts"]
            r, s0, s1 = vs.emit(v)
            assert s0.tobytes() == refs[v]["s0"] and s1.tobytes() == refs[v]["s1"]
            assert r.matches == refs[v]["stats"]["matches"] and r.victim_round_out == refs[v]["vr_out"]
    finally:
        vs.close()
