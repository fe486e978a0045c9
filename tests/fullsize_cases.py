"""BASELINE.json's configurations at their stated sizes, shared by the golden-vector generator
(tests/golden/make_fullsize_golden.py, run in the build container against the unmodified reference binary) and the
`-m gpu` full-size parity tests.  TEST INFRASTRUCTURE ONLY.

flags are the `mrzip -n` flags that pin the archive layout (chunk size, chunk_bytes, stream bufsize):
  -m is chosen so that ramsize/3 >= chunk size, i.e. the reference maps the chunk whole (rzip.c:974-982: no sliding mmap)
  and stream bufsize covers a chunk (stream.c:878-914: one block per stream per chunk).
"""
GIB = 1 << 30

CASES = {
    # configs[1]: 4 GiB tar-like, single chunk, 5-byte distances
    "tar4g": dict(kind="tar", n=4 * GIB, seed=2, flags=("-m", "130", "-p", "1")),
    # configs[2]'s data kind as one 4 GiB chunk (bench.py --workload vm)
    "vm4g": dict(kind="vm", n=4 * GIB, seed=2, flags=("-m", "130", "-p", "1")),
    # configs[3]: 8 GiB incompressible
    "random8g": dict(kind="random", n=8 * GIB, seed=4, flags=("-m", "260", "-p", "1")),
    # configs[4]: 16 GiB dense repeats, rzip stage alone at rzip level 9 (max_chain_len 128)
    "dense16g_R9": dict(kind="dense", n=16 * GIB, seed=5, flags=("-m", "500", "-p", "1", "-R", "9")),
    # a 4 GiB prefix-sized variant of the same (cheaper to re-check on the GPU box)
    "dense4g_R9": dict(kind="dense", n=4 * GIB, seed=5, flags=("-m", "130", "-p", "1", "-R", "9")),
    # configs[2]: 32 GiB VM-image-like, -w 41 -> 8 chunks of 4,299,161,600 B; -w 10 -> 33 chunks
    "vm32g_w41": dict(kind="vm", n=32 * GIB, seed=3, flags=("-w", "41", "-m", "140", "-p", "1")),
    # the multi-GPU weak-scaling input of bench.py: 8 tar-like pieces of one `-w 41` chunk each (piece k = seed 2 + k)
    "tar32g_w41": dict(kind="tarchunks", n=32 * GIB, seed=2, flags=("-w", "41", "-m", "140", "-p", "1")),
    "vm32g_w10": dict(kind="vm", n=32 * GIB, seed=3, flags=("-w", "10", "-m", "140", "-p", "1")),
}
