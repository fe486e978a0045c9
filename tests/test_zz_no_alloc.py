"""mrz_crc32_noalloc (include/mrz_gpu.h; no counterpart in the reference):
the chunk CRC32 with its scratch in a fixed file-scope device array - no cudaMalloc.  Called through the C-ABI on two
inputs in a row and compared with the existing entry point (the crc32 that mrz_rzip_chunk reports and writes into stream
0) and with zlib; sizes beyond the fixed maximum are rejected.  Collected last on purpose."""
import ctypes as C
import zlib

import numpy as np
import pytest

import helpers as H

pytestmark = pytest.mark.gpu


def test_crc32_noalloc_matches_existing_entry_point_twice():
    import torch
    import mrz_py
    L = mrz_py.lib()
    hidx = mrz_py.hash_index()
    out = torch.zeros(1, dtype=torch.int32, device="cuda:0")
    ctx = mrz_py.Context(0, 64 << 20)
    try:
        # second call smaller than the first: whatever the first left in the shared scratch must not matter
        for n, seed in ((48 * 1048576 + 12345, 5), (3 * 1048576 + 16384 * 7 + 77, 6), (0, 7), (100, 8)):
            data = H.gen_random(n, seed=seed) if n else np.zeros(0, dtype=np.uint8)
            res, _, _ = ctx.rzip_chunk(data, mrz_py.make_params(7, mrz_py.chunk_bytes(n), hidx, 0))     # the existing entry point
            dev = torch.from_numpy(data.copy()).to("cuda:0") if n else torch.zeros(16, dtype=torch.uint8, device="cuda:0")
            assert dev.data_ptr() % 16 == 0
            out.fill_(-1)
            rc = L.mrz_crc32_noalloc(0, C.c_void_p(dev.data_ptr()), n, C.c_void_p(out.data_ptr()), None)
            assert rc == 0, L.mrz_noalloc_last_error()
            torch.cuda.synchronize()
            got = int(out.item()) & 0xFFFFFFFF
            assert got == res.crc32 == (zlib.crc32(data.tobytes()) & 0xFFFFFFFF), n      # integer work: bit-exact
    finally:
        ctx.close()


def test_crc32_noalloc_rejects_sizes_beyond_the_fixed_scratch():
    import torch
    import mrz_py
    L = mrz_py.lib()
    dev = torch.zeros(64, dtype=torch.uint8, device="cuda:0")
    out = torch.zeros(1, dtype=torch.int32, device="cuda:0")
    rc = L.mrz_crc32_noalloc(0, C.c_void_p(dev.data_ptr()), (1 << 30) + 1, C.c_void_p(out.data_ptr()), None)
    assert rc == -5 and b"exceed" in L.mrz_noalloc_last_error()
    assert L.mrz_crc32_noalloc(0, C.c_void_p(dev.data_ptr() + 1), 10, C.c_void_p(out.data_ptr()), None) == -3
