"""Property test on the emulator build of the kernels (CPU only): made-up inputs of every texture the automaton treats
differently - random bytes, short periods (self-overlapping matches), copied spans, runs - at random sizes around the
31-byte minimum match and around batch / tile boundaries, random level and victim_round_in: streams, counters and the
carried state equal the oracle's, and the device replay gives the input back (tests/parity_common.check)."""
import os

import numpy as np
from hypothesis import HealthCheck, given, settings, strategies as st

from parity_common import check


@st.composite
def textured(draw):
    n = draw(st.one_of(st.integers(0, 200), st.integers(8000, 8400), st.integers(20000, 70000)))
    seed = draw(st.integers(0, 2 ** 31 - 1))
    rng = np.random.default_rng(seed)
    buf = rng.integers(0, 256, size=n, dtype=np.uint8)
    pieces = draw(st.integers(0, 6))
    for _ in range(pieces):
        if n < 64:
            break
        kind = draw(st.sampled_from(["period", "copy", "run", "alphabet"]))
        a = int(rng.integers(0, n - 32))
        ln = int(rng.integers(32, min(n - a, 5000) + 1))
        if kind == "period":
            p = int(rng.integers(1, 40))
            buf[a:a + ln] = np.resize(buf[a:a + p].copy(), ln)
        elif kind == "copy" and a > 64:
            src = int(rng.integers(0, a - 32))
            m = min(ln, a - src)
            buf[a:a + m] = buf[src:src + m]
        elif kind == "run":
            buf[a:a + ln] = buf[a]
        else:
            buf[a:a + ln] = rng.integers(97, 101, size=ln, dtype=np.uint8)
    return buf


@settings(max_examples=int(os.environ.get("MRZ_PROPERTY_EXAMPLES", "200")), deadline=None, derandomize=True, suppress_health_check=list(HealthCheck))
@given(data=textured(), level=st.integers(0, 9), vr=st.integers(0, 127))
def test_textured_inputs_match_the_oracle_and_round_trip(sim_ctx, data, level, vr):
    import mrz_py
    check(sim_ctx, data, level=level, vr_in=vr % mrz_py.CHAIN[level])
