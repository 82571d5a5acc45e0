"""random.h:9-15 (`random_double` over one shared mt19937) is replaced by Philox4x32-10 (device/pt_rng.h).  Known-answer
tests of the generator itself against the vectors published with Random123 (Salmon et al., SC'11: kat_vectors,
philox4x32 10 rounds), of the uniform mapping and of the keying: (seed; pixel, 56-bit global sample, bounce / stage / k,
index) must never alias over the ranges b200pt_render_start accepts."""
import ctypes as C

import numpy as np
import pytest

from oracle_lib import PORT_LIB

# counter[4], key[2] -> output[4]  (Random123 kat_vectors: "philox4x32 10")
KAT = [
    ((0x00000000, 0x00000000, 0x00000000, 0x00000000), (0x00000000, 0x00000000),
     (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
    ((0xffffffff, 0xffffffff, 0xffffffff, 0xffffffff), (0xffffffff, 0xffffffff),
     (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
    ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0),
     (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1)),
]


def _kat_arrays():
    inp = np.array([list(c) + list(k) for c, k, _ in KAT], dtype=np.uint32)
    exp = np.array([list(o) for _, _, o in KAT], dtype=np.uint32)
    return inp, exp


def _port():
    lib = C.CDLL(PORT_LIB)
    lib.pt_oracle_philox.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p]
    lib.pt_oracle_rng4.argtypes = [C.c_uint64, C.c_uint32, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_uint32, C.c_void_p,
                                   C.c_void_p]
    return lib


def _rng4(lib, seed, pixel, sample, bounce, stage, k, idx):
    out = np.zeros(4, dtype=np.float32)
    sw = C.c_uint32()
    lib.pt_oracle_rng4(seed, pixel, sample, bounce, stage, k, idx, out.ctypes.data, C.byref(sw))
    return out, sw.value


def test_philox_known_answers_cpu():
    inp, exp = _kat_arrays()
    out = np.zeros_like(exp)
    _port().pt_oracle_philox(inp.ctypes.data, len(inp), out.ctypes.data)
    assert np.array_equal(out, exp), [hex(x) for x in out.ravel()]


@pytest.mark.gpu
def test_philox_known_answers_gpu():
    """The same vectors through the device code, plus 10^5 random counters against the host compilation of pt_rng.h."""
    import pathtrace_b200 as pt
    from pathtrace_b200._capi import check
    inp, exp = _kat_arrays()
    rnd = np.random.default_rng(3).integers(0, 2 ** 32, size=(100000, 6), dtype=np.uint64).astype(np.uint32)
    allin = np.ascontiguousarray(np.concatenate([inp, rnd]))
    got = np.zeros((len(allin), 4), dtype=np.uint32)
    check(pt.lib().b200pt_philox4x32_10(0, allin.ctypes.data, len(allin), got.ctypes.data))
    assert np.array_equal(got[:len(exp)], exp)
    host = np.zeros_like(got)
    _port().pt_oracle_philox(allin.ctypes.data, len(allin), host.ctypes.data)
    assert np.array_equal(got, host)


def test_uniforms_are_24_bit_fractions_in_unit_interval():
    lib = _port()
    vals = np.concatenate([_rng4(lib, 7, p, s, 0, 0, 0, 0)[0] for p in range(64) for s in range(16)])
    assert vals.min() >= 0.0 and vals.max() < 1.0
    assert np.array_equal(vals * 16777216.0, np.round(vals * 16777216.0))  # k / 2^24 exactly
    assert abs(vals.mean() - 0.5) < 0.02


def test_stage_words_do_not_alias():
    """Every (bounce, stage, k) the render can ask for maps to its own stage word: max_bounces <= 4096,
    light_samples <= 1024 (b200pt_render_start), stages 0..6 (pt_rng.h)."""
    lib = _port()
    seen = {}
    for bounce in (0, 1, 2, 255, 256, 4095, 4096):
        for stage in range(7):
            for k in (0, 1, 3, 255, 256, 511, 512, 767, 768, 1023):
                sw = _rng4(lib, 0, 0, 0, bounce, stage, k, 0)[1]
                assert sw not in seen, (bounce, stage, k, seen[sw])
                seen[sw] = (bounce, stage, k)


def test_high_bits_of_the_global_sample_index_reach_the_counter():
    """config.sample_offset is 64 bits wide; two samples 2^32 (or 2^40, 2^55) apart must not draw the same numbers,
    and the medium-slot index (idx < 256) shares its counter word with those high bits without colliding."""
    lib = _port()
    base = _rng4(lib, 1, 5, 5, 0, 0, 0, 0)[0]
    for shift in (32, 33, 40, 55):
        far = _rng4(lib, 1, 5, 5 + (1 << shift), 0, 0, 0, 0)[0]
        assert not np.array_equal(base, far), shift
    a = _rng4(lib, 1, 5, 5 + (1 << 32), 0, 1, 0, 0)[0]   # sample_hi = 1, idx = 0
    b = _rng4(lib, 1, 5, 5, 0, 1, 0, 255)[0]             # sample_hi = 0, idx = 255
    c = _rng4(lib, 1, 5, 5 + (1 << 32), 0, 1, 0, 255)[0]
    assert not np.array_equal(a, b) and not np.array_equal(a, c) and not np.array_equal(b, c)
