"""CPU: numpy Philox4x32-10 against the Random123 known-answer vectors (kat_vectors: philox4x32 10)."""
import numpy as np

from oracle import philox


def _run(c, k):
    return [int(v[0]) for v in philox.philox4x32_10(*[np.array([x], np.uint32) for x in c],
                                                    *[np.array([x], np.uint32) for x in k])]


def test_known_answers():
    assert _run([0, 0, 0, 0], [0, 0]) == [0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8]
    f = 0xFFFFFFFF
    assert _run([f, f, f, f], [f, f]) == [0x408F276D, 0x41C83B0E, 0xA20BC7C6, 0x6D5451FD]
    assert _run([0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344], [0xA4093822, 0x299F31D0]) == [
        0xD16CFE09, 0x94FDCCEB, 0x5001E420, 0x24126EA1]


def test_uniform_range_and_layout():
    u = philox.reset_uniforms(42, 3, np.arange(1000), 21, 0)
    v = philox.reset_uniforms(42, 3, np.arange(1000), 21, 1)
    assert u.shape == (1000, 21) and u.dtype == np.float32
    assert u.min() >= 0.0 and u.max() < 1.0 and abs(u.mean() - 0.5) < 0.01
    assert not np.array_equal(u, v)
    assert np.array_equal(u[10:20], philox.reset_uniforms(42, 3, np.arange(10, 20), 21, 0))
