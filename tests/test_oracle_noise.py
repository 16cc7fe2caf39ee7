"""CPU: oracle/noise.py, the restated noise stream of the c2 / c5 kernel."""
import numpy as np

from oracle.noise import HotNoise, auto_shape, normal4_ck, vpt_of_shape
from oracle.highway import philox4x32_10


def test_philox_fixed_key_matches_the_generic_generator():
    # same permutation as the Random123-pinned generator with the kernel's compile-time key
    assert philox4x32_10([1, 2, 3, 4], (0xA4093822, 0x299F31D0)) == \
        philox4x32_10([1, 2, 3, 4], [0xA4093822, 0x299F31D0])
    z = normal4_ck(0, 256, 7, 99)
    assert len(z) == 4 and all(np.isfinite(z))


def test_stream_is_standard_normal_and_keyed_by_env_episode_vehicle():
    for shape in (1, 2, 4, 8, 16, 32):
        n = HotNoise(99, 7, 1, shape)
        z = np.array([n(s, v) for s in range(600) for v in range(21)])
        assert abs(z.mean()) < 0.03 and abs(z.std() - 1.0) < 0.03, shape
        assert n(5, 3) == HotNoise(99, 7, 1, shape)(5, 3)              # counter-based: stateless
        assert n(5, 3) != HotNoise(99, 8, 1, shape)(5, 3)              # other env
        assert n(5, 3) != HotNoise(99, 7, 2, shape)(5, 3)              # other episode
        assert n(5, 3) != n(5, 4) and n(5, 3) != n(6, 3)


def test_shape_tables_follow_the_engine():
    assert [vpt_of_shape(g) for g in (1, 2, 4, 8, 16, 32)] == [22, 11, 6, 3, 2, 1]
    assert auto_shape(4096) == 8 and auto_shape(65536) == 1 and auto_shape(65536, f64=True) == 2
    assert auto_shape(5) == 32
