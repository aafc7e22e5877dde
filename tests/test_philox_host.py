import numpy as np
from philox_ref import philox4x32_10, KAT


def test_numpy_philox_known_answers():
    for ctr, key, want in KAT:
        got = philox4x32_10(np.array([ctr], np.uint32), key)[0]
        assert tuple(int(v) for v in got) == want
