"""The bounce loop's hash (philox.cuh: rtm_bounce_rng, pcg3d + final xorshift) restated in numpy and checked on
the lattice inputs the kernels feed it: (sub-pixel index, sample<<9|depth, key).  The device function is compared
with this restatement in test_gpu_parity.py through the images' statistics (rays/sample, mean radiance)."""
import numpy as np


def bounce_rng(cx, cy, key):
    M, A = np.uint32(1664525), np.uint32(1013904223)
    with np.errstate(over="ignore"):
        x = cx * M + A
        y = cy * M + A
        z = (np.zeros_like(cx) + np.uint32(key)) * M + A
        x = x + y * z; y = y + z * x; z = z + x * y
        x ^= x >> np.uint32(16); y ^= y >> np.uint32(16); z ^= z >> np.uint32(16)
        x = x + y * z; y = y + z * x; z = z + x * y
        y = y ^ z
        x ^= x >> np.uint32(16); y ^= y >> np.uint32(16)
    return x, y


def test_bit_balance_and_serial_correlation_on_lattice_inputs():
    n = 1 << 21
    q = np.arange(n, dtype=np.uint32)
    zs = []
    for key in (0x5EED, 1, 0, 77):
        for c in (0, 1, 512, 1538, 99999):
            x, y = bounce_rng(q, np.full_like(q, c), key)
            for w in (x, y):
                for b in (0, 1, 5, 11, 31):
                    zs.append((((w >> np.uint32(b)) & 1).mean() - 0.5) / (0.5 / np.sqrt(n)))
            lo = (x & 0xFFF).astype(np.float64)          # the sampler's height takes its top bits from here
            zs.append(np.corrcoef(lo[:-1], lo[1:])[0, 1] * np.sqrt(n))
            u, v = x / 2.0 ** 32, y / 2.0 ** 32
            zs.append(np.corrcoef(u, v)[0, 1] * np.sqrt(n))
    zs = np.array(zs)
    assert np.abs(zs).max() < 4.5 and np.sqrt((zs ** 2).mean()) < 1.25, (np.abs(zs).max(), np.sqrt((zs ** 2).mean()))


def test_counter_axis_uniform():
    c = np.arange(1 << 20, dtype=np.uint32)
    x, _ = bounce_rng(np.full_like(c, 12345), c, 0x5EED)
    h, _ = np.histogram(x / 2.0 ** 32, 256)
    chi2 = ((h - h.mean()) ** 2 / h.mean()).sum() / 255
    assert 0.75 < chi2 < 1.3, chi2
    u = x / 2.0 ** 32
    assert abs(np.corrcoef(u[:-1], u[1:])[0, 1]) < 5e-3 and abs(np.corrcoef(u[:-512], u[512:])[0, 1]) < 5e-3
