"""numpy statements of the two sides of the ACNE model:
  reference_eps   what the reference's arithmetic does: |d|^2 - 1 of a bounce direction built by
                  src/Renderer.cpp:88-107 and normalised by src/Ray.h:67-72 (sqrtf of the float-rounded sum)
  model_eps       what kernels_radiance.cu: ref_normalize_eps draws from 32 random bits
TEST INFRASTRUCTURE."""
import numpy as np


def ref_normalize(v):
    m = np.sqrt((v * v).sum(-1).astype(np.float32)).astype(np.float64)   # (double) sqrtf((float) sum)
    return v / m[..., None]


def _cross(a, b):   # Ray.h:64-66 (the y component is written -a.x*b.z + a.z*b.x)
    return np.stack([a[:, 1] * b[:, 2] - a[:, 2] * b[:, 1], -a[:, 0] * b[:, 2] + a[:, 2] * b[:, 0],
                     a[:, 0] * b[:, 1] - a[:, 1] * b[:, 0]], 1)


def reference_eps(n, rng, r_lo=0.05, r_hi=0.3):
    w = ref_normalize(rng.normal(size=(n, 3)) * rng.uniform(r_lo, r_hi, (n, 1)))     # normal = Normalize(hit - centre)
    r1, r2 = 2 * np.pi * rng.uniform(size=n), rng.uniform(size=n)
    r2s = np.sqrt(r2)
    u = ref_normalize(_cross(np.tile([0.0, 1.0, 0.0], (n, 1)), w))
    v = _cross(w, u)
    d = ref_normalize(u * (np.cos(r1) * r2s)[:, None] + v * (np.sin(r1) * r2s)[:, None] + w * np.sqrt(1 - r2)[:, None])
    return (d * d).sum(-1) - 1.0


def model_eps(bits):
    bits = np.asarray(bits, np.uint32)
    s = ((bits & 255) + ((bits >> 8) & 255) + ((bits >> 16) & 255) + (bits >> 24)).astype(np.int32)
    eta = ((s - 510).astype(np.float32) * np.float32(3.586e-10)).astype(np.float32)
    m = np.sqrt((np.float32(1.0) + eta).astype(np.float32)).astype(np.float32)
    return (eta - np.float32(2.0) * (m - np.float32(1.0))).astype(np.float32)
