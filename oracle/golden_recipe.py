"""TEST INFRASTRUCTURE: deterministic regeneration of the large tensors of the default-config golden fixtures.

The default model (base.yaml: 16 levels x 8 features, 2^22-entry hashmap, 256-wide MLPs) has a 1.46 GB hash table and
~0.9 M MLP weights -- too large to commit.  The fixture stores a RECIPE instead (seed, shape, mean, std per tensor;
seed + amplitude for the table) and both sides -- oracle/make_golden.py in the authoring container and the GPU test --
rebuild the tensors with numpy's PCG64 generator, whose stream is stable across numpy versions and platforms.
"""
import numpy as np

BIG = 1024   # tensors with at least this many elements are regenerated from a recipe; smaller ones are stored


def normal_tensor(seed, shape, mean, std):
    rng = np.random.Generator(np.random.PCG64(int(seed)))
    return (rng.standard_normal(tuple(int(s) for s in shape), dtype=np.float32) * np.float32(std) + np.float32(mean))


def table_tensor(seed, n, amplitude):
    """U(-amplitude, amplitude) float32 [n]; generated in chunks so that the peak host memory stays near n * 4 bytes."""
    rng = np.random.Generator(np.random.PCG64(int(seed)))
    out = np.empty(int(n), dtype=np.float32)
    step = 1 << 24
    for lo in range(0, int(n), step):
        hi = min(lo + step, int(n))
        out[lo:hi] = (rng.random(hi - lo, dtype=np.float32) * np.float32(2.0) - np.float32(1.0)) * np.float32(amplitude)
    return out


def sample_positions(n, count, seed):
    """`count` distinct flat positions in [0, n) (sorted), for spot-checking a large gradient tensor."""
    rng = np.random.Generator(np.random.PCG64(int(seed)))
    if n <= count:
        return np.arange(n, dtype=np.int64)
    return np.unique(rng.integers(0, n, size=count, dtype=np.int64))
