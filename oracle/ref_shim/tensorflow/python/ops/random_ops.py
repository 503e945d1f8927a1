"""Seeded normal draws: NumPy's generator, not TensorFlow's -- same distribution, different numbers."""
import numpy as np


def random_normal(shape, mean=0.0, stddev=1.0, dtype=None, seed=None, name=None):
    return (np.random.default_rng(seed).standard_normal(tuple(shape)) * stddev + mean).astype(np.float32)
