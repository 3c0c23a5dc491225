"""The seeded logits rows of the sampling fixtures (shared by the generator, which needs /root/reference, and by the tests, which do not)."""
import numpy as np


def row(i: int, vocab: int = 4096) -> np.ndarray:
    rng = np.random.default_rng(10_000 + i)
    scale = [0.5, 2.0, 6.0][i % 3]                       # flat, medium and peaked distributions
    return (rng.standard_normal(vocab) * scale).astype(np.float32)
