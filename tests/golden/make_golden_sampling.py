#!/usr/bin/env python3
"""tests/golden/sampling_golden.npz: token ids the REFERENCE's own python/sampling.py returns (build container: imports it from
/root/reference/python) for seeded logits rows; np.random.seed(i) fixes the uniform np.random.choice draws, and the same u is
recorded so the oracle / the device sampler can be driven with it.  Rows are regenerated from their seeds (numpy PCG64)."""
import os, sys
import numpy as np
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, "/root/reference/python")
import sampling as ref_sampling  # noqa: E402
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from tests.golden.make_golden_sampling_rows import row  # noqa: E402

CASES = [(1.0, 0.8), (0.7, 0.5), (1.3, 0.95), (1.0, 1.0), (0.0, 0.8), (0.5, 0.0)]
N_ROWS, VOCAB = 1000, 4096


def main():
    out = {"cases": np.float32(CASES), "n_rows": np.int64(N_ROWS), "vocab": np.int64(VOCAB)}
    for c, (T, P) in enumerate(CASES):
        toks, us = np.empty(N_ROWS, np.int64), np.empty(N_ROWS, np.float64)
        for i in range(N_ROWS):
            np.random.seed(i)
            us[i] = np.random.random_sample()            # the draw np.random.choice will make after the same seed
            np.random.seed(i)
            toks[i] = ref_sampling.sample_logits(row(i).copy(), T, P)
        out[f"tokens_{c}"], out[f"u_{c}"] = toks, us
        print(T, P, toks[:8])
    np.savez_compressed(os.path.join(HERE, "sampling_golden.npz"), **out)


if __name__ == "__main__":
    main()
