"""TEST INFRASTRUCTURE ONLY -- numpy restatement of the reference's sampler (python/sampling.py:10-52) with the one random draw made
explicit.  np.random.choice(a, p=p) draws a single uniform u from the global generator and returns
cumsum(p).searchsorted(u, side="right") after normalising the cumulative sum by its last element (numpy/random/mtrand.pyx, `choice`);
here u is an argument, everything else follows sampling.py line by line.  Pinned to the reference's own sampling.py in
tests/test_oracle.py (build container) and to tests/golden/sampling_golden.npz everywhere.
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this module.
"""
import numpy as np


def softmax(x: np.ndarray) -> np.ndarray:                      # sampling.py:5-8
    x = x - x.max(axis=-1, keepdims=True)
    e = np.exp(x)
    return e / e.sum(axis=-1, keepdims=True)


def sample_logits(logits: np.ndarray, temperature: float, top_p: float, u: float) -> int:
    """One row of float32 logits -> token id; u in [0, 1) replaces np.random.choice's draw."""
    probs = softmax(np.asarray(logits, np.float32).copy())      # :13-16
    if top_p == 0.0:                                            # :25-26
        top_p = 1.0
    if temperature == 0.0:                                      # :38-39
        return int(np.argmax(probs))
    if top_p < 1.0:                                             # :41-45
        sorted_probs = np.sort(probs)[::-1]
        cumulative_probs = np.cumsum(sorted_probs)
        cutoff = float(sorted_probs[np.argmax(cumulative_probs > top_p)])
        probs[probs < cutoff] = 0
    if temperature != 1.0:                                      # :47-48
        probs = np.power(probs, 1.0 / temperature)
    probs = probs / np.sum(probs)                               # :50
    cdf = np.asarray(probs, np.float64).cumsum()                # np.random.choice
    cdf /= cdf[-1]
    return int(cdf.searchsorted(u, side="right"))


def boundary_distance(logits: np.ndarray, temperature: float, top_p: float, u: float) -> float:
    """|u - nearest cdf step| of the final draw: a device / numpy disagreement is only acceptable when this is at rounding level."""
    probs = softmax(np.asarray(logits, np.float32).copy())
    if top_p == 0.0:
        top_p = 1.0
    if top_p < 1.0:
        sp = np.sort(probs)[::-1]
        cutoff = float(sp[np.argmax(np.cumsum(sp) > top_p)])
        probs[probs < cutoff] = 0
    if temperature not in (0.0, 1.0):
        probs = np.power(probs, 1.0 / temperature)
    cdf = np.asarray(probs / np.sum(probs), np.float64).cumsum()
    cdf /= cdf[-1]
    return float(np.abs(cdf - u).min())
