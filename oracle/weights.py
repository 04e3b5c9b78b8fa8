"""
TEST INFRASTRUCTURE.  Deterministic, RNG-library-independent weights for parity tests: every floating-point entry of
a state dict is filled from numpy's PCG64 stream seeded with crc32(key), scaled like torch's default Linear init
(uniform(-1/sqrt(fan_in), 1/sqrt(fan_in))); LayerNorm weights are 1 + U(-0.1, 0.1).  Statistics buffers
(to_k / to_eq / k_mean / k_std / n_periodicity / permutations / positional_encoding) are left untouched.
The same function is applied to the reference model (oracle/make_golden.py) and to the product model (tests), so no
83 MB weight file has to be committed.
"""
import zlib

import numpy as np
import torch

_KEEP = ("to_k.", "to_eq.", ".k_mean", ".k_std", "n_periodicity", "permutations", "permutation_prefactors",
         "positional_encoding")


def deterministic_fill_(sd, seed: int = 0, bias_scale: float = 1.0):
    """In-place on the tensors of `sd` (a state dict or dict of tensors). Returns sd."""
    for key in sd.keys():
        t = sd[key]
        if not torch.is_floating_point(t) or any(s in key for s in _KEEP):
            continue
        rng = np.random.Generator(np.random.PCG64(zlib.crc32(key.encode()) + seed))
        u = rng.random(t.numel()).astype(np.float64) * 2.0 - 1.0
        is_norm = ("norm" in key.split(".")[-2]) if "." in key else False
        if is_norm and key.endswith("weight"):
            v = 1.0 + 0.1 * u
        elif is_norm:
            v = 0.1 * u
        elif t.dim() >= 2:
            v = u / np.sqrt(t.shape[-1])
        else:  # Linear bias / in_proj_bias
            v = bias_scale * 0.05 * u
        with torch.no_grad():
            t.copy_(torch.from_numpy(v.astype(np.float32)).view(t.shape))
    return sd
