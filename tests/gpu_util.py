"""Helpers for the -m gpu parity tests: torch owns device memory, every compute
call goes through the C ABI (ctypes) of libkafka_b200.so."""
import numpy as np
import torch

from kafka_inferenceengine_b200 import _lib
from kafka_inferenceengine_b200.device import padded

DEV = "cuda:0"


def planar(a, ld=None):
    """numpy (n, k) -> device (k, ld) float64 planes, zero padded."""
    a = np.asarray(a, dtype=np.float64)
    if a.ndim == 1:
        a = a[:, None]
    n, k = a.shape
    ld = ld or padded(n)
    t = torch.zeros((k, ld), dtype=torch.float64, device=DEV)
    t[:, :n] = torch.from_numpy(np.ascontiguousarray(a.T)).to(DEV)
    return t


def unplanar(t, n):
    return t[:, :n].t().contiguous().cpu().numpy()


def pack_blocks(blocks, ld=None):
    """numpy (n, P, P) symmetric -> device packed planar (L, ld)."""
    n, P, _ = blocks.shape
    r, c = np.tril_indices(P)
    return planar(blocks[:, r, c], ld)


def unpack_blocks(t, n, P):
    low = unplanar(t, n)
    out = np.zeros((n, P, P))
    r, c = np.tril_indices(P)
    out[:, r, c] = low
    out[:, c, r] = low
    return out


def random_spd(rng, n, P, scale=1.0):
    G = rng.standard_normal((n, P, P)) * 0.5
    return scale * (4.0 * np.eye(P)[None] + G @ np.transpose(G, (0, 2, 1)))


def stream():
    return _lib.stream_ptr()


def rel(a, b, floor=0.0):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return np.abs(a - b) / np.maximum(np.abs(b), floor if floor else 1e-300)
