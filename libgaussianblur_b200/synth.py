"""Deterministic synthetic u8 inputs: px[i] = fmix32(lo(i) ^ hi(i)*0x9E3779B9 ^ seed) >> 24
(murmur3 finaliser), identical to the generator the tests and the CPU baseline use."""
import numpy as np


def synth_u8(shape, seed: int) -> np.ndarray:
    n = int(np.prod(shape))
    i = np.arange(n, dtype=np.uint64)
    h = (i & np.uint64(0xFFFFFFFF)).astype(np.uint32)
    hi = (i >> np.uint64(32)).astype(np.uint32)
    h = h ^ (hi * np.uint32(0x9E3779B9)) ^ np.uint32(seed & 0xFFFFFFFF)
    h ^= h >> np.uint32(16)
    h *= np.uint32(0x85EBCA6B)
    h ^= h >> np.uint32(13)
    h *= np.uint32(0xC2B2AE35)
    h ^= h >> np.uint32(16)
    return (h >> np.uint32(24)).astype(np.uint8).reshape(shape)
