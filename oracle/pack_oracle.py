"""CPU oracle: lossless frame packing for checkpoints (TEST INFRASTRUCTURE — see oracle/__init__.py).

No counterpart in the reference (README.md:24 lists observation compression as a TODO; the reference never persists
its replay).  Restates the format of ape-x-dqn_b200/csrc/pack.cu in numpy, byte for byte: every frame of F bytes
(nb = F/16 blocks) is XORed with its predecessor in slot order — every 64th frame of the packed range is a key frame
and kept as is — and only the 16-byte blocks with a non-zero byte are stored behind a bitmap of ceil(nb/32)
little-endian u32 words (padded to 16 bytes).
"""
from __future__ import annotations

import numpy as np

KEY_EVERY = 64


def pack_frames(frames: np.ndarray):
    """frames uint8 [count, F] -> (sizes uint32 [count], offsets uint64 [count], packed uint8 [total])."""
    frames = np.ascontiguousarray(frames, np.uint8)
    count, F = frames.shape
    assert F % 16 == 0
    nb = F // 16
    nw = (nb + 31) // 32
    head = (nw * 4 + 15) // 16 * 16
    parts, sizes = [], np.zeros(count, np.uint32)
    for j in range(count):
        key = j % KEY_EVERY == 0
        x = frames[j] if key else frames[j] ^ frames[j - 1]
        blocks = x.reshape(nb, 16)
        keep = np.ones(nb, bool) if key else blocks.any(axis=1)
        bits = np.zeros(nw * 32, np.uint8)
        bits[:nb] = keep
        words = (bits.reshape(nw, 32).astype(np.uint64) << np.arange(32, dtype=np.uint64)).sum(axis=1).astype("<u4")
        rec = np.zeros(head, np.uint8)
        rec[:nw * 4] = words.view(np.uint8)
        parts.append(np.concatenate([rec, blocks[keep].reshape(-1)]))
        sizes[j] = parts[-1].size
    offsets = np.concatenate([[0], np.cumsum(sizes.astype(np.uint64))[:-1]]).astype(np.uint64)
    return sizes, offsets, (np.concatenate(parts) if parts else np.zeros(0, np.uint8))


def unpack_frames(sizes, offsets, packed, F: int) -> np.ndarray:
    count = len(sizes)
    nb = F // 16
    nw = (nb + 31) // 32
    head = (nw * 4 + 15) // 16 * 16
    out = np.zeros((count, F), np.uint8)
    for j in range(count):
        rec = packed[int(offsets[j]):int(offsets[j]) + int(sizes[j])]
        words = rec[:nw * 4].view("<u4")
        keep = ((words[:, None] >> np.arange(32, dtype=np.uint32)) & 1).reshape(-1)[:nb].astype(bool)
        x = np.zeros((nb, 16), np.uint8)
        x[keep] = rec[head:].reshape(-1, 16)
        x = x.reshape(-1)
        out[j] = x if j % KEY_EVERY == 0 else x ^ out[j - 1]
    return out
