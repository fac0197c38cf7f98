"""CPU decode of the binary request frames (TEST INFRASTRUCTURE; the format is this repository's, include/die.h --
the reference speaks JSON, src/worker.py:92-100, so this side of the parity is "unpinned": it checks the kernel against
an independent struct-based reading of the same bytes, and against the JSON request the frame stands for)."""
import struct

import numpy as np

MAGIC = 0x31454944
HEADER = struct.Struct("<IIQHHI")


def decode(buf: bytes, offs, dim: int):
    """-> list of (status, request_id, queue, key bytes, row fp32[dim]) per frame."""
    out = []
    for a, b in zip(offs[:-1], offs[1:]):
        a, b = int(a), int(b)
        bad = (1, 0, -1, b"", np.zeros(dim, np.float32))
        if b - a < HEADER.size or a % 4:
            out.append(bad)
            continue
        magic, size, rid, key_len, queue, n_floats = HEADER.unpack_from(buf, a)
        need = HEADER.size + (key_len + 3) // 4 * 4 + 4 * n_floats
        if magic != MAGIC or size != b - a or need != b - a:
            out.append(bad)
            continue
        if n_floats > dim:
            out.append((2, 0, -1, b"", np.zeros(dim, np.float32)))
            continue
        k0 = a + HEADER.size
        key = bytes(buf[k0:k0 + key_len])
        x0 = k0 + (key_len + 3) // 4 * 4
        row = np.zeros(dim, np.float32)
        row[:n_floats] = np.frombuffer(buf, dtype="<f4", count=n_floats, offset=x0)
        out.append((0, rid, queue, key, row))
    return out
