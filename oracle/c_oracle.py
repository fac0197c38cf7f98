"""ctypes access to the C side of the oracle (TEST INFRASTRUCTURE).  `make -C oracle` builds it."""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))


def _load(name):
    path = os.path.join(_HERE, "_build", name)
    if not os.path.exists(path):
        import subprocess
        subprocess.run(["make", "-C", _HERE], check=True, capture_output=True)
    return ctypes.CDLL(path)


def encode(keys):
    enc = [k.encode() for k in keys]
    off = np.zeros(len(enc) + 1, dtype=np.int64)
    if enc:
        np.cumsum([len(e) for e in enc], out=off[1:])
    data = np.frombuffer(b"".join(enc) + b"\0" * 16, dtype=np.uint8).copy()
    return data, off


def route(key_bytes, key_off, n_shards, routable=None, healthy=None, failover=True):
    """C oracle of route_request over CSR keys -> (shard int32[B], digests uint8[B,16])."""
    lib = _load("liboracle.so")
    B = len(key_off) - 1
    shard = np.empty(B, dtype=np.int32)
    dig = np.empty((B, 16), dtype=np.uint8)
    r = None if routable is None else np.ascontiguousarray(routable, dtype=np.uint8)
    h = None if healthy is None or len(healthy) == 0 else np.ascontiguousarray(healthy, dtype=np.int32)
    vp = ctypes.c_void_p
    lib.oracle_route(vp(key_bytes.ctypes.data), vp(key_off.ctypes.data), ctypes.c_int64(B), ctypes.c_int32(n_shards),
                     vp(None if r is None else r.ctypes.data), vp(None if h is None else h.ctypes.data),
                     ctypes.c_int32(0 if h is None else len(h)), ctypes.c_int32(1 if failover else 0),
                     vp(shard.ctypes.data), vp(dig.ctypes.data))
    return shard, dig


def kernel_md5_host(key_bytes, key_off, n_shards):
    """The product kernel's MD5 + mod code compiled for the host -> (digest words uint32[B,4], shard int32[B])."""
    lib = _load("libmd5kernelhost.so")
    B = len(key_off) - 1
    dig = np.empty((B, 4), dtype=np.uint32)
    shard = np.empty(B, dtype=np.int32)
    vp = ctypes.c_void_p
    lib.kernel_md5_host(vp(key_bytes.ctypes.data), vp(key_off.ctypes.data), ctypes.c_int64(B), ctypes.c_int32(n_shards),
                        vp(dig.ctypes.data), vp(shard.ctypes.data))
    return dig, shard
