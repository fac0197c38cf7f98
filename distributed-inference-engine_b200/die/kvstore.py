"""GPU-resident key-value cache; drop-in for the reference's KVCache (src/kvstore.py).

The table, its LRU order, TTL checks and hit/miss/eviction counters live in HBM and are driven by
the K2/K3 kernels (csrc/kvstore.cu) through ``die_kv_*``.  Keys are identified on the device by the
128-bit MD5 of their UTF-8 bytes (computed by the K1 kernel); values of arbitrary Python type are
kept on the host in a handle table and the device stores the 64-bit handle -- that is API-edge
bookkeeping, the cache logic itself never runs on the CPU.

``DeviceKV`` is the tensor-level API (digests in, hit flags / handles out, no host sync);
``KVCache`` mirrors the reference class method for method.
"""
from __future__ import annotations

import ctypes
import math
import time
from collections import OrderedDict
from dataclasses import dataclass
from typing import Any, Callable, Dict, Generic, List, Optional, Sequence, TypeVar

import torch

from . import _cabi, ops
from ._cabi import check, ptr, require_cuda, stream_ptr

T = TypeVar("T")


@dataclass
class CacheEntry(Generic[T]):
    """Same record as the reference's CacheEntry (src/kvstore.py:17-24)."""
    value: T
    created_at: float
    last_accessed: float
    access_count: int = 0
    ttl: Optional[float] = None


class DeviceKV:
    """Owns one ``die_kv_t``.  All tensor arguments are CUDA tensors; digests are int32 [B, 4]."""

    def __init__(self, max_size: int, default_ttl: Optional[float] = None, _handle: Optional[int] = None,
                 policy: str = "lru"):
        self._lib = require_cuda()
        self.max_size = max_size
        self.policy = policy
        self._owned = _handle is None
        if _handle is None:
            h = ctypes.c_void_p()
            check(self._lib.die_kv_create(max_size, _cabi.POLICY_IDS[policy],
                                          math.nan if default_ttl is None else float(default_ttl), ctypes.byref(h)))
            _handle = h.value
        self._h = _handle

    def __del__(self):
        if getattr(self, "_owned", False) and getattr(self, "_h", None):
            self._lib.die_kv_destroy(self._h)
            self._h = None

    def get_batch(self, digests: torch.Tensor, now: float, want_slots: bool = False):
        B = digests.shape[0]
        dev = digests.device
        hit = torch.empty(B, dtype=torch.uint8, device=dev)
        value = torch.zeros(B, dtype=torch.int64, device=dev)
        slot = torch.empty(B, dtype=torch.int32, device=dev) if want_slots else None
        check(self._lib.die_kv_get_batch(self._h, ptr(digests), B, now, ptr(hit), ptr(value), ptr(slot), stream_ptr()))
        return (hit, value, slot) if want_slots else (hit, value)

    def set_batch(self, digests: torch.Tensor, values: Optional[torch.Tensor], now: float,
                  ttl: Optional[torch.Tensor] = None, deferred: int = 0, want_dropped: bool = False):
        """-> (slot int32[M], dropped int64[2M] | None).  Status is read with ``sync_status``."""
        M = digests.shape[0]
        dev = digests.device
        slot = torch.empty(M, dtype=torch.int32, device=dev)
        dropped = torch.empty(2 * M, dtype=torch.int64, device=dev) if want_dropped else None
        check(self._lib.die_kv_set_batch(self._h, ptr(digests), ptr(values), ptr(ttl), M, now, int(deferred),
                                         ptr(slot), ptr(dropped), stream_ptr()))
        return slot, dropped

    def delete_batch(self, digests: torch.Tensor):
        B = digests.shape[0]
        deleted = torch.empty(B, dtype=torch.uint8, device=digests.device)
        value = torch.zeros(B, dtype=torch.int64, device=digests.device)
        check(self._lib.die_kv_delete_batch(self._h, ptr(digests), B, ptr(deleted), ptr(value), stream_ptr()))
        return deleted, value

    def contains_batch(self, digests: torch.Tensor, now: float):
        B = digests.shape[0]
        out = torch.empty(B, dtype=torch.uint8, device=digests.device)
        value = torch.zeros(B, dtype=torch.int64, device=digests.device)
        check(self._lib.die_kv_contains_batch(self._h, ptr(digests), B, now, ptr(out), ptr(value), stream_ptr()))
        return out, value

    def len_sync(self, now: float):
        """-> (len, purged handles list)."""
        cap = max(int(self.stats()["size"]), 1)
        purged = torch.empty(cap, dtype=torch.int64, device="cuda")
        n_purged, n = ctypes.c_uint64(), ctypes.c_uint64()
        check(self._lib.die_kv_len_sync(self._h, now, ptr(purged), cap, ctypes.byref(n_purged), ctypes.byref(n),
                                        stream_ptr()))
        return int(n.value), purged[: n_purged.value].tolist()

    def clear(self) -> None:
        check(self._lib.die_kv_clear(self._h, stream_ptr()))

    def stats(self) -> Dict[str, int]:
        out = (ctypes.c_uint64 * 8)()
        check(self._lib.die_kv_stats_sync(self._h, out, stream_ptr()))
        names = ("hits", "misses", "evictions", "size", "max_size", "capacity", "tombstones", "lru_log_span")
        return dict(zip(names, (int(x) for x in out)))

    def sync_status(self):
        """-> (status, n_dropped, slow_path_batches); status 4 = the reference raised StopIteration."""
        st, nd, slow = ctypes.c_int32(), ctypes.c_uint64(), ctypes.c_uint32()
        check(self._lib.die_kv_sync_status(self._h, ctypes.byref(st), ctypes.byref(nd), ctypes.byref(slow),
                                           stream_ptr()))
        return int(st.value), int(nd.value), int(slow.value)

    def maintain(self, force: bool = False) -> None:
        check(self._lib.die_kv_maintain_sync(self._h, 1 if force else 0, stream_ptr()))

    def dump(self):
        """Live entries in LRU order (head first), never-stamped entries last.  Host tensors."""
        cap = max(int(self.stats()["size"]), 1)
        dig = torch.empty((cap, 4), dtype=torch.int32)
        val = torch.empty(cap, dtype=torch.int64)
        created = torch.empty(cap, dtype=torch.float64)
        last = torch.empty(cap, dtype=torch.float64)
        cnt = torch.empty(cap, dtype=torch.int32)
        ttl = torch.empty(cap, dtype=torch.float64)
        stamped = torch.empty(cap, dtype=torch.uint8)
        lfu = torch.empty(cap, dtype=torch.int32)
        n = ctypes.c_uint64()
        check(self._lib.die_kv_dump_sync(self._h, cap, ptr(dig), ptr(val), ptr(created), ptr(last), ptr(cnt), ptr(ttl),
                                         ptr(stamped), ctypes.byref(n), stream_ptr(), ptr(lfu)))
        k = int(n.value)
        out = {"digests": dig[:k], "values": val[:k], "created_at": created[:k], "last_accessed": last[:k],
               "access_count": cnt[:k], "ttl": ttl[:k], "stamped": stamped[:k]}
        if self.policy == "lfu":
            out["lfu_count"] = lfu[:k]
        return out

    def snapshot(self) -> torch.Tensor:
        """The whole table (entries, eviction order, counters) as one flat uint8 host tensor (die_kv_snapshot_sync)."""
        nbytes = int(self._lib.die_kv_snapshot_bytes(self._h))
        image = torch.empty(nbytes, dtype=torch.uint8)
        check(self._lib.die_kv_snapshot_sync(self._h, ptr(image), nbytes, stream_ptr()))
        return image

    def restore(self, image: torch.Tensor) -> None:
        """Load an image taken from a table of the same max_size and policy."""
        image = image.contiguous()
        check(self._lib.die_kv_restore_sync(self._h, ptr(image), image.numel(), stream_ptr()))


def digests_of(keys: Sequence[str]) -> torch.Tensor:
    """MD5 digests of the keys, computed by the K1 kernel: int32 [B, 4] on the device."""
    _, dig, _ = ops.route_md5(ops.KeyBatch.from_strings(keys), 1)
    return dig


class KVCache:
    """Same constructor, methods, attributes and error behaviour as the reference's KVCache
    (src/kvstore.py:26-236), for the three eviction policies "lru", "lfu" and "fifo".  Extra keyword-only ``clock``
    replaces ``time.time`` (used by the parity tests); ``save`` / ``load`` are the persistence the reference's
    README.md:14 promises and never wrote."""

    def __init__(self, max_size: int = 1000, eviction_policy: str = "lru", default_ttl: Optional[float] = None,
                 *, clock: Optional[Callable[[], float]] = None):
        self.max_size = max_size
        self.eviction_policy = eviction_policy.lower()
        self.default_ttl = default_ttl
        if self.eviction_policy not in _cabi.POLICY_IDS:
            # the reference accepts any string and silently behaves as FIFO (kvstore.py:99 "else"); a typo is refused here
            raise ValueError(f"eviction_policy={eviction_policy!r}: expected 'lru', 'lfu' or 'fifo'")
        self._clock = clock or time.time
        self._kv = DeviceKV(max_size, default_ttl, policy=self.eviction_policy)
        self._objects: Dict[int, Any] = {}     # handle -> Python value
        self._keys: Dict[bytes, str] = {}      # digest -> key string (only to present .cache / LRU dumps)
        self._next_handle = 1

    # -- helpers ---------------------------------------------------------------------------------
    def _digests(self, keys: Sequence[str]) -> torch.Tensor:
        dig = digests_of(keys)
        return dig

    def _remember(self, keys: Sequence[str], dig: torch.Tensor) -> None:
        raw = dig.cpu().numpy().tobytes()
        for i, k in enumerate(keys):
            self._keys[raw[16 * i:16 * i + 16]] = k

    def _new_handles(self, values: Sequence[Any]) -> List[int]:
        hs = []
        for v in values:
            h = self._next_handle
            self._next_handle += 1
            self._objects[h] = v
            hs.append(h)
        return hs

    def _drop(self, handles) -> None:
        for h in handles:
            self._objects.pop(int(h), None)

    def _set_many(self, keys: List[str], values: List[Any], ttls: List[Optional[float]], deferred: int) -> None:
        if not keys:
            return
        now = float(self._clock())
        dig = self._digests(keys)
        self._remember(keys, dig)
        handles = self._new_handles(values)
        hv = torch.tensor(handles, dtype=torch.int64, device="cuda")
        tt = torch.tensor([math.nan if t is None else float(t) for t in ttls], dtype=torch.float64, device="cuda")
        slot, dropped = self._kv.set_batch(dig, hv, now, tt, deferred, want_dropped=True)
        status, n_dropped, _ = self._kv.sync_status()
        self._drop(dropped[:n_dropped].tolist())
        applied = slot.tolist()
        # duplicates inside the batch: only the last applied op of a key still owns a table entry
        last_of = {}
        for i, k in enumerate(keys):
            if applied[i] >= 0:
                last_of[k] = i
        for i, k in enumerate(keys):
            if applied[i] < 0 or last_of.get(k) != i:
                self._objects.pop(handles[i], None)
        if status == _cabi.DIE_ERR_STOP_ITERATION:
            raise StopIteration  # the reference's next(iter(OrderedDict())) at kvstore.py:89 / next(iter(cache)) :101
        if status == _cabi.DIE_ERR_VALUE_ERROR:
            raise ValueError("min() arg is an empty sequence")  # LFU with nothing counted, kvstore.py:95

    # -- reference API ---------------------------------------------------------------------------------
    def set(self, key: str, value: Any, ttl: Optional[float] = None, batch: bool = False) -> None:
        self._set_many([key], [value], [ttl], 2 if batch else 0)

    def get(self, key: str, default: Any = None) -> Any:
        now = float(self._clock())
        hit, value = self._kv.get_batch(self._digests([key]), now)
        code, h = int(hit[0]), int(value[0])
        if code == 1:
            return self._objects[h]
        if code == 2:
            self._objects.pop(h, None)
        return default

    def batch_get(self, keys: List[str]) -> Dict[str, Any]:
        if not keys:
            return {}
        now = float(self._clock())
        hit, value = self._kv.get_batch(self._digests(keys), now)
        out: Dict[str, Any] = {}
        for k, c, h in zip(keys, hit.tolist(), value.tolist()):
            if c == 2:
                self._objects.pop(h, None)
            out[k] = self._objects[h] if c == 1 else None
        return out

    def batch_set(self, items: Dict[str, Any], ttl: Optional[float] = None) -> None:
        keys = list(items)
        self._set_many(keys, [items[k] for k in keys], [ttl] * len(keys), 1)

    def delete(self, key: str) -> bool:
        deleted, value = self._kv.delete_batch(self._digests([key]))
        if int(deleted[0]):
            self._objects.pop(int(value[0]), None)
            return True
        return False

    def clear(self) -> None:
        self._kv.clear()
        self._objects.clear()
        self._keys.clear()

    def get_stats(self) -> Dict[str, Any]:
        s = self._kv.stats()
        total = s["hits"] + s["misses"]
        return {"hits": s["hits"], "misses": s["misses"], "hit_rate": s["hits"] / total if total > 0 else 0,
                "size": s["size"], "max_size": self.max_size, "evictions": s["evictions"],
                "eviction_policy": self.eviction_policy}

    def __contains__(self, key: str) -> bool:
        out, value = self._kv.contains_batch(self._digests([key]), float(self._clock()))
        code = int(out[0])
        if code == 2:
            self._objects.pop(int(value[0]), None)
        return code == 1

    def __len__(self) -> int:
        n, purged = self._kv.len_sync(float(self._clock()))
        self._drop(purged)
        return n

    # -- attributes callers of the reference read (examples/kvstore_demo.py:41,135-141,162-163) --------
    @property
    def hits(self) -> int:
        return self._kv.stats()["hits"]

    @property
    def misses(self) -> int:
        return self._kv.stats()["misses"]

    @property
    def evictions(self) -> int:
        return self._kv.stats()["evictions"]

    def _dump_entries(self):
        d = self._kv.dump()
        raw = d["digests"].numpy().tobytes()
        out = []
        for i in range(d["values"].numel()):
            key = self._keys.get(raw[16 * i:16 * i + 16])
            ttl = float(d["ttl"][i])
            entry = CacheEntry(value=self._objects.get(int(d["values"][i])), created_at=float(d["created_at"][i]),
                               last_accessed=float(d["last_accessed"][i]), access_count=int(d["access_count"][i]),
                               ttl=None if math.isnan(ttl) else ttl)
            out.append((key, entry, bool(d["stamped"][i])))
        return out

    @property
    def cache(self) -> Dict[str, CacheEntry]:
        """Snapshot of the table as ``{key: CacheEntry}`` (LRU order, head first)."""
        return {k: e for k, e, _ in self._dump_entries()}

    @property
    def _access_count(self) -> Dict[str, int]:
        """LFU: snapshot of the reference's private use-count dict, in the order the keys were first counted."""
        d = self._kv.dump()
        raw = d["digests"].numpy().tobytes()
        return {self._keys.get(raw[16 * i:16 * i + 16]): int(d["lfu_count"][i])
                for i in range(d["values"].numel()) if bool(d["stamped"][i])}

    def save(self, path: str) -> None:
        """Persist the cache: the device table's image plus the Python values its handles stand for."""
        import pickle
        with open(path, "wb") as f:
            pickle.dump({"image": self._kv.snapshot().numpy(), "objects": self._objects, "keys": self._keys,
                         "next_handle": self._next_handle, "max_size": self.max_size,
                         "eviction_policy": self.eviction_policy, "default_ttl": self.default_ttl}, f)

    def load(self, path: str) -> None:
        """Restore a cache saved by ``save`` from a KVCache of the same max_size and eviction_policy."""
        import pickle
        with open(path, "rb") as f:
            d = pickle.load(f)
        if (d["max_size"], d["eviction_policy"]) != (self.max_size, self.eviction_policy):
            raise ValueError("the saved cache has another max_size / eviction_policy")
        self._kv.restore(torch.from_numpy(d["image"]))
        self._objects, self._keys, self._next_handle = d["objects"], d["keys"], d["next_handle"]

    @property
    def _access_order(self) -> "OrderedDict[str, float]":
        """Snapshot of the LRU list (oldest first), like the reference's private attribute."""
        return OrderedDict((k, e.last_accessed) for k, e, stamped in self._dump_entries() if stamped)


KVStore = KVCache
create_kv_store = KVCache
