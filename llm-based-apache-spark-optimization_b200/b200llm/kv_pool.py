"""One KV pool for several engines on a GPU (BASELINE.json config 5: duckdb-nsql-7B and Llama-3.2-3B co-served;
the reference calls both models from one process, Flask/app.py:102-107 and :160-166).

The pool is one allocation cut into BLOCKS.  An engine's KV page is a SLOT: slot_bytes = block_bytes / k, the largest
k >= 1 for which a page (b200llm_kv_page_bytes: all layers, K and V, 16 tokens) still fits, rounded down to the slot
alignment.  The engine with the largest page gets k = 1 (a block is exactly one of its pages); engines with smaller
pages pack k of them into a block (Llama-3.2-3B pages of 1.75 MiB go 4 to an 8 MiB block of the 7B).  Blocks move
between engines as demand moves: an engine takes a free block when it runs out of slots and gives a block back as
soon as all its slots are free, so neither model's KV is capped by a static partition -- only by what the other one
holds right now.  The library sees none of this: page ids are slot indices (b200llm_bind_kv_pool_shared).

Host-side bookkeeping only (plain Python, no torch beyond the allocation).  Thread-safe: the engines' scheduler
threads allocate concurrently.
"""
from __future__ import annotations

import threading
from typing import Dict, List


class SharedKVPool:
    def __init__(self, n_blocks: int, block_bytes: int, device: int = 0, allocate: bool = True):
        if n_blocks < 1 or block_bytes < 256 or block_bytes % 256:
            raise ValueError("SharedKVPool: need >= 1 block of a multiple of 256 bytes")
        self.n_blocks, self.block_bytes, self.device = n_blocks, block_bytes, device
        self.buffer = None
        if allocate:
            import torch
            self.buffer = torch.zeros(n_blocks * block_bytes, dtype=torch.uint8, device=device)
        self._free_blocks: List[int] = list(range(n_blocks - 1, -1, -1))
        self._lock = threading.RLock()
        self._views: Dict[str, "PoolView"] = {}
        self.peak_blocks_in_use = 0

    @property
    def nbytes(self) -> int:
        return self.n_blocks * self.block_bytes

    def data_ptr(self) -> int:
        return self.buffer.data_ptr()

    def view(self, name: str, page_bytes: int, align: int) -> "PoolView":
        """The pool as engine `name` sees it: slots of block_bytes / k bytes, k as large as a `page_bytes` page allows;
        every slot a multiple of `align` (one (kv head, 16-token) tile, itself a multiple of 256 bytes)."""
        with self._lock:
            if name in self._views:
                raise ValueError(f"SharedKVPool: '{name}' is attached already")
            if page_bytes > self.block_bytes:
                raise ValueError(f"SharedKVPool: a page of '{name}' ({page_bytes} B) exceeds the block ({self.block_bytes} B)")
            # the largest k whose slot (block / k, exactly) still holds a page and keeps the tile alignment
            k = self.block_bytes // page_bytes
            while k > 1 and (self.block_bytes % k or (self.block_bytes // k) % align):
                k -= 1
            slot = self.block_bytes // k
            if slot < page_bytes or slot % align:
                raise ValueError(f"SharedKVPool: cannot fit pages of '{name}' ({page_bytes} B, aligned to {align}) into "
                                 f"{self.block_bytes}-byte blocks")
            v = PoolView(self, name, k, slot)
            self._views[name] = v
            return v

    # ---- blocks (called by the views, under the pool lock) -------------------------------------------------
    def _take_block(self) -> int:
        if not self._free_blocks:
            return -1
        b = self._free_blocks.pop()
        self.peak_blocks_in_use = max(self.peak_blocks_in_use, self.n_blocks - len(self._free_blocks))
        return b

    def _give_block(self, b: int) -> None:
        self._free_blocks.append(b)

    def occupancy(self) -> dict:
        """Blocks held per engine and slots in use inside them, now."""
        with self._lock:
            out = {"blocks": self.n_blocks, "block_bytes": self.block_bytes, "free_blocks": len(self._free_blocks),
                   "peak_blocks_in_use": self.peak_blocks_in_use, "engines": {}}
            for name, v in self._views.items():
                out["engines"][name] = {"blocks_held": len(v._blocks), "slots_per_block": v.k, "slot_bytes": v.slot_bytes,
                                        "slots_in_use": v.slots_in_use, "bytes_in_use": v.slots_in_use * v.slot_bytes}
            return out


class PoolView:
    """Page allocator of one engine on a shared pool.  Looks like the engine's own free-page stack (pop / append /
    len), with page ids = slot indices in units of this engine's slot size."""

    def __init__(self, pool: SharedKVPool, name: str, k: int, slot_bytes: int):
        self.pool, self.name, self.k, self.slot_bytes = pool, name, k, slot_bytes
        self._blocks: Dict[int, List[int]] = {}     # block -> its free slots (page ids)
        self._used: Dict[int, int] = {}             # block -> slots handed out
        self.slots_in_use = 0
        # a slot index addresses slot_bytes units from the pool base, so block b's slots start at b * block_bytes / slot_bytes:
        # only exact when slot_bytes divides block_bytes; otherwise the engine would address with a per-block offset
        if pool.block_bytes % slot_bytes:
            raise ValueError(f"SharedKVPool: slot of {slot_bytes} B does not divide the {pool.block_bytes}-byte block")
        self.slots_per_block_stride = pool.block_bytes // slot_bytes

    @property
    def n_slots(self) -> int:
        """Upper bound of page ids + 1 (what the engine sizes its reference counts for)."""
        return self.pool.n_blocks * self.slots_per_block_stride

    def __len__(self) -> int:
        with self.pool._lock:
            return sum(len(f) for f in self._blocks.values()) + len(self.pool._free_blocks) * self.k

    def pop(self) -> int:
        with self.pool._lock:
            for b, free in self._blocks.items():
                if free:
                    self._used[b] += 1
                    self.slots_in_use += 1
                    return free.pop()
            b = self.pool._take_block()
            if b < 0:
                raise IndexError("shared KV pool exhausted")
            base = b * self.slots_per_block_stride
            self._blocks[b] = [base + j for j in range(self.k - 1, 0, -1)]
            self._used[b] = 1
            self.slots_in_use += 1
            return base

    def append(self, page: int) -> None:
        with self.pool._lock:
            b = page // self.slots_per_block_stride
            self._blocks[b].append(page)
            self._used[b] -= 1
            self.slots_in_use -= 1
            if self._used[b] == 0:              # the whole block is free again: back to the pool, for whoever needs it
                del self._blocks[b], self._used[b]
                self.pool._give_block(b)
