"""Python host of the decode engine: PyTorch allocates device memory (weights, paged KV
pool, workspace) and hands raw pointers to the C ABI (include/b200llm.h); everything
that computes runs in libb200llm.so.  Replaces what Ollama's runner does for the
reference's `ollama.generate` calls (Flask/app.py:102-107,160-166; FastAPI/app.py:85-90,
105-111): prefill the prompt, then decode token by token with the sampler on the device.
"""
from __future__ import annotations

import os

import ctypes as C
import threading
from dataclasses import dataclass
from typing import Iterable, List, Optional, Sequence

import numpy as np
import torch

from . import _lib
from .geometry import Geometry
from .weights import PackedWeights

PAGE_TOKENS = 16
MAX_DECODE_STEPS = 4096


@dataclass
class SamplingParams:
    """Ollama `options` subset honoured by the engine (SURVEY.md section 5 "Config / flags").
    temperature 0 = greedy, lowest id wins ties (the parity mode of BASELINE.json)."""
    temperature: float = 0.0
    top_k: int = 40
    top_p: float = 0.9
    repeat_penalty: float = 1.0
    seed: int = 0

    def c_struct(self) -> _lib.Sampling:
        return _lib.Sampling(float(self.temperature), int(self.top_k), float(self.top_p),
                             float(self.repeat_penalty), int(self.seed) & 0xFFFFFFFFFFFFFFFF)


@dataclass
class GenTiming:
    prefill_ms: float = 0.0
    decode_ms: float = 0.0
    prefill_tokens: int = 0
    decode_steps: int = 0
    kernel_launches: int = 0


class Engine:
    def __init__(self, geo: Geometry, weights: PackedWeights, device: int = 0, max_lanes: int = 8,
                 max_ctx: int = 4096, max_rows: int = 64, kv_pages: Optional[int] = None, tp_rank: int = 0,
                 tp_size: int = 1, nccl_id: Optional[bytes] = None, use_graphs: bool = True, use_pdl: bool = False,
                 attn_splits: int = 0, use_megakernel: bool = True, use_tc: bool = True, kv_pool=None, name: str = "",
                 lean_weights: bool = False):
        if not torch.cuda.is_available():
            raise _lib.B200LLMError(_lib.ENODEV, "no CUDA device: b200llm has no CPU fallback")
        self.geo, self.weights, self.device = geo, weights, device
        if use_megakernel and not any(k >= _lib.W_QKV_ST for k, _, _ in weights.tensors):
            # the persistent kernels stream their own copies of the matrices: tile-major for the
            # tcgen05 kernel, stage-major for the mma.sync one (geometries the former cannot tile)
            from .weights import add_stage_major, add_tile_major, tile_major_ok
            if tile_major_ok(geo, tp_size) and os.environ.get("B200LLM_PERSISTENT", "tc") == "tc":
                add_tile_major(weights)     # (tensor parallel: of this rank's shards; used once tp.connect_peers ran)
            elif tp_size == 1:
                add_stage_major(weights)
        if lean_weights:
            # keep the tile-major matrices only: the persistent kernel (1-4 lanes) and the tcgen05 GEMM path (everything else)
            # both stream them; the row-major ones, which only the GEMV kernels read, are dropped -- half the weight footprint
            from .weights import drop_row_major
            drop_row_major(weights)
        self.max_lanes, self.max_rows = max_lanes, max_rows
        self.max_pages_per_lane = (max_ctx + PAGE_TOKENS - 1) // PAGE_TOKENS
        torch.cuda.set_device(device)
        self.stream = torch.cuda.Stream(device)
        self._id_buf = C.create_string_buffer(nccl_id, 128) if nccl_id else None
        cfg = _lib.Config(_lib.ABI_VERSION, geo.d, geo.n_layers, geo.n_q, geo.n_kv, geo.head_dim, weights.ffn_cfg, geo.vocab,
                          geo.rms_eps, PAGE_TOKENS, max_lanes, self.max_pages_per_lane, max_rows, tp_rank, tp_size,
                          device, int(use_graphs), int(use_pdl), attn_splits, int(use_megakernel), int(use_tc), self.stream.cuda_stream,
                          C.cast(self._id_buf, C.c_void_p) if self._id_buf else None)
        self.cfg = cfg
        self._h = C.c_void_p()
        _lib.check(_lib.lib.b200llm_create(C.byref(cfg), C.byref(self._h)))
        nbytes = C.c_size_t()
        _lib.check(_lib.lib.b200llm_workspace_bytes(C.byref(cfg), C.byref(nbytes)))
        self.workspace = torch.empty(nbytes.value, dtype=torch.uint8, device=device)
        _lib.check(_lib.lib.b200llm_kv_page_bytes(C.byref(cfg), C.byref(nbytes)))
        self.page_bytes = nbytes.value
        descs = weights.descs()
        _lib.check(_lib.lib.b200llm_bind_weights(self._h, descs, len(descs)))
        if kv_pool is not None:
            # pages come from a pool shared with other engines on this GPU (kv_pool.SharedKVPool): a page is one slot of it
            tile = PAGE_TOKENS * geo.head_dim * 2
            self._free_pages = kv_pool.view(name or geo.name, self.page_bytes, tile)
            kv_pages = self._free_pages.n_slots
            self.kv_pool = kv_pool.buffer
            _lib.check(_lib.lib.b200llm_bind_kv_pool_shared(self._h, kv_pool.data_ptr(), kv_pool.nbytes, self._free_pages.slot_bytes))
        else:
            if kv_pages is None:
                kv_pages = max_lanes * self.max_pages_per_lane
            self.kv_pool = torch.zeros(kv_pages * self.page_bytes, dtype=torch.uint8, device=device)
            _lib.check(_lib.lib.b200llm_bind_kv_pool(self._h, self.kv_pool.data_ptr(), self.kv_pool.numel()))
            self._free_pages = list(range(kv_pages - 1, -1, -1))
        self.n_pages = kv_pages
        torch.cuda.synchronize(device)
        _lib.check(_lib.lib.b200llm_bind_workspace(self._h, self.workspace.data_ptr(), self.workspace.numel()))
        self._page_ref = [0] * kv_pages      # lanes and prefix-cache entries holding each page
        self.on_page_pressure = None         # callable(n_pages_short), set by the prefix cache
        self._lane_pages: List[List[int]] = [[] for _ in range(max_lanes)]
        self._free_lanes = list(range(max_lanes - 1, -1, -1))
        self.pos = [0] * max_lanes
        self.lock = threading.RLock()

    # ---- lanes and pages -------------------------------------------------------------
    def open_lane(self) -> int:
        with self.lock:
            if not self._free_lanes:
                raise _lib.B200LLMError(_lib.ESTATE, "no free lane")
            lane = self._free_lanes.pop()
            self.pos[lane] = 0
            return lane

    def close_lane(self, lane: int) -> None:
        with self.lock:
            self.release_pages(self._lane_pages[lane])
            self._lane_pages[lane] = []
            self._free_lanes.append(lane)

    # ---- shared KV pages (prefix cache) ------------------------------------------------
    def lane_pages(self, lane: int) -> List[int]:
        return list(self._lane_pages[lane])

    def retain_pages(self, pages: Sequence[int]) -> None:
        with self.lock:
            for p in pages:
                self._page_ref[p] += 1

    def release_pages(self, pages: Sequence[int]) -> None:
        with self.lock:
            for p in reversed(list(pages)):
                self._page_ref[p] -= 1
                if self._page_ref[p] == 0:
                    self._free_pages.append(p)

    def attach_prefix(self, lane: int, pages: Sequence[int], ids: Sequence[int] = ()) -> None:
        """Start a fresh lane on KV pages that already hold its first len(pages) * 16 tokens (computed by
        an earlier request with the same token prefix).  The pages are shared read-only: the lane's own
        tokens go to pages it allocates after them, so continue with prefill(lane, rest, pos0=len(pages) * 16).
        `ids` are those first tokens: the last 64 of them seed the lane's repeat-penalty window."""
        with self.lock:
            if self._lane_pages[lane]:
                raise _lib.B200LLMError(_lib.ESTATE, "attach_prefix: the lane already has pages")
            if len(pages) > self.max_pages_per_lane:
                raise _lib.B200LLMError(_lib.EINVAL, "attach_prefix: prefix longer than the lane context")
            self.retain_pages(pages)
            self._lane_pages[lane] = list(pages)
            if pages:
                arr = (C.c_int32 * len(pages))(*pages)
                _lib.check(_lib.lib.b200llm_set_page_table(self._h, lane, arr, len(pages)))
            n = len(pages) * PAGE_TOKENS
            tail = np.ascontiguousarray(np.asarray(ids[:n], dtype=np.int32)[-64:])
            _lib.check(_lib.lib.b200llm_set_history(self._h, lane, tail.ctypes.data_as(C.POINTER(C.c_int32)), len(tail),
                                                    max(0, n - len(tail))))
            self.pos[lane] = n

    def _ensure_pages(self, lane: int, n_tokens: int) -> None:
        need = (n_tokens + PAGE_TOKENS - 1) // PAGE_TOKENS
        pages = self._lane_pages[lane]
        if need > self.max_pages_per_lane:
            raise _lib.B200LLMError(_lib.EINVAL, f"{n_tokens} tokens exceed the lane context "
                                                 f"({self.max_pages_per_lane * PAGE_TOKENS})")
        if need <= len(pages):
            return
        if need - len(pages) > len(self._free_pages) and self.on_page_pressure is not None:
            self.on_page_pressure(need - len(pages) - len(self._free_pages))   # the prefix cache gives pages back
        if need - len(pages) > len(self._free_pages):
            raise _lib.B200LLMError(_lib.ENOMEM, "KV pool exhausted")
        try:
            while len(pages) < need:
                pages.append(self._free_pages.pop())
                self._page_ref[pages[-1]] = 1
        except IndexError:      # a shared pool: another engine took the blocks in the meantime
            arr = (C.c_int32 * len(pages))(*pages)
            _lib.check(_lib.lib.b200llm_set_page_table(self._h, lane, arr, len(pages)))
            raise _lib.B200LLMError(_lib.ENOMEM, "KV pool exhausted") from None
        arr = (C.c_int32 * len(pages))(*pages)
        _lib.check(_lib.lib.b200llm_set_page_table(self._h, lane, arr, len(pages)))

    def set_sampling(self, lane: int, sp: SamplingParams) -> None:
        s = sp.c_struct()
        _lib.check(_lib.lib.b200llm_set_sampling(self._h, lane, C.byref(s)))

    def set_lane(self, lane: int, next_token: int, pos: int) -> None:
        """Teacher forcing / resume: overwrite the lane's next input token and position."""
        self._ensure_pages(lane, pos + 1)
        _lib.check(_lib.lib.b200llm_set_lane(self._h, lane, int(next_token), int(pos)))
        self.pos[lane] = pos

    # ---- the hot path ----------------------------------------------------------------
    def prefill(self, lane: int, ids: Sequence[int], pos0: int = 0) -> None:
        ids = np.ascontiguousarray(ids, dtype=np.int32)
        self._ensure_pages(lane, pos0 + len(ids))
        _lib.check(_lib.lib.b200llm_prefill(self._h, lane, ids.ctypes.data_as(C.POINTER(C.c_int32)), len(ids), pos0))
        self.pos[lane] = pos0 + len(ids)

    def decode(self, lanes: Sequence[int], n_steps: int, read: bool = True) -> Optional[np.ndarray]:
        """n_steps decode steps for `lanes` as one batch; returns int32 [n_steps, len(lanes)]."""
        lanes_arr = (C.c_int32 * len(lanes))(*lanes)
        for ln in lanes:
            self._ensure_pages(ln, self.pos[ln] + n_steps)
        _lib.check(_lib.lib.b200llm_decode(self._h, lanes_arr, len(lanes), n_steps))
        for ln in lanes:
            self.pos[ln] += n_steps
        return self.read_tokens(n_steps, len(lanes)) if read else None

    def read_tokens(self, n_steps: int, n_lanes: int) -> np.ndarray:
        out = np.empty((n_steps, n_lanes), dtype=np.int32)
        _lib.check(_lib.lib.b200llm_read_tokens(self._h, out.ctypes.data_as(C.POINTER(C.c_int32)), n_steps, n_lanes))
        return out

    def read_lane(self, lane: int):
        t, p = C.c_int32(), C.c_int32()
        _lib.check(_lib.lib.b200llm_read_lane(self._h, lane, C.byref(t), C.byref(p)))
        return t.value, p.value

    def logits(self, row: int = 0, n: Optional[int] = None) -> np.ndarray:
        n = n or (self.geo.vocab // self.cfg.tp_size)
        out = np.empty(n, dtype=np.float32)
        _lib.check(_lib.lib.b200llm_step_logits(self._h, row, out.ctypes.data_as(C.POINTER(C.c_float)), n))
        return out

    def sync(self) -> None:
        _lib.check(_lib.lib.b200llm_sync(self._h))

    def timing(self) -> GenTiming:
        t = _lib.Timing()
        _lib.check(_lib.lib.b200llm_last_timing(self._h, C.byref(t)))
        return GenTiming(t.prefill_ms, t.decode_ms, t.prefill_tokens, t.decode_steps, t.kernel_launches)

    # ---- one request, start to finish -------------------------------------------------
    def generate_ids(self, prompt_ids: Sequence[int], max_new: int, sampling: Optional[SamplingParams] = None,
                     stop_ids: Iterable[int] = (), chunk: int = 64):
        """Prefill + decode of one sequence.  Returns (new token ids, GenTiming).  With stop ids
        the decode runs in chunks of `chunk` steps and is cut at the first stop token."""
        stop = set(int(s) for s in stop_ids)
        with self.lock:
            lane = self.open_lane()
            try:
                self.set_sampling(lane, sampling or SamplingParams())
                self.prefill(lane, prompt_ids)
                first, _ = self.read_lane(lane)
                t = self.timing()
                toks = [first]
                tm = GenTiming(t.prefill_ms, 0.0, t.prefill_tokens, 0, 0)
                while len(toks) < max_new and toks[-1] not in stop:
                    n = min(max_new - len(toks), chunk if stop else MAX_DECODE_STEPS)
                    out = self.decode([lane], n)[:, 0]
                    t = self.timing()
                    tm.decode_ms += t.decode_ms
                    tm.decode_steps += n
                    tm.kernel_launches += t.kernel_launches
                    for tok in out.tolist():
                        toks.append(tok)
                        if tok in stop:
                            break
                    else:
                        continue
                    break
                return toks, tm
            finally:
                self.close_lane(lane)

    def close(self) -> None:
        if self._h:
            _lib.lib.b200llm_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
