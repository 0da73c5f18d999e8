"""sgl_kernel.kvcache — KVCacheManager with the reference API (tests/test_kv_cache_cpp.py:279-373),
storage and scatter/gather on the GPU.

Host side: a free-list page allocator with the reference's exact hand-out order
(PythonKVCacheReference.__init__/allocate_pages, tests/test_kv_cache_cpp.py:44-59: pages come out 0,1,2,…;
an unsatisfiable request returns an EMPTY int32 tensor and allocates nothing).
Device side: pool [max_pages, L, 2, H, page_size, D]; scatter_kv / gather_kv move rows [T, H*D] between
caller tensors and (page, slot, layer) with one CTA per token (wf_kv_scatter / wf_kv_gather).
Unlike the reference's Python class, the slot index is honoured (page_size tokens per page).
"""
from __future__ import annotations

import torch

import wf_bitnet
from wf_bitnet import ops as _ops


def check_kernel_available() -> bool:
    return wf_bitnet.available()


class PageAllocator:
    """Host free list with the reference's semantics (tests/test_kv_cache_cpp.py:44-70)."""

    def __init__(self, max_pages: int):
        self.max_pages = max_pages
        self._free = list(range(max_pages - 1, -1, -1))

    def num_free_pages(self) -> int:
        return len(self._free)

    def allocate_page(self) -> int:
        return self._free.pop() if self._free else -1

    def allocate_pages(self, num_pages: int) -> torch.Tensor:
        if num_pages < 0 or len(self._free) < num_pages:
            return torch.empty(0, dtype=torch.int32)
        return torch.tensor([self._free.pop() for _ in range(num_pages)], dtype=torch.int32)

    def free_page(self, page_id: int) -> None:
        if 0 <= page_id < self.max_pages:
            self._free.append(int(page_id))

    def free_pages(self, page_ids: torch.Tensor) -> None:
        for p in page_ids.tolist():
            self.free_page(p)

    free_pages_batch = free_pages


class KVCacheManager(PageAllocator):
    def __init__(self, num_layers: int, num_heads: int, head_dim: int, page_size: int, max_pages: int,
                 dtype: torch.dtype = torch.float32, device: str = "cuda"):
        if min(num_layers, num_heads, head_dim, page_size, max_pages) <= 0:
            raise ValueError("all KVCacheManager dimensions must be positive")
        super().__init__(max_pages)
        self.num_layers, self.num_heads, self.head_dim = num_layers, num_heads, head_dim
        self.page_size = page_size
        self.dtype = dtype
        self.kv = _ops.PagedKV(max_pages, num_layers, num_heads, page_size, head_dim, dtype=dtype, device=device)

    # ---- data movement (device) ---------------------------------------------------------------
    def _dev(self, t: torch.Tensor, dtype=None) -> torch.Tensor:
        t = t.to(device=self.kv.pool.device, dtype=dtype) if dtype is not None else t.to(self.kv.pool.device)
        return t.contiguous()

    def scatter_kv(self, k_in: torch.Tensor, v_in: torch.Tensor, page_indices: torch.Tensor, slot_indices: torch.Tensor,
                   layer_id: int) -> None:
        T = page_indices.shape[0]
        if k_in.numel() != T * self.num_heads * self.head_dim or v_in.numel() != k_in.numel():
            raise ValueError("k_in / v_in must hold [num_tokens, num_heads*head_dim] elements")
        if not 0 <= layer_id < self.num_layers:
            raise ValueError(f"layer_id {layer_id} out of range")
        _ops.kv_scatter(self.kv, self._dev(k_in.reshape(T, -1), self.dtype), self._dev(v_in.reshape(T, -1), self.dtype),
                        self._dev(page_indices, torch.int32), self._dev(slot_indices, torch.int32), layer_id)

    def gather_kv(self, k_out: torch.Tensor, v_out: torch.Tensor, page_indices: torch.Tensor, slot_indices: torch.Tensor,
                  layer_id: int) -> None:
        """In-place into the caller's buffers (reference: tests/test_kv_cache_cpp.py:328-335). CPU buffers are
        filled through a device staging copy."""
        T = page_indices.shape[0]
        if not 0 <= layer_id < self.num_layers:
            raise ValueError(f"layer_id {layer_id} out of range")
        dev = self.kv.pool.device
        direct = k_out.is_cuda and v_out.is_cuda and k_out.is_contiguous() and v_out.is_contiguous() and k_out.dtype == self.dtype
        kd = k_out if direct else torch.empty(T, self.num_heads * self.head_dim, dtype=self.dtype, device=dev)
        vd = v_out if direct else torch.empty(T, self.num_heads * self.head_dim, dtype=self.dtype, device=dev)
        _ops.kv_gather(self.kv, kd, vd, self._dev(page_indices, torch.int32), self._dev(slot_indices, torch.int32), layer_id)
        if not direct:
            k_out.copy_(kd.reshape(k_out.shape))
            v_out.copy_(vd.reshape(v_out.shape))
