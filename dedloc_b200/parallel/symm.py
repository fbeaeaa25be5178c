"""Symmetric peer memory over CUDA IPC (SURVEY 5.8 data plane).

Each peer (one process per GPU) allocates ONE device buffer with an identical internal layout and publishes
its 64-byte IPC handle (through the DHT record of the peer, see collab/optimizer.py). Every other peer maps it
once; the fused all-reduce kernel then addresses `peer_base + region_offset` directly with P2P loads/stores
over NVLink. This replaces hivemind's CPU shared-memory tensors + gRPC streams.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, Optional

import torch

from .. import lib as native


class _CudaArray:
    """Minimal __cuda_array_interface__ holder so torch can view raw device memory without copying."""

    def __init__(self, ptr: int, nbytes: int, owner):
        self.__cuda_array_interface__ = {
            "shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3, "strides": None,
        }
        self._owner = owner  # keeps the allocation alive as long as a view exists


class MappingLeases:
    """Which peer mappings are still worth keeping. A mapping of a peer that left pins that peer's whole buffer on ITS
    GPU until every importer closes it (CUDA IPC keeps the allocation alive for the importers), so a trainer that is
    killed and restarted a few times would fill its GPU with the buffers of its dead predecessors. Every use of a
    mapping renews its lease; `expired()` names the handles nobody used for `ttl` seconds."""

    def __init__(self, ttl: float):
        self.ttl = float(ttl)
        self._last_used: Dict[bytes, float] = {}

    def touch(self, handles, now: float) -> None:
        for h in handles:
            if h:
                self._last_used[bytes(h)] = now

    def expired(self, now: float, keep=()) -> list:
        keep = {bytes(h) for h in keep if h}
        out = [h for h, t in self._last_used.items() if now - t > self.ttl and h not in keep]
        for h in out:
            del self._last_used[h]
        return out


class SymmetricBuffer:
    """One device allocation that can be mapped by peer processes through a CUDA IPC handle."""

    def __init__(self, nbytes: int, device: Optional[int] = None):
        if device is None:
            device = torch.cuda.current_device()
        self.device = device
        self.nbytes = int(nbytes)
        ptr = C.c_void_p()
        handle = C.create_string_buffer(native.IPC_HANDLE_BYTES)
        with torch.cuda.device(device):
            native.check(native.lib().dedloc_symm_alloc(self.nbytes, C.byref(ptr), handle), "symm_alloc")
        self.ptr = int(ptr.value)
        self.handle = bytes(handle.raw)
        self._peers: Dict[bytes, int] = {}
        self._bytes = torch.as_tensor(_CudaArray(self.ptr, self.nbytes, self), device=f"cuda:{device}")

    def view(self, offset: int, numel: int, dtype: torch.dtype) -> torch.Tensor:
        esz = torch.empty((), dtype=dtype).element_size()
        return self._bytes[offset:offset + numel * esz].view(dtype)

    def open_peer(self, handle: bytes) -> int:
        """Map a peer's buffer (idempotent); returns the device pointer valid in THIS process."""
        if handle == self.handle:
            return self.ptr
        if handle not in self._peers:
            ptr = C.c_void_p()
            with torch.cuda.device(self.device):
                native.check(native.lib().dedloc_symm_open(handle, C.byref(ptr)), "symm_open")
            self._peers[handle] = int(ptr.value)
        return self._peers[handle]

    def close_peer(self, handle: bytes) -> None:
        ptr = self._peers.pop(handle, None)
        if ptr is not None:
            with torch.cuda.device(self.device):
                native.lib().dedloc_symm_close(ptr)

    def close(self) -> None:
        for h in list(self._peers):
            self.close_peer(h)
        # the allocation itself lives until process exit: peers may still have it mapped
