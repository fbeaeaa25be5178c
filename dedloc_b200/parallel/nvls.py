"""NVLS data plane: CUDA VMM allocations bound to NVSwitch multicast objects.

`multimem.ld_reduce` lets the switch add the same address range of every member's memory and hand back ONE copy;
`multimem.st` / `multimem.red` store to (or add into) all members at once. Per GPU and direction an all-reduce then
moves about N bytes instead of 2 (P-1)/P N. That needs memory allocated with the virtual-memory-management driver
API (`cuMemCreate`, not cudaMalloc), exported as a POSIX file descriptor, and a multicast object that every member
adds its device to and binds its allocation to (`cuMulticastCreate / AddDevice / BindMem`).

Pieces:
  * `FdServer`  one Unix-domain socket per peer that hands out file descriptors (SCM_RIGHTS) and runs the small
                "everybody has added its device / bound its memory" barriers of a group set-up. Its path travels in the
                peer's matchmaking blob, like the CUDA-IPC handle of the symmetric buffer.
  * `NvlsWindow` the VMM allocation of one peer: [payload | result] of one flat vector plus a flag page.
  * `NvlsGroup`  the multicast object of one group composition and this peer's mapping of it; cached per composition
                by the caller (elastic membership: a new composition costs a few milliseconds once).
Driver calls go through cuda-python (`cuda.bindings.driver`); nothing here touches the CUDA-IPC symmetric buffer.
Only plain `multimem.*` loads / stores / reductions are issued on multicast addresses (no TMA).
"""
from __future__ import annotations

import json
import logging
import os
import socket
import struct
import tempfile
import threading
import time
from typing import Dict, List, Optional, Sequence

logger = logging.getLogger(__name__)

try:  # cuda-python is optional: without it (or without a multicast-capable device) the P2P kernels are used
    from cuda.bindings import driver as cu
except Exception:  # pragma: no cover
    cu = None


class NvlsError(RuntimeError):
    pass


def _ck(res, what: str):
    err, rest = res[0], res[1:]
    if int(err) != 0:
        raise NvlsError(f"{what} failed: {err!r}")
    return rest[0] if len(rest) == 1 else rest


def multicast_supported(device: int) -> bool:
    if cu is None:
        return False
    try:
        _ck(cu.cuInit(0), "cuInit")
        dev = _ck(cu.cuDeviceGet(device), "cuDeviceGet")
        v = _ck(cu.cuDeviceGetAttribute(cu.CUdevice_attribute.CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev), "attr")
        return bool(v)
    except Exception:
        return False


# --------------------------------------------------------------------------------------------------- fd exchange
class FdServer:
    """Serves file descriptors and counting barriers to the other peers of this host over a Unix-domain socket."""

    def __init__(self):
        self.path = os.path.join(tempfile.gettempdir(), f"dedloc_nvls_{os.getpid()}_{os.urandom(4).hex()}.sock")
        self._fds: Dict[str, int] = {}
        self._counts: Dict[str, int] = {}
        self._cv = threading.Condition()
        self._sock = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
        self._sock.bind(self.path)
        self._sock.listen(64)
        self._stop = False
        threading.Thread(target=self._serve, daemon=True).start()

    def publish(self, key: str, fd: int) -> None:
        with self._cv:
            self._fds[key] = fd
            self._cv.notify_all()

    def _serve(self):
        while not self._stop:
            try:
                conn, _ = self._sock.accept()
            except OSError:
                return
            threading.Thread(target=self._handle, args=(conn,), daemon=True).start()

    def _handle(self, conn: socket.socket):
        try:
            with conn:
                req = json.loads(conn.recv(4096).decode())
                timeout = float(req.get("timeout", 20.0))
                if req["op"] == "fd":
                    with self._cv:
                        ok = self._cv.wait_for(lambda: req["key"] in self._fds, timeout)
                        fd = self._fds.get(req["key"])
                    if not ok:
                        conn.sendall(b"E")
                        return
                    socket.send_fds(conn, [b"F"], [fd])
                elif req["op"] == "arrive":      # counting barrier: answers once `n` members have arrived at `key`
                    with self._cv:
                        self._counts[req["key"]] = self._counts.get(req["key"], 0) + 1
                        self._cv.notify_all()
                        ok = self._cv.wait_for(lambda: self._counts.get(req["key"], 0) >= int(req["n"]), timeout)
                    conn.sendall(b"K" if ok else b"E")
        except Exception as e:  # pragma: no cover
            logger.debug("nvls fd server: %r", e)

    def close(self):
        self._stop = True
        try:
            self._sock.close()
            os.unlink(self.path)
        except OSError:
            pass


def fetch_fd(path: str, key: str, timeout: float = 20.0) -> int:
    with socket.socket(socket.AF_UNIX, socket.SOCK_STREAM) as s:
        s.settimeout(timeout + 5)
        s.connect(path)
        s.sendall(json.dumps({"op": "fd", "key": key, "timeout": timeout}).encode())
        msg, fds, _, _ = socket.recv_fds(s, 16, 1)
        if msg != b"F" or not fds:
            raise NvlsError(f"peer at {path} has no descriptor '{key}'")
        return fds[0]


def arrive(path: str, key: str, n: int, timeout: float = 20.0) -> None:
    with socket.socket(socket.AF_UNIX, socket.SOCK_STREAM) as s:
        s.settimeout(timeout + 5)
        s.connect(path)
        s.sendall(json.dumps({"op": "arrive", "key": key, "n": n, "timeout": timeout}).encode())
        if s.recv(1) != b"K":
            raise NvlsError(f"barrier '{key}' timed out")


# --------------------------------------------------------------------------------------------------- VMM memory
def _alloc_prop(device: int):
    prop = cu.CUmemAllocationProp()
    prop.type = cu.CUmemAllocationType.CU_MEM_ALLOCATION_TYPE_PINNED
    prop.location.type = cu.CUmemLocationType.CU_MEM_LOCATION_TYPE_DEVICE
    prop.location.id = device
    prop.requestedHandleTypes = cu.CUmemAllocationHandleType.CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR
    return prop


def _map_rw(handle, size: int, device: int) -> int:
    va = _ck(cu.cuMemAddressReserve(size, 0, 0, 0), "cuMemAddressReserve")
    _ck(cu.cuMemMap(va, size, 0, handle, 0), "cuMemMap")
    acc = cu.CUmemAccessDesc()
    acc.location.type = cu.CUmemLocationType.CU_MEM_LOCATION_TYPE_DEVICE
    acc.location.id = device
    acc.flags = cu.CUmemAccess_flags.CU_MEM_ACCESS_FLAGS_PROT_READWRITE
    _ck(cu.cuMemSetAccess(va, size, [acc], 1), "cuMemSetAccess")
    return int(va)


def _mc_prop(n_devices: int, size: int):
    prop = cu.CUmulticastObjectProp()
    prop.numDevices = n_devices
    prop.size = size
    prop.handleTypes = cu.CUmemAllocationHandleType.CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR
    prop.flags = 0
    return prop


class NvlsWindow:
    """One peer's multicast-bindable memory: [flags 4 KiB | payload 2 B/elem | result 2 B/elem], rounded up to the
    multicast granularity. `ptr` is the ordinary (unicast) address in this process."""

    FLAG_BYTES = 4096

    def __init__(self, n_elems: int, device: int, server: FdServer):
        if not multicast_supported(device):
            raise NvlsError("this device has no multicast (NVLS) support")
        import torch

        self.device, self.n, self.server = device, int(n_elems), server
        with torch.cuda.device(device):
            torch.cuda.current_stream().synchronize()        # make sure the primary context exists
            gran = _ck(cu.cuMulticastGetGranularity(_mc_prop(2, 1 << 21),
                                                    cu.CUmulticastGranularity_flags.CU_MULTICAST_GRANULARITY_RECOMMENDED),
                       "cuMulticastGetGranularity")
            need = self.FLAG_BYTES + 4 * self.n
            self.size = (need + gran - 1) // gran * gran
            self.handle = _ck(cu.cuMemCreate(self.size, _alloc_prop(device), 0), "cuMemCreate")
            self.ptr = _map_rw(self.handle, self.size, device)
            _ck(cu.cuMemsetD8(cu.CUdeviceptr(self.ptr), 0, self.size), "cuMemsetD8")
        self.off_flags, self.off_payload, self.off_result = 0, self.FLAG_BYTES, self.FLAG_BYTES + 2 * self.n
        self.key = f"win-{os.getpid()}-{id(self):x}"
        fd = _ck(cu.cuMemExportToShareableHandle(
            self.handle, cu.CUmemAllocationHandleType.CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0), "export window")
        server.publish(self.key, int(fd))

    def tensor(self, offset: int, numel: int, dtype):
        import torch

        from .symm import _CudaArray

        esz = torch.empty((), dtype=dtype).element_size()
        raw = torch.as_tensor(_CudaArray(self.ptr + offset, numel * esz, self), device=f"cuda:{self.device}")
        return raw.view(dtype)


class NvlsGroup:
    """The multicast object of one group composition, as seen by this member.

    Set-up (every member calls the constructor with the same ordered `socks`): member 0 creates the object and
    publishes its descriptor; everybody imports it, adds its device, waits until all have (binding before that is an
    error), binds its window, waits again, and maps the object. `mc_ptr` then addresses ALL members' windows at
    once: loads through it must be `multimem.ld_reduce`, stores `multimem.st` / `multimem.red`."""

    def __init__(self, window: NvlsWindow, socks: Sequence[str], me: int, group_key: str, timeout: float = 30.0):
        import torch

        self.window, self.P, self.me = window, len(socks), me
        size, dev = window.size, window.device
        leader = socks[0]
        with torch.cuda.device(dev):
            if me == 0:
                self.mc = _ck(cu.cuMulticastCreate(_mc_prop(self.P, size)), "cuMulticastCreate")
                fd = _ck(cu.cuMemExportToShareableHandle(
                    self.mc, cu.CUmemAllocationHandleType.CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0), "export mc")
                window.server.publish("mc-" + group_key, int(fd))
            else:
                fd = fetch_fd(leader, "mc-" + group_key, timeout)
                self.mc = _ck(cu.cuMemImportFromShareableHandle(
                    fd, cu.CUmemAllocationHandleType.CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR), "import mc")
                os.close(fd)
            cudev = _ck(cu.cuDeviceGet(dev), "cuDeviceGet")
            _ck(cu.cuMulticastAddDevice(self.mc, cudev), "cuMulticastAddDevice")
            arrive(leader, "added-" + group_key, self.P, timeout)
            _ck(cu.cuMulticastBindMem(self.mc, 0, window.handle, 0, size, 0), "cuMulticastBindMem")
            arrive(leader, "bound-" + group_key, self.P, timeout)
            self.mc_ptr = _map_rw(self.mc, size, dev)
        self.round = 0        # the barrier counters in the flag page only ever grow


# --------------------------------------------------------------------------------------------------- the operation
class NvlsAllreduce:
    """Weighted mean of a flat fp32 vector across a group, reduced inside the NVSwitch (csrc/nvls_allreduce.cu).

    One instance per peer and vector length. `group()` builds (once per composition) the multicast object;
    `allreduce()` encodes `src * weight` as fp16 / bf16 into the window and launches the kernel. The result lands in
    the window as bf16 and, if asked, in `out16` / `out32`."""

    def __init__(self, n_elems: int, device: int, server: Optional[FdServer] = None):
        import torch

        from .. import lib as native

        self.native = native
        self.server = server or FdServer()
        self.window = NvlsWindow(n_elems, device, self.server)
        self.device, self.n = device, int(n_elems)
        self._groups: Dict[str, NvlsGroup] = {}
        self.status = torch.zeros(16, dtype=torch.int32, device=f"cuda:{device}")
        self.payload16 = self.window.tensor(self.window.off_payload, 2 * self.n, torch.uint8)
        self.result = self.window.tensor(self.window.off_result, self.n, torch.bfloat16)
        self._aux = torch.zeros(1, dtype=torch.float32, device=f"cuda:{device}")
        props = torch.cuda.get_device_properties(device)
        self.ctas = props.multi_processor_count * 4

    @property
    def sock(self) -> str:
        return self.server.path

    def group(self, socks: Sequence[str], me: int, salt: str = "", timeout: float = 30.0) -> NvlsGroup:
        """`socks`: the members' FdServer paths in group order; `salt` tells apart several windows of the same peers
        (every member must pass the same string)."""
        import hashlib

        key = hashlib.sha1(("|".join(socks) + "#" + salt).encode()).hexdigest()[:16]
        g = self._groups.get(key)
        if g is None:
            g = self._groups[key] = NvlsGroup(self.window, list(socks), me, key, timeout)
        return g

    def allreduce(self, src, weight: float, group: NvlsGroup, out16=None, out32=None, bf16_payload: bool = False,
                  timeout_s: float = 20.0, stream=None) -> None:
        """result = sum over members of (src_i * weight_i); pass normalised weights for a weighted mean."""
        import ctypes as C

        import torch

        from ..ops import flat_ops
        from ..utils.compression import CompressionType

        native = self.native
        codec = CompressionType.BFLOAT16 if bf16_payload else CompressionType.FLOAT16
        with torch.cuda.device(self.device):
            flat_ops.encode(codec, src, float(weight), self.payload16, self._aux)
            group.round += 1
            p = native.NvlsParams()
            p.mc_base, p.uc_base, p.n = group.mc_ptr, self.window.ptr, self.n
            p.off_flags, p.off_payload, p.off_result = self.window.off_flags, self.window.off_payload, self.window.off_result
            p.P, p.me, p.round, p.ctas = group.P, group.me, group.round, self.ctas
            p.payload_is_bf16 = int(bf16_payload)
            p.timeout_ns = int(timeout_s * 1e9)
            p.status = self.status.data_ptr()
            p.out16 = out16.data_ptr() if out16 is not None else None
            p.out32 = out32.data_ptr() if out32 is not None else None
            st = stream if stream is not None else torch.cuda.current_stream(self.device)
            native.check(native.lib().dedloc_nvls_allreduce_mean(C.byref(p), st.cuda_stream), "nvls_allreduce")
