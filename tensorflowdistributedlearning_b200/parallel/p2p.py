"""Symmetric heap + peer-to-peer communicator (host side of ``csrc/p2p_allreduce.cu``).

Every rank allocates identically-sized buffers from the native heap (plain ``cudaMalloc`` so they
can be exported with ``cudaIpcGetMemHandle``), all-gathers the 64-byte handles through the
bootstrap process group (the *only* use of torch.distributed on the gradient path) and maps every
peer's buffer.  The resulting pointer tables feed the fused all-reduce + optimizer kernels.

This replaces the variable mirroring and cross-tower plumbing that MirroredStrategy does inside
TensorFlow for the reference (model.py:115-119).
"""
from __future__ import annotations

import ctypes
import os
from typing import List, Optional

import torch
import torch.distributed as dist

from .. import _native as N


_FD_ROUND = [0]  # descriptor exchanges so far (identical on every tower: socket names must pair up)


def exchange_fds(fd: Optional[int], rank: int, world: int, group=None, root: Optional[int] = None) -> List[Optional[int]]:
    """Pass file descriptors between the towers of one node over abstract UNIX sockets (SCM_RIGHTS).

    ``root=None``: all-gather -- every tower sends ``fd`` to every other tower; returns the list indexed by rank (own
    entry = ``fd``).  ``root=r``: broadcast -- only tower ``r`` sends; every other tower's list has just entry ``r``.
    Received descriptors are new descriptors of this process (the caller owns them).  Collective: all towers call it
    in the same order."""
    import socket
    import struct

    _FD_ROUND[0] += 1
    base = f"\0tdl_fd_{os.environ.get('MASTER_PORT', '0')}_{os.getuid()}_{_FD_ROUND[0]}"
    out: List[Optional[int]] = [None] * world
    if root is None or root == rank:
        out[rank] = fd
    senders = [r for r in range(world) if r != rank and (root is None or r == root)]
    srv = None
    if senders:
        srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
        srv.bind(f"{base}_{rank}")
        srv.listen(world)
    dist.barrier(group=group)  # every listener is up
    try:
        if root is None or root == rank:
            for r in range(world):
                if r == rank:
                    continue
                with socket.socket(socket.AF_UNIX, socket.SOCK_STREAM) as c:
                    c.connect(f"{base}_{r}")
                    socket.send_fds(c, [struct.pack("i", rank)], [int(fd)])
        for _ in senders:
            conn, _addr = srv.accept()
            with conn:
                msg, fds, _flags, _a = socket.recv_fds(conn, 4, 1)
                if len(msg) != 4 or len(fds) != 1:
                    raise RuntimeError("descriptor exchange: short message from a peer tower")
                out[struct.unpack("i", msg)[0]] = fds[0]
    finally:
        if srv is not None:
            srv.close()
    dist.barrier(group=group)
    return out


class _RawCuda:
    """Zero-copy view of a raw device pointer for torch.as_tensor."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


class SymmetricBuffer:
    """One allocation per rank, same size everywhere, mapped into every peer."""

    def __init__(self, nbytes: int, group=None, emulate_world: int = 0, multicast: bool = False):
        """``multicast=True`` allocates from this library's CUDA-VMM heap (``csrc/vmm_heap.cu``: cuMemCreate + POSIX
        descriptors passed between the towers + one cuMulticast object) so that the buffer also gets an NVLS multicast
        mapping (``self.mc_ptr``; 0 when the driver / fabric does not expose one).  ``TDL_SYMM_BACKEND=torch`` uses
        torch's symmetric-memory module for it instead (the earlier path; also the fallback when the driver has no
        VMM descriptors).  Without ``multicast`` the buffer comes from the plain cudaIpc heap."""
        lib = N.lib()
        self.nbytes = int(nbytes)
        self.group = group
        self.emulated = emulate_world > 0
        self.mc_ptr = 0
        self._torch_tensor = None
        self._torch_handle = None
        self._vmm_own = self._vmm_mc = None
        self._vmm_peers: List[ctypes.c_void_p] = []
        backend = os.environ.get("TDL_SYMM_BACKEND", "vmm")  # "torch": torch's symmetric memory (the round-1 path)
        if (multicast and backend != "torch" and not self.emulated and dist.is_initialized()
                and dist.get_world_size(group) > 1 and self._init_vmm(group)):
            return
        if multicast and not self.emulated and dist.is_initialized() and dist.get_world_size(group) > 1:
            import torch.distributed._symmetric_memory as symm

            self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
            t = symm.empty(self.nbytes, dtype=torch.uint8, device=torch.device("cuda", torch.cuda.current_device()))
            hdl = symm.rendezvous(t, group if group is not None else dist.group.WORLD)
            t.zero_()
            self._torch_tensor, self._torch_handle = t, hdl
            self.ptrs = [int(p) for p in hdl.buffer_ptrs]
            try:
                self.mc_ptr = int(hdl.multicast_ptr)
            except Exception:  # noqa: BLE001
                self.mc_ptr = 0
            self._own, self._opened = [], []
            self._holders = [_RawCuda(p, self.nbytes) for p in self.ptrs]
            torch.cuda.synchronize()
            dist.barrier(group=group)
            return
        self.world = emulate_world if self.emulated else (dist.get_world_size(group) if dist.is_initialized() else 1)
        self.rank = 0 if self.emulated else (dist.get_rank(group) if dist.is_initialized() else 0)
        self._own: List[int] = []
        self._opened: List[int] = []
        self.ptrs: List[int] = []
        if self.emulated or self.world == 1:
            for _ in range(self.world):
                p = ctypes.c_void_p()
                N.check(lib.tdl_symm_alloc(ctypes.c_int64(self.nbytes), ctypes.byref(p)), "tdl_symm_alloc")
                self._own.append(p.value)
            self.ptrs = list(self._own)
        else:
            p = ctypes.c_void_p()
            N.check(lib.tdl_symm_alloc(ctypes.c_int64(self.nbytes), ctypes.byref(p)), "tdl_symm_alloc")
            self._own.append(p.value)
            h = (ctypes.c_ubyte * 64)()
            N.check(lib.tdl_ipc_export(p, h), "tdl_ipc_export")
            handles: List[Optional[bytes]] = [None] * self.world
            dist.all_gather_object(handles, bytes(h), group=group)
            for r in range(self.world):
                if r == self.rank:
                    self.ptrs.append(p.value)
                else:
                    q = ctypes.c_void_p()
                    buf = (ctypes.c_ubyte * 64).from_buffer_copy(handles[r])
                    N.check(lib.tdl_ipc_open(buf, ctypes.byref(q)), "tdl_ipc_open")
                    self._opened.append(q.value)
                    self.ptrs.append(q.value)
        self._holders = [_RawCuda(p, self.nbytes) for p in self.ptrs]

    def _init_vmm(self, group) -> bool:
        """This library's own heap for multicast buffers (csrc/vmm_heap.cu): cuMemCreate allocations shared between the
        towers as POSIX descriptors, every peer mapped, and one NVLS multicast object over all of them.  Returns False
        (nothing allocated) when the driver lacks VMM descriptors -- the caller falls back."""
        lib = N.lib()
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        vmm_ok, mc_ok, gran = ctypes.c_int(0), ctypes.c_int(0), ctypes.c_int64(0)
        N.check(lib.tdl_vmm_query(ctypes.c_int(self.world), ctypes.byref(vmm_ok), ctypes.byref(mc_ok),
                                  ctypes.byref(gran)), "tdl_vmm_query")
        # every tower must take the same branch: agree on the minimum
        caps = torch.tensor([vmm_ok.value, mc_ok.value, -gran.value], dtype=torch.int64, device="cuda")
        dist.all_reduce(caps, op=dist.ReduceOp.MIN, group=group)
        vmm_all, mc_all, gran_all = int(caps[0]), int(caps[1]), int(-caps[2])
        if not vmm_all:
            return False
        padded = (self.nbytes + gran_all - 1) // gran_all * gran_all
        own, ptr, fd = ctypes.c_void_p(), ctypes.c_void_p(), ctypes.c_int(-1)
        N.check(lib.tdl_vmm_alloc(ctypes.c_int64(padded), ctypes.byref(own), ctypes.byref(ptr), ctypes.byref(fd)),
                "tdl_vmm_alloc")
        self._vmm_own, self._vmm_padded = own, padded
        fds = exchange_fds(fd.value, self.rank, self.world, group)
        self.ptrs = []
        for r in range(self.world):
            if r == self.rank:
                self.ptrs.append(ptr.value)
                continue
            a, q = ctypes.c_void_p(), ctypes.c_void_p()
            N.check(lib.tdl_vmm_import(ctypes.c_int(fds[r]), ctypes.c_int64(padded), ctypes.byref(a), ctypes.byref(q)),
                    "tdl_vmm_import")
            self._vmm_peers.append(a)
            self.ptrs.append(q.value)
        if mc_all:
            mc, mfd = ctypes.c_void_p(), ctypes.c_int(-1)
            if self.rank == 0:
                N.check(lib.tdl_vmm_mc_create(ctypes.c_int(self.world), ctypes.c_int64(padded), ctypes.byref(mc),
                                              ctypes.byref(mfd)), "tdl_vmm_mc_create")
            got = exchange_fds(mfd.value if self.rank == 0 else None, self.rank, self.world, group, root=0)
            if self.rank != 0:
                N.check(lib.tdl_vmm_mc_import(ctypes.c_int(got[0]), ctypes.c_int64(padded), ctypes.byref(mc)),
                        "tdl_vmm_mc_import")
            N.check(lib.tdl_vmm_mc_add_device(mc), "tdl_vmm_mc_add_device")
            dist.barrier(group=group)  # every device is registered before anyone binds memory
            mptr = ctypes.c_void_p()
            N.check(lib.tdl_vmm_mc_bind_map(mc, own, ctypes.byref(mptr)), "tdl_vmm_mc_bind_map")
            self._vmm_mc, self.mc_ptr = mc, int(mptr.value)
        self._own, self._opened = [], []
        self._holders = [_RawCuda(p, self.nbytes) for p in self.ptrs]
        self.tensor(torch.uint8).zero_()
        torch.cuda.synchronize()
        dist.barrier(group=group)
        return True

    def tensor(self, dtype=torch.float32, rank: Optional[int] = None) -> torch.Tensor:
        """This rank's buffer (or, in emulation, the buffer of ``rank``) as a flat tensor of ``dtype``."""
        r = self.rank if rank is None else rank
        if self._torch_tensor is not None and r == self.rank:
            return self._torch_tensor.view(dtype)
        t = torch.as_tensor(self._holders[r], device=torch.device("cuda", torch.cuda.current_device()))
        return t.view(dtype)

    def ptr_table(self):
        arr = (ctypes.c_void_p * self.world)()
        for r, p in enumerate(self.ptrs):
            arr[r] = p
        return arr

    def close(self):
        lib = N.lib()
        if self._vmm_mc is not None:
            lib.tdl_vmm_mc_free(self._vmm_mc)
            self._vmm_mc = None
        for a in self._vmm_peers:
            lib.tdl_vmm_free(a)
        self._vmm_peers = []
        if self._vmm_own is not None:
            lib.tdl_vmm_free(self._vmm_own)
            self._vmm_own = None
        for q in self._opened:
            lib.tdl_ipc_close(ctypes.c_void_p(q))
        for p in self._own:
            lib.tdl_symm_free(ctypes.c_void_p(p))
        self._opened, self._own = [], []


class P2PComm:
    """Pointer tables + signal pads for the fused kernels.  ``emulate_rank`` selects which emulated rank
    this communicator acts as when all buffers live on one GPU (single-GPU tests)."""

    def __init__(self, grad: SymmetricBuffer, param: Optional[SymmetricBuffer], group=None,
                 emulate_rank: Optional[int] = None):
        lib = N.lib()
        self.grad, self.param = grad, param
        self.world = grad.world
        self.emulated = grad.emulated
        self.rank = emulate_rank if self.emulated else grad.rank
        self.flags = SymmetricBuffer(int(lib.tdl_p2p_flag_bytes()), group, emulate_world=self.world if self.emulated else 0)
        self.epoch = torch.zeros(256, dtype=torch.int32, device="cuda")
        out = ctypes.c_void_p()
        N.check(lib.tdl_p2p_comm_create(ctypes.c_int(self.world), ctypes.c_int(self.rank), grad.ptr_table(),
                                        param.ptr_table() if param is not None else None, self.flags.ptr_table(),
                                        N.ptr(self.epoch), ctypes.c_int(int(self.emulated)), ctypes.byref(out)),
                "tdl_p2p_comm_create")
        self.handle = out
        self.wire = None
        # the device-side wait limit follows the host-side collective timeout (a hard-coded 2 s limit killed
        # every tower when one of them stalled in its input pipeline); TDL_P2P_TIMEOUT_S overrides it on its own
        limit = float(os.environ.get("TDL_P2P_TIMEOUT_S", os.environ.get("TDL_COLLECTIVE_TIMEOUT_S", "600")))
        N.check(lib.tdl_p2p_comm_set_timeout(self.handle, ctypes.c_double(limit)), "tdl_p2p_comm_set_timeout")
        self._debug_seen = None
        if os.environ.get("TDL_P2P_DEBUG", "0") == "1":
            # protocol assertions on every signal-pad poll (epochs monotone per (CTA, peer), at most one barrier ahead)
            self._debug_seen = torch.zeros(256 * 8, dtype=torch.int32, device="cuda")
            N.check(lib.tdl_p2p_comm_set_debug(self.handle, N.ptr(self._debug_seen)), "tdl_p2p_comm_set_debug")
        self.has_multicast = bool(grad.mc_ptr and param is not None and param.mc_ptr)
        if self.has_multicast:
            N.check(lib.tdl_p2p_comm_set_multicast(self.handle, ctypes.c_void_p(grad.mc_ptr),
                                                   ctypes.c_void_p(param.mc_ptr)), "tdl_p2p_comm_set_multicast")
        if not self.emulated and dist.is_initialized() and self.world > 1:
            torch.cuda.synchronize()
            dist.barrier(group=group)

    def set_wire(self, wire: SymmetricBuffer) -> None:
        """Attach the bf16 wire buffers (one per rank, gradient-bucket element offsets): enables ``fused_step_wire``."""
        self.wire = wire
        N.check(N.lib().tdl_p2p_comm_set_wire(self.handle, wire.ptr_table()), "tdl_p2p_comm_set_wire")

    def fused_step_wire(self, offset: int, n: int, *, kind: int, zero_grad: bool, m1, m2, lr_state, a: float, b: float,
                        c: float, grad_scale: float = 1.0, mode: int = 0, max_blocks: int = 64, phases: int = 3):
        """The fused step with gradients crossing NVLink as bf16: phase 1 packs (and clears) this rank's fp32 bucket into
        its wire buffer, phase 2 reduces every rank's wire copy in fp32 and runs the optimizer (``phases=3`` = both)."""
        rc = N.lib().tdl_p2p_fused_step_wire(self.handle, ctypes.c_int64(offset), ctypes.c_int64(n), ctypes.c_int(mode),
                                             ctypes.c_int(kind), ctypes.c_int(int(zero_grad)), N.ptr(m1), N.ptr(m2),
                                             N.ptr(lr_state), ctypes.c_float(a), ctypes.c_float(b), ctypes.c_float(c),
                                             ctypes.c_float(grad_scale), ctypes.c_int(max_blocks), ctypes.c_int(phases),
                                             N.stream_ptr())
        N.check(rc, "tdl_p2p_fused_step_wire")

    def fused_step(self, offset: int, n: int, *, kind: int, grad_is_bf16: bool, zero_grad: bool, m1, m2, lr_state,
                   a: float, b: float, c: float, grad_scale: float = 1.0, mode: int = 0, max_blocks: int = 64):
        rc = N.lib().tdl_p2p_fused_step(self.handle, ctypes.c_int64(offset), ctypes.c_int64(n), ctypes.c_int(mode),
                                        ctypes.c_int(kind), ctypes.c_int(int(grad_is_bf16)),
                                        ctypes.c_int(int(zero_grad)), N.ptr(m1), N.ptr(m2), N.ptr(lr_state),
                                        ctypes.c_float(a), ctypes.c_float(b), ctypes.c_float(c),
                                        ctypes.c_float(grad_scale), ctypes.c_int(max_blocks), N.stream_ptr())
        N.check(rc, "tdl_p2p_fused_step")

    def broadcast_params_(self, n: int, src: int = 0, offset: int = 0, max_blocks: int = 64) -> None:
        """X2 (initial parameter mirror) over NVLink: every rank pulls ``src``'s parameters with a copy kernel bracketed
        by the signal-pad barriers (``tdl_p2p_broadcast``); all ranks must call it."""
        rc = N.lib().tdl_p2p_broadcast(self.handle, ctypes.c_int(src), ctypes.c_int64(offset), ctypes.c_int64(n),
                                       ctypes.c_int(max_blocks), N.stream_ptr())
        N.check(rc, "tdl_p2p_broadcast")

    def all_reduce_mean_(self, staging: torch.Tensor, offset: int, n: int, max_blocks: int = 8):
        """In-place mean over ranks of grad[offset:offset+n] (fp32), one-shot.  ``staging`` is fp32 scratch."""
        self.fused_step(offset, n, kind=0, grad_is_bf16=False, zero_grad=False, m1=staging, m2=None, lr_state=None,
                        a=0.0, b=0.0, c=0.0, mode=1, max_blocks=max_blocks)

    def timed_out(self) -> Optional[dict]:
        """None, or who a kernel of this communicator gave up waiting for: {'rank', 'cta', 'peer', 'epoch'}.  Reads
        host memory only, so it still answers after the trap has left the CUDA context unusable."""
        if self.handle is None:
            return None
        out = (ctypes.c_int * 5)()
        N.lib().tdl_p2p_comm_status(self.handle, out)
        if not out[0]:
            return None
        why = {1: "timeout", 2: "signal went backwards", 3: "signal too far ahead"}.get(out[0], str(out[0]))
        return {"rank": out[1], "cta": out[2], "peer": out[3], "epoch": out[4], "why": why}

    def close(self):
        if self.handle is not None:
            N.lib().tdl_p2p_comm_destroy(self.handle)
            self.handle = None
        self.flags.close()
