"""Symmetric peer-memory arena: the B200 replacement for Horovod's fusion buffer + NCCL/MPI ops.

Every rank allocates identical *segments* with ``cudaMalloc`` inside ``libb200dp``; the 64-byte
CUDA IPC handles are exchanged once over the gloo control plane and each peer's segment is mapped
into this process, after which collectives are single kernels that load/store peer memory over
NVLink (``csrc/comm.cu``).  Reference call sites being served: ``hvd.init()``
(``horovod/tensorflow_mnist.py:90``), gradient allreduce (``:133-136``), variable broadcast (``:143``),
metric averaging (``horovod/tensorflow_mnist_gpu.py:153``).

``SymmetricArena.emulated(n)`` builds all ``n`` ranks on ONE GPU (cooperative launches, blockIdx.y
= rank) so the flag protocol, the reductions and the fused Adam can be tested on a single B200.
"""
from __future__ import annotations

import ctypes
import math
import os
from typing import Dict, Iterable, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist

from ..ops import kernels as K
from . import autoselect
from .collectives import ReduceOp

CH_GENERIC = 15
CH_BCAST = 14
CH_ADASUM = 13
CH_GATHER = 12


class _Raw:
    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


def _round_up(x: int, a: int) -> int:
    return (x + a - 1) // a * a


class SymmetricArena:
    def __init__(self, handle, rank: int, nranks: int, emulated: bool, device: torch.device):
        self.h = handle
        self.rank, self.n, self.emulated, self.device = rank, nranks, emulated, device
        self.seg_bytes: List[int] = []
        self._views: Dict[Tuple[int, int], torch.Tensor] = {}
        self._staging: Optional[int] = None
        self._staging_elems = 0
        self._parity = 0
        self._closed = False
        pad = K.lib().b200_comm_pad_bytes()
        self.pad_seg = self.alloc(_round_up(pad, 256))
        assert self.pad_seg == 0

    # ------------------------------------------------------------------ construction
    @classmethod
    def create(cls, world, nbytes: Optional[int] = None) -> "SymmetricArena":
        h = ctypes.c_void_p()
        K.check(K.lib().b200_comm_create(world.rank, world.size, 0, ctypes.byref(h)), "comm_create")
        self = cls(h, world.rank, world.size, False, world.device)
        # How long a kernel spins at a peer-memory barrier before it reports the missing rank and traps.  Peers wait
        # there while rank 0 alone writes a checkpoint or runs an export, so the default sits well above such pauses.
        ms = os.environ.get("B200DP_COLLECTIVE_TIMEOUT_MS")
        if ms is None and os.environ.get("B200DP_COLLECTIVE_TIMEOUT_S"):
            ms = float(os.environ["B200DP_COLLECTIVE_TIMEOUT_S"]) * 1e3
        self.set_timeout_ms(float(ms) if ms is not None else 300e3)
        return self

    @classmethod
    def emulated(cls, nranks: int, device: Optional[torch.device] = None) -> "SymmetricArena":
        h = ctypes.c_void_p()
        K.check(K.lib().b200_comm_create(0, nranks, 1, ctypes.byref(h)), "comm_create(emulated)")
        return cls(h, 0, nranks, True, device or torch.device("cuda", torch.cuda.current_device()))

    def set_timeout_ms(self, ms: float) -> None:
        K.check(K.lib().b200_comm_set_timeout_ms(self.h, float(ms)), "set_timeout")

    def set_debug_skew_ns(self, max_ns: int) -> None:
        """Race hunting: every barrier arrival is delayed by a pseudo-random time below ``max_ns``."""
        K.check(K.lib().b200_comm_set_debug_skew_ns(self.h, int(max_ns)), "set_debug_skew")

    def set_debug_poison(self, on: bool) -> None:
        """Race hunting: the fused kernels fill a gradient buffer they have consumed with NaN instead of zeros."""
        K.check(K.lib().b200_comm_set_debug_poison(self.h, int(bool(on))), "set_debug_poison")

    def alloc(self, nbytes: int) -> int:
        """Collective: every rank must call with the same size, in the same order."""
        nbytes = _round_up(int(nbytes), 256)
        seg = ctypes.c_int(-1)
        hbuf = ctypes.create_string_buffer(64)
        K.check(K.lib().b200_comm_seg_alloc(self.h, nbytes, ctypes.byref(seg), hbuf), "seg_alloc")
        if not self.emulated and self.n > 1:
            handles: List[bytes] = [b""] * self.n
            dist.all_gather_object(handles, hbuf.raw)
            blob = b"".join(handles)
            K.check(K.lib().b200_comm_seg_open(self.h, seg.value, ctypes.c_char_p(blob)), "seg_open")
            dist.barrier()
        self.seg_bytes.append(nbytes)
        return seg.value

    # ------------------------------------------------------------------ NVLS (multicast) segments
    def nvls_supported(self) -> bool:
        return (not self.emulated) and self.n > 1 and bool(K.lib().b200_nvls_supported())

    def _exchange_fds(self, mine: List[int]) -> List[List[int]]:
        """All-to-all of POSIX file descriptors over AF_UNIX + SCM_RIGHTS (abstract socket names, no files)."""
        import socket
        self._fdx_round = getattr(self, "_fdx_round", 0) + 1
        base = f"\0b200dp-{os.environ.get('MASTER_PORT', '0')}-{self._fdx_round}"
        srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
        srv.bind(f"{base}-{self.rank}")
        srv.listen(self.n)
        dist.barrier()
        for p in range(self.n):
            if p == self.rank:
                continue
            c = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
            c.connect(f"{base}-{p}")
            socket.send_fds(c, [bytes([self.rank])], list(mine))
            c.close()
        got: List[List[int]] = [[] for _ in range(self.n)]
        got[self.rank] = list(mine)
        for _ in range(self.n - 1):
            conn, _ = srv.accept()
            msg, fds, _, _ = socket.recv_fds(conn, 16, 8)
            got[msg[0]] = list(fds)
            conn.close()
        srv.close()
        dist.barrier()
        return got

    def alloc_multicast(self, nbytes: int) -> int:
        """Collective: a symmetric segment bound to an NVSwitch multicast object (plus the usual peer mappings)."""
        if not self.nvls_supported():
            raise K.B200Error("NVLS multicast is not supported by this device/driver/world")
        seg, mem_fd, mc_fd = ctypes.c_int(-1), ctypes.c_int(-1), ctypes.c_int(-1)
        actual = ctypes.c_size_t(0)
        K.check(K.lib().b200_comm_mcseg_alloc(self.h, _round_up(int(nbytes), 256), ctypes.byref(seg), ctypes.byref(mem_fd),
                                              ctypes.byref(mc_fd), ctypes.byref(actual)), "mcseg_alloc")
        mine = [mem_fd.value] + ([mc_fd.value] if self.rank == 0 else [])
        got = self._exchange_fds(mine)
        mem_fds = (ctypes.c_int * self.n)(*[g[0] if g else -1 for g in got])
        K.check(K.lib().b200_comm_mcseg_open(self.h, seg.value, mem_fds, got[0][1] if self.rank != 0 else -1), "mcseg_open")
        os.close(mem_fd.value)
        if self.rank == 0:
            os.close(mc_fd.value)
        dist.barrier()                       # every device joined the team before anyone binds
        K.check(K.lib().b200_comm_mcseg_bind(self.h, seg.value), "mcseg_bind")
        dist.barrier()
        self.seg_bytes.append(int(actual.value))
        self.mc_segs = getattr(self, "mc_segs", set()) | {seg.value}
        return seg.value

    def ptr(self, seg: int, peer: Optional[int] = None) -> int:
        p = K.lib().b200_comm_seg_ptr(self.h, seg, self.rank if peer is None else peer)
        if not p:
            raise K.B200Error(f"segment {seg} / peer {peer} not mapped")
        return int(p)

    def bytes_view(self, seg: int, peer: Optional[int] = None) -> torch.Tensor:
        """uint8 tensor aliasing a whole segment.  ``peer`` selects another rank's copy — only meaningful in an
        emulated communicator (all copies on this GPU); real peer memory is reached by kernels and by
        :meth:`pull_from_peer`, never through torch tensors (torch would snapshot foreign-device memory)."""
        if peer is not None and peer != self.rank and not self.emulated:
            raise K.B200Error("peer views exist only in emulated communicators; use pull_from_peer()")
        key = (seg, self.rank if peer is None else peer)
        v = self._views.get(key)
        if v is None:
            v = torch.as_tensor(_Raw(self.ptr(seg, key[1]), self.seg_bytes[seg]), device=self.device)
            self._views[key] = v
        return v

    def view(self, seg: int, offset: int, numel: int, dtype: torch.dtype, peer: Optional[int] = None) -> torch.Tensor:
        es = torch.empty((), dtype=dtype).element_size()
        return self.bytes_view(seg, peer)[offset:offset + numel * es].view(dtype)

    def close(self) -> None:
        if self._closed:
            return
        self._closed = True
        self._views.clear()
        try:
            torch.cuda.synchronize()
        except Exception:
            pass
        K.lib().b200_comm_destroy(self.h)

    # ------------------------------------------------------------------ fused step
    def fused_allreduce_adam(self, *, channel: int, nblocks: int, nthreads: int, reduce: bool, grad_seg: int,
                             grad_cur_off: int, grad_other_off: int, state_seg: int, master_off: int, m_off: int, v_off: int,
                             w16_seg: int, w16_off: int, ctrl_seg: int, step_off: int, hyper_off: int, scale: float,
                             ranges: Sequence[Tuple[int, int, int, int]], stream: Optional[int] = None) -> None:
        n = len(ranges)
        LL = ctypes.c_longlong * n
        II = ctypes.c_int * n
        st = ctypes.c_void_p(stream if stream is not None else torch.cuda.current_stream().cuda_stream)
        K.check(K.lib().b200_fused_allreduce_adam(
            self.h, channel, nblocks, nthreads, int(reduce), grad_seg, grad_cur_off, grad_other_off, state_seg, master_off,
            m_off, v_off, w16_seg, w16_off, ctrl_seg, step_off, hyper_off, float(scale), n,
            LL(*[r[0] for r in ranges]), LL(*[r[1] for r in ranges]), II(*[r[2] for r in ranges]), II(*[r[3] for r in ranges]), st),
            "fused_allreduce_adam")

    # ------------------------------------------------------------------ generic collectives (staged)
    def _ensure_staging(self, elems: int) -> None:
        if self._staging is not None and elems <= self._staging_elems:
            return
        if self._staging is not None:
            raise K.B200Error(f"staging buffer of {self._staging_elems} elements too small for {elems}; "
                              f"raise B200DP_FUSION_THRESHOLD")
        from .optimizer import fusion_threshold_bytes
        cap = max(fusion_threshold_bytes() // 4, 1 << 16)
        self._staging_elems = _round_up(cap, 64)
        # [src parity 0][src parity 1][dst][adasum ping][adasum pong][scratch]
        # B200DP_BCAST=nvls puts the staging buffer behind an NVSwitch multicast object: large broadcasts then leave
        # the root once and are replicated by the switch
        self._staging_mc = bool(os.environ.get("B200DP_BCAST", "") == "nvls" and self.nvls_supported())
        alloc = self.alloc_multicast if self._staging_mc else self.alloc
        self._staging = alloc(self._staging_elems * 4 * 5 + (1 << 20))

    def _launch_shape(self, numel: int) -> Tuple[int, int]:
        nb = max(1, min(32, numel // 8192))
        return nb, 512

    def _stage_chunks(self, numel: int) -> Iterable[Tuple[int, int]]:
        # never ask for more than one chunk: a tensor above the staging capacity is what this loop exists for
        self._ensure_staging(min(numel, self._staging_elems) if self._staging is not None else min(numel, 1 << 24))
        off = 0
        while off < numel:
            n = min(self._staging_elems, numel - off)
            yield off, n
            off += n

    def allreduce_(self, tensor: torch.Tensor, op: ReduceOp, segments: Optional[Iterable[int]] = None) -> torch.Tensor:
        if self.emulated:
            raise K.B200Error("use the explicit emulated test helpers; hvd.allreduce_ needs a real communicator")
        if not tensor.is_floating_point():
            cpu = tensor.cpu()
            dist.all_reduce(cpu)
            if op is ReduceOp.Average:
                cpu //= self.n
            tensor.copy_(cpu)
            return tensor
        flat = tensor.reshape(-1) if tensor.is_contiguous() else tensor.contiguous().view(-1)
        if op is ReduceOp.Adasum:
            self._adasum_(flat, list(segments) if segments is not None else [flat.numel()])
        else:
            scale = 1.0 / self.n if op is ReduceOp.Average else 1.0
            for off, n in self._stage_chunks(flat.numel()):
                n4 = _round_up(n, 4)
                src_off = self._parity * self._staging_elems * 4
                dst_off = 2 * self._staging_elems * 4
                self._parity ^= 1
                src = self.view(self._staging, src_off, n4, torch.float32)
                src[:n].copy_(flat[off:off + n])
                if n4 != n:
                    src[n:].zero_()
                algo = autoselect.choose(n4 * 4, self.n)
                nb, nt = self._launch_shape(n4)
                K.check(K.lib().b200_allreduce(self.h, CH_GENERIC, nb, nt, algo, self._staging, src_off, self._staging, dst_off,
                                               n4, scale, ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)), "allreduce")
                flat[off:off + n].copy_(self.view(self._staging, dst_off, n, torch.float32))
        if flat.data_ptr() != tensor.data_ptr():
            tensor.copy_(flat.view_as(tensor))
        return tensor

    def broadcast_(self, tensor: torch.Tensor, root: int) -> torch.Tensor:
        if self.emulated:
            raise K.B200Error("hvd.broadcast_ needs a real communicator")
        raw = tensor.contiguous().view(-1).view(torch.uint8) if tensor.is_contiguous() else None
        if raw is None:
            tmp = tensor.contiguous()
            self.broadcast_(tmp, root)
            tensor.copy_(tmp)
            return tensor
        nbytes = raw.numel()
        for off, n in self._stage_chunks((nbytes + 3) // 4):
            b0, b1 = off * 4, min(nbytes, (off + n) * 4)
            n4 = _round_up(n, 4)
            src_off = self._parity * self._staging_elems * 4
            dst_off = 2 * self._staging_elems * 4
            self._parity ^= 1
            if self.rank == root:
                self.bytes_view(self._staging)[src_off:src_off + (b1 - b0)].copy_(raw[b0:b1])
            nb, nt = self._launch_shape(n4)
            st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
            if getattr(self, "_staging_mc", False) and n4 * 4 >= (1 << 20):
                K.check(K.lib().b200_broadcast_nvls(self.h, CH_BCAST, max(nb, 64), nt, root, self._staging, src_off, self._staging,
                                                    dst_off, n4, st), "broadcast_nvls")
            else:
                K.check(K.lib().b200_broadcast(self.h, CH_BCAST, nb, nt, root, self._staging, src_off, self._staging, dst_off, n4,
                                               st), "broadcast")
            if self.rank != root:
                raw[b0:b1].copy_(self.bytes_view(self._staging)[dst_off:dst_off + (b1 - b0)])
        return tensor

    def adasum_levels(self, seg: int, ping_off: int, pong_off: int, scratch_off: int, seg_start: Sequence[int],
                      seg_numel: Sequence[int], channel: int = CH_ADASUM, nblocks: int = 16,
                      stream: Optional[int] = None) -> int:
        """Run all log2(n) Adasum levels on a vector living at ``ping_off`` of ``seg``.
        Returns the byte offset (ping or pong) holding the result."""
        ns = len(seg_start)
        LL = ctypes.c_longlong * ns
        cur, nxt = ping_off, pong_off
        st = ctypes.c_void_p(stream if stream is not None else torch.cuda.current_stream().cuda_stream)
        for level in range(int(math.log2(self.n))):
            K.check(K.lib().b200_adasum_level(self.h, channel, nblocks, level, seg, cur, seg, nxt, seg, scratch_off, ns,
                                              LL(*seg_start), LL(*seg_numel), st), "adasum_level")
            cur, nxt = nxt, cur
        return cur

    def _adasum_(self, flat: torch.Tensor, seg_numel: List[int]) -> None:
        lim = 32                                   # tensors per kernel call (B200_MAX_SEGMENTS)
        if len(seg_numel) > lim:
            # a bucket with more tensors than one call takes: reduce it in runs of `lim` tensors
            off = 0
            for i in range(0, len(seg_numel), lim):
                part = seg_numel[i:i + lim]
                m = sum(part)
                self._adasum_(flat[off:off + m], part)
                off += m
            return
        n = flat.numel()
        self._ensure_staging(n)
        if n > self._staging_elems:
            raise K.B200Error("Adasum tensor larger than the staging buffer; raise B200DP_FUSION_THRESHOLD")
        E = self._staging_elems * 4
        ping, pong, scratch = 3 * E, 4 * E, 5 * E
        self.view(self._staging, ping, n, torch.float32).copy_(flat)
        starts, off = [], 0
        for s in seg_numel:
            starts.append(off)
            off += s
        res = self.adasum_levels(self._staging, ping, pong, scratch, starts, seg_numel)
        flat.copy_(self.view(self._staging, res, n, torch.float32))

    def pull_from_peer(self, seg: int, off: int, nbytes: int, peer: int) -> None:
        """One-sided: local[off:off+nbytes] = peer[off:off+nbytes] (cudaMemcpy over the IPC mapping)."""
        K.check(K.lib().b200_comm_pull_from_peer(self.h, seg, off, nbytes, peer,
                                                 ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)), "pull_from_peer")

    def allgather_slices_(self, seg: int, off: int, numel: int, nblocks: int = 16) -> None:
        K.check(K.lib().b200_allgather_slices(self.h, CH_GATHER, nblocks, 512, seg, off, numel,
                                              ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)), "allgather_slices")
