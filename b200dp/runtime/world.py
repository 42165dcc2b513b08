"""World bootstrap: rank/size discovery, device pinning, control-plane process group.

Replaces ``hvd.init()`` + the MPI wire-up it triggers (reference call sites
``horovod/tensorflow_mnist.py:90``, ``horovod/tensorflow_mnist_gpu.py:93``) and the
device pinning blocks (``horovod/tensorflow_mnist.py:152-155``,
``horovod/tensorflow_mnist_gpu.py:95-101``).

Design: one process per GPU.  The *control plane* (handle exchange, object
broadcast, CPU tensors) is a gloo process group over TCP; the *data plane* for
CUDA tensors is the peer-memory arena in :mod:`b200dp.parallel.arena`
(hand-written sm_100a kernels over NVLink), never NCCL.
"""
from __future__ import annotations

import datetime
import os
import socket
import threading
from dataclasses import dataclass, field
from typing import Optional

import torch
import torch.distributed as dist

_ENV_RANK = ("B200DP_RANK", "RANK", "OMPI_COMM_WORLD_RANK", "PMI_RANK")
_ENV_SIZE = ("B200DP_SIZE", "WORLD_SIZE", "OMPI_COMM_WORLD_SIZE", "PMI_SIZE")
_ENV_LOCAL_RANK = ("B200DP_LOCAL_RANK", "LOCAL_RANK", "OMPI_COMM_WORLD_LOCAL_RANK")
_ENV_LOCAL_SIZE = ("B200DP_LOCAL_SIZE", "LOCAL_WORLD_SIZE", "OMPI_COMM_WORLD_LOCAL_SIZE")


def _env_int(names, default):
    for n in names:
        v = os.environ.get(n)
        if v not in (None, ""):
            return int(v)
    return default


@dataclass
class World:
    rank: int = 0
    size: int = 1
    local_rank: int = 0
    local_size: int = 1
    device: torch.device = field(default_factory=lambda: torch.device("cpu"))
    backend: str = "single"          # "single" | "gloo" | "p2p"
    host: str = field(default_factory=socket.gethostname)
    owns_pg: bool = False
    arena: object = None             # parallel.arena.SymmetricArena when backend == "p2p"

    @property
    def is_cuda(self) -> bool:
        return self.device.type == "cuda"


_WORLD: Optional[World] = None
_LOCK = threading.Lock()


class NotInitializedError(RuntimeError):
    pass


def is_initialized() -> bool:
    return _WORLD is not None


def world() -> World:
    if _WORLD is None:
        raise NotInitializedError("b200dp has not been initialized; call b200dp.init() first")
    return _WORLD


def init(device: Optional[str] = None, *, timeout_s: float = 600.0, arena_bytes: Optional[int] = None) -> World:
    """Join the world created by the launcher.  Idempotent, like ``hvd.init()``.

    ``device``: "cuda", "cpu" or None (cuda when available).  Rank/size come
    from the launcher's environment: ``b200run`` / ``torchrun`` / ``mpirun``
    variable names are all understood.
    """
    global _WORLD
    with _LOCK:
        if _WORLD is not None:
            return _WORLD
        rank = _env_int(_ENV_RANK, 0)
        size = _env_int(_ENV_SIZE, 1)
        local_rank = _env_int(_ENV_LOCAL_RANK, rank)
        local_size = _env_int(_ENV_LOCAL_SIZE, size)
        if not (0 <= rank < size):
            raise ValueError(f"rank {rank} outside world of size {size}")

        want_cuda = device in (None, "cuda") and torch.cuda.is_available() and os.environ.get("B200DP_FORCE_CPU", "0") != "1"
        if device == "cuda" and not torch.cuda.is_available():
            raise RuntimeError("device='cuda' requested but no CUDA device is visible")
        if want_cuda:
            ndev = torch.cuda.device_count()
            # one GPU per process chosen by the local rank (reference: visible_device_list=str(local_rank))
            dev_index = local_rank % max(ndev, 1)
            torch.cuda.set_device(dev_index)
            dev = torch.device("cuda", dev_index)
        else:
            dev = torch.device("cpu")

        w = World(rank=rank, size=size, local_rank=local_rank, local_size=local_size, device=dev)
        if size > 1:
            if not dist.is_initialized():
                os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
                os.environ.setdefault("MASTER_PORT", "29512")
                dist.init_process_group(
                    backend="gloo", rank=rank, world_size=size,
                    timeout=datetime.timedelta(seconds=timeout_s))
                w.owns_pg = True
            w.backend = "gloo"
        _WORLD = w
        if size > 1 and w.is_cuda:
            _require_single_node(w)
            from ..parallel.arena import SymmetricArena
            w.arena = SymmetricArena.create(w, nbytes=arena_bytes)
            w.backend = "p2p"
        return w


def _require_single_node(w: World) -> None:
    """The CUDA data plane maps every peer's memory with CUDA IPC over one NVSwitch domain, so all ranks must share
    a host and fit the kernels' rank table.  Checked here so a multi-pod launch fails with a sentence instead of an
    opaque cudaIpcOpenMemHandle error (set B200DP_FORCE_CPU=1 for the gloo path across pods)."""
    hosts = [None] * w.size
    dist.all_gather_object(hosts, w.host)
    if len(set(hosts)) != 1:
        raise RuntimeError(f"b200dp: the GPU data plane is single-node (one NVSwitch domain), but ranks run on {sorted(set(hosts))}; "
                           f"use one worker pod with all GPUs, or B200DP_FORCE_CPU=1 / --cpu for the gloo path")
    max_ranks = 8                    # B200_MAX_RANKS in include/b200dp.h
    if w.size > max_ranks:
        raise RuntimeError(f"b200dp: world size {w.size} exceeds the {max_ranks} ranks the peer-memory kernels address")


def shutdown() -> None:
    global _WORLD
    with _LOCK:
        w = _WORLD
        if w is None:
            return
        if w.arena is not None:
            # Peers keep each other's segments mapped (CUDA IPC) and rank 0 may still be pulling sharded optimizer
            # state out of them for its final checkpoint/export: nobody frees until everybody is done.  A rank that
            # died never arrives; the launcher's fail-fast teardown (or the gloo timeout) ends the wait.
            try:
                torch.cuda.synchronize()
                if dist.is_initialized():
                    dist.barrier()
            except Exception:
                pass
            try:
                w.arena.close()
            except Exception:
                pass
        if w.owns_pg and dist.is_initialized():
            try:
                dist.destroy_process_group()
            except Exception:
                pass
        _WORLD = None


def rank() -> int:
    return world().rank


def size() -> int:
    return world().size


def local_rank() -> int:
    return world().local_rank


def local_size() -> int:
    return world().local_size


def device() -> torch.device:
    return world().device


def nccl_built() -> bool:
    """Capability probe kept under Horovod's name (``horovod/tensorflow_mnist.py:127``).

    Horovod answers "was the GPU collective path compiled in?".  Here that path
    is the sm_100a peer-memory library, so the probe is true when that library
    is present and a CUDA device is usable.
    """
    if not torch.cuda.is_available():
        return False
    from ..ops import _lib
    return _lib.available()


def mpi_built() -> bool:
    return False


def gloo_built() -> bool:
    return dist.is_gloo_available()


def barrier() -> None:
    w = world()
    if w.size == 1:
        return
    if w.arena is not None:
        torch.cuda.synchronize()
    dist.barrier()
