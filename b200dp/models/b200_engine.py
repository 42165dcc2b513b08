"""The sm_100a training engine for both MNIST CNN families (``mnist_cnn`` as coded in the reference scripts, and
``keras128``, the Keras-example shape BASELINE.json's prose names).

One training step (the reference's ``mon_sess.run(train_op)``, ``horovod/tensorflow_mnist.py:168-171``;
Keras ``fit`` inner loop, ``horovod/tensorflow_mnist_gpu.py:173``) is a fixed sequence of hand-written
kernels replayed from a CUDA graph:

  fwd : conv1+bias+ReLU+pool (CUDA cores, K=25)  ->  conv2 implicit GEMM + bias+ReLU+pool (tcgen05)
        ->  dense GEMM, split-K into fp32 (tcgen05)  ->  head: bias+ReLU+dropout+logits+softmax-xent fwd/bwd
  bwd : head wgrad  ->  dense wgrad (tcgen05, written straight into the gradient arena)
        ->  dense dgrad (tcgen05) -> unpool+ReLU mask -> conv2 wgrad / dgrad (tcgen05) -> conv1 wgrad
  comm: per bucket ONE kernel = NVLink allreduce + 1/size + Adam + bf16 weight refresh; the dense
        bucket's kernel runs on a second stream underneath the convolution backward.

``keras128`` runs the same DAG with its own geometry: conv1 3x3 VALID + ReLU (CUDA cores, no pool) -> conv2 3x3 VALID
implicit GEMM + ReLU + pool (tcgen05, 24x24 output walked in 8x16 tiles) -> dropout 0.25 -> dense 9216x128 -> head
(dropout 0.5); the feature dropout's backward is folded into the unpool kernel (a dropped activation is a zero there).

All persistent state (fp32 master weights, Adam moments, two gradient buffers, bf16 working
weights, step counter, hyper-parameters, loss ring) lives in symmetric peer-mapped segments.
"""
from __future__ import annotations

import ctypes
import os
from typing import Any, Dict, List, Optional, Tuple

import numpy as np
import torch

from ..ops import kernels as K
from ..parallel import autoselect
from ..parallel import collectives as C
from ..parallel.arena import SymmetricArena
from ..parallel.buckets import plan_arena_buckets
from ..parallel.optimizer import Adam, DistributedAdam
from ..runtime import world as _w
from ..train.engine import Engine, StepResult, _as_dist_adam
from .spec import ModelSpec, get_spec, init_flat

bf16 = torch.bfloat16
RING = 1024
CTRL_STEP, CTRL_HYPER, CTRL_EVAL_LOSS, CTRL_EVAL_CORRECT, CTRL_LOSS_RING, CTRL_CORRECT_RING = 0, 16, 64, 80, 256, 256 + RING * 4
CTRL_POW, CTRL_COEF = 96, 128          # beta1^t / beta2^t (2 doubles); this step's Adam coefficients (5 floats)
CTRL_BYTES = 256 + RING * 8


class _LazyScalar:
    """Device scalar read on demand (keeps the training loop free of per-step syncs)."""

    def __init__(self, t: torch.Tensor, scale: float = 1.0):
        self.t, self.scale, self._v = t, scale, None

    def _get(self):
        if self._v is None:
            self._v = self.t.item() * self.scale
            self.t = None
        return self._v

    def __float__(self):
        return float(self._get())

    def __int__(self):
        return int(self._get())

    def item(self):
        return self._get()


class _HostScalar:
    """A step's loss / correct count in the pinned host ring the step's last kernel writes (no D2H copy in the stream):
    reading it waits for that step's completion event, nothing else."""

    def __init__(self, ring: torch.Tensor, slot: int, event: torch.cuda.Event, scale: float = 1.0):
        self.ring, self.slot, self.event, self.scale, self._v = ring, slot, event, scale, None

    def _get(self):
        if self._v is None:
            self.event.synchronize()
            self._v = self.ring[self.slot].item() * self.scale
        return self._v

    def __float__(self):
        return float(self._get())

    def __int__(self):
        return int(self._get())

    def item(self):
        return self._get()


class _PendingState:
    """Handle of an asynchronous state snapshot (B200Engine.state_async)."""

    def __init__(self, host: torch.Tensor, event: torch.cuda.Event, T: int, meta: Dict[str, Any]):
        self.host, self.event, self.T, self.meta = host, event, T, meta

    def result(self) -> Dict[str, Any]:
        self.event.synchronize()
        T, h = self.T, self.host
        return {"tensors": {"params": h[:T], "adam_m": h[T:2 * T], "adam_v": h[2 * T:3 * T]}, "meta": self.meta}


def _env_flag(name: str, default: str = "1") -> bool:
    return os.environ.get(name, default) not in ("0", "false", "no")


class B200Engine(Engine):
    def __init__(self, spec: Optional[ModelSpec] = None, optimizer=None, *, seed: int = 0, batch_size: int = 128,
                 use_graph: Optional[bool] = None, dropout: bool = True, use_tc: Optional[bool] = None,
                 comm_blocks: int = 296, arena: Optional[SymmetricArena] = None, **_ignored):
        self.spec = spec or get_spec()
        if self.spec.arch not in ("mnist_cnn", "keras128"):
            raise NotImplementedError(f"B200Engine implements mnist_cnn and keras128; use engine='torch' for {self.spec.arch}")
        self.keras = self.spec.arch == "keras128"
        if not torch.cuda.is_available():
            raise RuntimeError("B200Engine needs a CUDA device")
        K.lib()                                     # fail loudly if the extension is missing
        self.device = _w.world().device if _w.is_initialized() and _w.world().is_cuda else torch.device("cuda", torch.cuda.current_device())
        self.seed = seed
        self.cap = int(batch_size)
        self.dropout = dropout
        self.use_graph = _env_flag("B200DP_GRAPH") if use_graph is None else use_graph
        tc = _env_flag("B200DP_TC") if use_tc is None else use_tc
        self.tc_dense = tc and _env_flag("B200DP_TC_DENSE")
        self.tc_conv = tc and _env_flag("B200DP_TC_CONV")
        self.comm_blocks = comm_blocks
        self.global_step = 0
        self.launches_per_step = 0
        self._nvtx = _env_flag("B200DP_NVTX", "0")      # NVTX ranges around the phases of a step (for nsys / ncu --nvtx)
        self.sample_comm_timing = False                  # True: the next step runs eagerly with CUDA events around its collectives
        self._comm_events: List[Tuple[torch.cuda.Event, torch.cuda.Event]] = []
        self._timing_now = False

        L = self.spec.layers
        self.C1 = L[0].args["cout"]
        self.K1 = L[0].args["k"]
        if self.keras:
            # conv(VALID) conv(VALID) pool dropout flatten dense dropout dense
            self.C2, self.K2, self.PAD2 = L[1].args["cout"], L[1].args["k"], 0
            self.H1 = 28 - self.K1 + 1                     # conv1 output (no pool): 26
            self.H2 = self.H1 - self.K2 + 1                # conv2 output: 24
            self.feat_drop = L[3].args["rate"]             # Dropout(0.25) on the pooled features, Philox layer 0
            self.head_layer = 1                            # Dropout(0.5) after dense: Philox layer 1
        else:
            # conv(SAME) pool conv(SAME) pool flatten dense dropout dense
            self.C2, self.K2, self.PAD2 = L[2].args["cout"], L[2].args["k"], L[2].args["k"] // 2
            self.H1 = 14                                   # conv1 output after its pool
            self.H2 = 14
            self.feat_drop = 0.0
            self.head_layer = 0
        self.HP = self.H2 // 2                             # pooled conv2 output: 7 / 12
        self.D = L[5].args["units"]
        self.F = self.HP * self.HP * self.C2
        self.NC = self.spec.num_classes
        self.drop_rate = L[6].args["rate"]
        if self.tc_conv and (self.C1 % 32 or self.C2 % 64):
            self.tc_conv = False
        if self.tc_dense and (self.F % 8 or self.D % 64):
            self.tc_dense = False

        # ---- symmetric state -------------------------------------------------------------
        w = _w.world() if _w.is_initialized() else None
        if arena is not None:
            self.arena = arena
        elif w is not None and w.arena is not None:
            self.arena = w.arena
        else:
            self.arena = self._solo_arena()
        self.n = self.arena.n
        self.rank = self.arena.rank
        T = self.spec.total
        self.T = T
        a = self.arena
        self.seg_ctrl = a.alloc(CTRL_BYTES)
        self.seg_state = a.alloc(T * 4 * 3)
        d0 = _as_dist_adam(optimizer)
        want = os.environ.get("B200DP_ALLREDUCE_ALGO") or (d0.algo if d0 else "auto")
        self.nvls = bool(self.n > 1 and want in ("nvls", "auto") and a.nvls_supported()
                         and (want == "nvls" or autoselect.nvls_min_bytes(self.n) is not None))
        galloc = a.alloc_multicast if self.nvls else a.alloc
        self.seg_grad = galloc(T * 4 * 3 + (1 << 16))        # parity 0, parity 1, adasum pong, adasum scratch
        self.seg_w16 = galloc(T * 2)
        self.master = a.view(self.seg_state, 0, T, torch.float32)
        self.m = a.view(self.seg_state, T * 4, T, torch.float32)
        self.v = a.view(self.seg_state, 2 * T * 4, T, torch.float32)
        self.grads = [a.view(self.seg_grad, p * T * 4, T, torch.float32) for p in (0, 1)]
        self.w16 = a.view(self.seg_w16, 0, T, bf16)
        self.step_dev = a.view(self.seg_ctrl, CTRL_STEP, 1, torch.int32)
        self.hyper_dev = a.view(self.seg_ctrl, CTRL_HYPER, 5, torch.float64)
        self.pow_dev = a.view(self.seg_ctrl, CTRL_POW, 2, torch.float64)
        # the first kernel of a step multiplies the beta powers once and leaves the step's Adam coefficients in the control
        # block; the optimizer kernels read them instead of running two fp64 pow() each
        self._coef_cache = not a.emulated and _env_flag("B200DP_COEF_CACHE")
        if self._coef_cache:
            K.check(K.lib().b200_comm_set_coef_cache(a.h, self.seg_ctrl, CTRL_COEF), "set_coef_cache")
        self.eval_loss = a.view(self.seg_ctrl, CTRL_EVAL_LOSS, 1, torch.float32)
        self.eval_correct = a.view(self.seg_ctrl, CTRL_EVAL_CORRECT, 1, torch.int32)
        self.loss_ring = a.view(self.seg_ctrl, CTRL_LOSS_RING, RING, torch.float32)
        self.correct_ring = a.view(self.seg_ctrl, CTRL_CORRECT_RING, RING, torch.int32)

        # host mirror of the two rings: the last kernel of every step stores its slot here (mapped pinned memory)
        self.host_loss = torch.zeros(RING, dtype=torch.float32).pin_memory()
        self.host_correct = torch.zeros(RING, dtype=torch.int32).pin_memory()
        self._mirror = not a.emulated and _env_flag("B200DP_HOST_MIRROR")
        # B200DP_POISON=1 (race hunting): whatever the protocol says nobody reads any more is overwritten with
        # NaN — consumed gradient buffers by the fused kernels, the activations the peers gather by a fill after the
        # step's last barrier — so a late reader turns the loss into NaN instead of training on stale or zero data
        self.poison = _env_flag("B200DP_POISON", "0")
        if self.poison:
            a.set_debug_poison(True)
        self.master.copy_(init_flat(self.spec, seed, "cpu").to(self.device))
        K.cast_f32_to_bf16(self.master, self.w16)
        self._sharded_ranges: List[Tuple[int, int]] = []
        self._shards_stale = False      # True once a two-shot step ran: peers' slices of master/m/v are newer than ours

        # ---- activations (plain device memory) --------------------------------------------
        B, dev = self.cap, self.device
        e = lambda *s, dt=bf16: torch.empty(*s, device=dev, dtype=dt)
        # two input buffers, one per gradient-buffer parity: the H2D copy of step t+1 (copy stream) runs underneath
        # step t, whose graph reads the other buffer
        self._xs = [torch.zeros(B, 784, device=dev) for _ in range(2)]
        self._ys = [torch.zeros(B, device=dev, dtype=torch.int32) for _ in range(2)]
        self.copy_stream = torch.cuda.Stream(device=dev)
        self._in_ready = [torch.cuda.Event() for _ in range(2)]
        self._buf_free = [torch.cuda.Event() for _ in range(2)]
        H1, H2, HP = self.H1, self.H2, self.HP
        self.a1 = e(B, H1, H1, self.C1)
        self.am1 = None if self.keras else e(B, H1, H1, self.C1, dt=torch.uint8)
        self.c2 = None if self.tc_conv else e(B, H2, H2, self.C2)
        # a2 (flattened conv features) and dz are what the peers gather for the fused dense wgrad: they live in a symmetric
        # segment [a2 | dz | gathered a2 slices of all ranks | gathered dz of all ranks]
        self.ACT_A2, self.ACT_DZ = 0, B * self.F * 2
        self.ACT_A2G = self.ACT_DZ + B * self.D * 2
        self.ACT_DZG = self.ACT_A2G + B * self.F * 2
        self.seg_act = a.alloc(self.ACT_DZG + self.n * B * self.D * 2)
        self.a2 = a.view(self.seg_act, self.ACT_A2, B * self.F, bf16).view(B, HP, HP, self.C2)
        self.dz = a.view(self.seg_act, self.ACT_DZ, B * self.D, bf16).view(B, self.D)
        self.am2 = e(B, HP, HP, self.C2, dt=torch.uint8)
        self.z = torch.zeros(B, self.D, device=dev)
        self.h = e(B, self.D)
        self.dlogits, self.logits = e(B, self.NC, dt=torch.float32), e(B, self.NC, dt=torch.float32)
        self.dpool, self.dy2, self.da1 = e(B, HP, HP, self.C2), e(B, H2, H2, self.C2), e(B, H1, H1, self.C1)
        self._pinned = [(torch.empty(B, 784).pin_memory(), torch.empty(B, dtype=torch.int32).pin_memory(), torch.cuda.Event())
                        for _ in range(3)]
        self._pin_i = 0
        # (a middle priority for the dense-update branch was measured and lost: 112.4 -> 119.7 us at N=1, 122.8 -> 128.9 at
        # N=2 — the HBM-bound Adam sweep then runs in front of the conv2 weight gradient and slows the main chain beside it)
        self.comm_stream = torch.cuda.Stream(device=dev)
        self.side_stream = torch.cuda.Stream(device=dev)
        # The step's kernels each want most of the machine, so whichever launches first holds the SMs.  The main chain
        # (forward, dgrads, conv1 wgrad) IS the critical path: it is captured on a high-priority stream, so its blocks are
        # placed first whenever SM resources free up, and the weight-gradient / optimizer / collective branches fill in
        # behind it (kernel nodes of a graph inherit the priority of the stream they were captured on).
        self.main_stream = torch.cuda.Stream(device=dev, priority=-5) if _env_flag("B200DP_PRIO") else None
        self._graphs: Dict[Tuple[int, int], torch.cuda.CUDAGraph] = {}
        self._warm: set = set()
        self.dist: Optional[DistributedAdam] = None
        self.adam: Optional[Adam] = None
        self.set_optimizer(optimizer)
        if _env_flag("B200DP_L2_PERSIST", "0"):
            # experiment: pin the fp32 master/m/v sweep target in L2 for the streams the fused Adam kernels run on
            # (the dense bucket's kernel runs on comm_stream; a graph capture gives the main chain a stream of its own)
            for st in [s for s in (torch.cuda.current_stream(), self.comm_stream) if s.cuda_stream != 0]:
                K.check(K.lib().b200_stream_l2_persist(ctypes.c_void_p(self.master.data_ptr()), T * 4 * 3,
                                                       float(os.environ.get("B200DP_L2_PERSIST_RATIO", "1.0")),
                                                       ctypes.c_void_p(st.cuda_stream)), "l2_persist")
        torch.cuda.synchronize()

    @property
    def x(self) -> torch.Tensor:
        """Device input buffer the next step (or an eval) reads."""
        return self._xs[self.global_step & 1]

    @property
    def y(self) -> torch.Tensor:
        return self._ys[self.global_step & 1]

    def _solo_arena(self) -> SymmetricArena:
        import ctypes
        h = ctypes.c_void_p()
        K.check(K.lib().b200_comm_create(0, 1, 0, ctypes.byref(h)), "comm_create")
        return SymmetricArena(h, 0, 1, False, self.device)

    # ------------------------------------------------------------------------------------
    def pv(self, name: str, buf: torch.Tensor) -> torch.Tensor:
        p = self.spec.param(name)
        return buf[p.offset:p.offset + p.numel].view(p.shape)

    def set_optimizer(self, optimizer) -> None:
        self.dist = _as_dist_adam(optimizer)
        self.adam = self.dist.base if self.dist else (optimizer if isinstance(optimizer, Adam) else None)
        if self.adam is not None:
            a = self.adam
            self.hyper_dev.copy_(torch.tensor([a.lr, a.beta1, a.beta2, a.epsilon, 0.0 if a.eps_mode == "tf" else 1.0],
                                              dtype=torch.float64))
            self._set_step_dev()
        self._graphs.clear()
        self._plan_comm()

    def _set_step_dev(self) -> None:
        """Device step counter and the beta-power accumulators that go with it (beta^step, like TensorFlow's beta1_power /
        beta2_power variables): written whenever the step or the betas change outside the kernels."""
        self.step_dev.fill_(self.global_step)
        b1, b2 = (self.adam.beta1, self.adam.beta2) if self.adam is not None else (0.9, 0.999)
        self.pow_dev.copy_(torch.tensor([b1 ** self.global_step, b2 ** self.global_step], dtype=torch.float64))

    def set_lr(self, lr: float) -> None:
        """The kernels read the learning rate from device memory at run time, so a new value needs no graph re-capture."""
        super().set_lr(lr)
        self.hyper_dev[0:1].fill_(float(lr))

    def _plan_comm(self) -> None:
        """Static bucket plan: which ranges each comm kernel reduces, and with which algorithm."""
        buckets = plan_arena_buckets(self.spec, [["dense.bias", "out."], ["dense.kernel"], ["conv1."], ["conv2."]])
        self.reduce = bool(self.dist is not None and self.n > 1)
        adasum = self.reduce and self.dist.op is C.Adasum
        forced = self.dist.algo if self.dist else None
        self._sharded_ranges = []

        def rng(b):
            algo = autoselect.choose(b.numel * 4, self.n, forced, nvls_ok=self.nvls) if (self.reduce and not adasum) else 0
            if algo in (autoselect.TWOSHOT, autoselect.NVLS):
                self._sharded_ranges.append((b.start, b.numel))
            return (b.start, b.numel, algo, 1)
        by = {b.name: b for b in buckets}
        # Fused dense path (default): no gradient buffer and no allreduce for dense.kernel (98 % of the parameters) at all.
        # Every rank gathers the peers' dz and its own feature slice of their activations, runs the wgrad GEMM over the
        # GLOBAL batch for the rows it owns, applies Adam in the GEMM epilogue and stores the bf16 weights into every rank.
        # The small tensors are reduced by two one-shot kernels: conv2/out/dense.bias underneath the rest of backward, and
        # conv1 (832 values) as the only exposed collective of the step.  B200DP_DENSE_FUSED=0 or a forced algorithm
        # (oneshot/twoshot/nvls) brings back the gradient-buffer paths.
        dk0 = self.spec.param("dense.kernel")
        self.dense_fused = bool(self.tc_dense and self.tc_conv and not adasum and (self.reduce or self.n == 1) and self.adam is not None
                                and forced in (None, "auto") and not os.environ.get("B200DP_ALLREDUCE_ALGO")
                                and self.F % (8 * self.n) == 0 and self.D % 64 == 0 and dk0.offset % 4 == 0
                                and _env_flag("B200DP_DENSE_FUSED"))
        if self.dense_fused:
            if self.n > 1:
                self._sharded_ranges.append((dk0.offset, dk0.numel))
            self.gemm_rs = False
            self.comm_dense = []
            one = lambda b: (b.start, b.numel, 0, 1)
            self.comm_rest = [one(by["conv2."]), one(by["dense.bias+out."])]
            self.comm_tail = [one(by["conv1."])]
            if self.reduce and _env_flag("B200DP_MERGE_TAIL"):
                # across ranks ONE kernel reduces all small tensors at the end of the step: its barrier is the step's
                # closing barrier anyway, and the peers' 250 KB arrive by bulk copies in a single NVLink round trip, so a
                # second collective (with a barrier of its own, contending with the dense update) buys nothing
                self.comm_tail = [one(by["conv1."]), one(by["conv2."]), one(by["dense.bias+out."])]
                self.comm_rest = []
                if _env_flag("B200DP_PUSH_TAIL"):
                    # ... and the ranks do not fetch each other's slices after the barrier: each stores its own into the
                    # peers' inboxes in front of its flag (one one-way latency instead of two round trips)
                    bpp = (self.n * sum(r[1] for r in self.comm_tail) * 4 + 255) // 256 * 256
                    if getattr(self, "seg_inbox", None) is None or self._inbox_bpp < bpp:
                        self.seg_inbox, self._inbox_bpp = self.arena.alloc(2 * bpp), bpp
                    K.check(K.lib().b200_comm_set_inbox(self.arena.h, 2, self.seg_inbox, self._inbox_bpp), "set_inbox")
            self.adasum = False
            self.scale = 1.0 / self.n if (self.reduce and self.dist.op is C.Average) else 1.0
            self._set_mirror(2)
            return
        self.comm_tail = []
        self.comm_dense = [rng(by["dense.kernel"])]
        # dense.kernel's gradient is produced by plain stores (GEMM epilogue), never accumulated: nothing to clear
        self.comm_dense = [(r[0], r[1], r[2], 0) for r in self.comm_dense]
        # GEMM -> reduce-scatter fusion (B200DP_GEMM_RS=0 disables): the dense wgrad GEMM adds its tiles straight into the
        # owner ranks' buffers over NVLink (bulk reduce-stores), so the fused kernel finds its slice already summed
        # (range algo 3) and skips the peer reads.  +7 % step throughput at N=8 over the peer-read two-shot, see DESIGN.md.
        dk = self.spec.param("dense.kernel")
        self.gemm_rs = bool(self.reduce and not adasum and self.tc_dense and self.comm_dense[0][2] == autoselect.TWOSHOT
                            and self.F % self.n == 0 and self.comm_dense[0][0] == dk.offset and self.comm_dense[0][1] == dk.numel
                            and _env_flag("B200DP_GEMM_RS") and not _env_flag("B200DP_DETERMINISTIC", "0"))
        if self.gemm_rs:
            self.comm_dense = [(dk.offset, dk.numel, 3, 1)]
        self.comm_rest = [rng(by["dense.bias+out."]), rng(by["conv1."]), rng(by["conv2."])]
        self.adasum = adasum
        self.scale = 1.0 / self.n if (self.reduce and self.dist.op is C.Average) else 1.0
        self._set_mirror(0 if adasum else 1)

    def _set_mirror(self, channel: int) -> None:
        """Name the channel whose fused kernel ends the step: it also stores the step's loss/correct slot to the host ring."""
        if not self._mirror:
            return
        K.check(K.lib().b200_comm_set_step_mirror(self.arena.h, channel, self.seg_ctrl, CTRL_LOSS_RING, CTRL_CORRECT_RING, RING,
                                                  ctypes.c_void_p(self.host_loss.data_ptr()),
                                                  ctypes.c_void_p(self.host_correct.data_ptr())), "set_step_mirror")

    # ------------------------------------------------------------------------------------ kernels
    def _forward(self, B: int, training: bool) -> None:
        w = self.w16
        x, a1, a2, am2 = self.x[:B], self.a1[:B], self.a2[:B], self.am2[:B]
        ring = dict(loss_ring=self.loss_ring if training else None, correct_ring=self.correct_ring if training else None)
        if self.keras:
            K.conv1_valid_fwd(x, self.pv("conv1.kernel", w), self.pv("conv1.bias", w), a1,
                              self.step_dev if training else None, bump_step=(2 if self._coef_cache else 1) if training else 0, **ring)
        else:
            K.conv1_pool_fwd(x, self.pv("conv1.kernel", w), self.pv("conv1.bias", w), a1, self.am1[:B],
                             self.step_dev if training else None, bump_step=(2 if self._coef_cache else 1) if training else 0, **ring)
        k2, b2 = self.pv("conv2.kernel", w), self.pv("conv2.bias", w)
        if self.tc_conv:
            K.conv_tc_fwd(a1, k2, b2, a2, am2, relu=True, pool=True, pad=self.PAD2)
        else:
            K.conv_direct_fwd(a1, k2, b2, self.c2[:B], self.PAD2, relu=True)
            K.maxpool_fwd(self.c2[:B], a2, am2)
        if training and self.dropout and self.feat_drop > 0:
            # Keras Dropout(0.25) between the pool and Flatten, in place; its backward rides in the unpool kernel
            K.dropout_bf16(a2, a2, self.feat_drop, self.seed, self.step_dev, 0)
        a2f = a2.view(B, self.F)
        wd = self.pv("dense.kernel", w)
        if self.tc_dense:
            # split-K: fill the machine, but every split adds one more fp32 red.add per output address — past ~32 adds the
            # L2 atomics on the small [B,D] accumulator cost more than the extra CTAs buy (keras128: 12.0 us at 144 splits,
            # 5.2 us at 24; profiles/kernel_bench_keras128.json)
            split = max(1, min(self.F // 64, (148 * 2) // max(1, (self.D // 64) * ((B + 127) // 128)), 32))
            K.gemm_tc(a2f, wd, self.z[:B], B, self.D, self.F, b_mn_major=True, out_mode=2, split_k=split, block_n=64)
        else:
            K.gemm_simple(a2f, wd, self.z[:B], B, self.D, self.F)

    def _head(self, B: int, training: bool) -> None:
        w = self.w16
        if training:
            K.head_fwd_bwd(self.z[:B], self.pv("dense.bias", w), self.pv("out.kernel", w), self.pv("out.bias", w), self.y[:B],
                           self.h[:B], self.dz[:B], self.dlogits[:B], None, self.loss_ring, self.correct_ring,
                           drop_rate=self.drop_rate if self.dropout else 0.0, training=True, seed=self.seed,
                           step_ptr=self.step_dev, layer_id=self.head_layer, zero_z=True, ring=RING)
        else:
            K.head_fwd_bwd(self.z[:B], self.pv("dense.bias", w), self.pv("out.kernel", w), self.pv("out.bias", w), self.y[:B],
                           self.h[:B], None, None, self.logits[:B], self.eval_loss, self.eval_correct,
                           drop_rate=0.0, training=False, seed=self.seed, step_ptr=None, zero_z=True, ring=0)

    def _backward_and_update(self, B: int, par: int) -> None:
        """Backward as a small DAG over three streams (captured as graph branches):

            main : dense dgrad -> unpool (+conv2 bias grad) -> conv2 dgrad -> conv1 wgrad -> comm(rest)
            side1: head wgrad, dense wgrad                    -> comm(dense bucket), underneath the conv backward
            side2: conv2 wgrad (needs dy2 only)
        """
        w, g = self.w16, self.grads[par]
        main = torch.cuda.current_stream()
        s1, s2 = self.comm_stream, self.side_stream
        a2f = self.a2[:B].view(B, self.F)
        gd = self.pv("dense.kernel", g)
        wd = self.pv("dense.kernel", w)
        dpool = self.dpool[:B].view(B, self.F)
        ev_head = torch.cuda.Event()
        ev_head.record(main)
        # ---- side 1: gradients that only need the head's outputs
        s1.wait_event(ev_head)
        with torch.cuda.stream(s1):
            K.head_wgrad(self.h[:B], self.dz[:B], self.dlogits[:B], self.pv("out.kernel", g), self.pv("out.bias", g),
                         self.pv("dense.bias", g))
            if self.gemm_rs:
                K.gemm_tc_reduce_scatter(self.arena.h, a2f, self.dz[:B], self.seg_grad,
                                         par * self.T * 4 + self.spec.param("dense.kernel").offset * 4, self.n,
                                         self.F, self.D, B, a_mn_major=True, b_mn_major=True, block_n=128)
            elif self.tc_dense:
                K.gemm_tc(a2f, self.dz[:B], gd, self.F, self.D, B, a_mn_major=True, b_mn_major=True, out_mode=0, block_n=128)
            else:
                K.gemm_simple(a2f, self.dz[:B], gd, self.F, self.D, B, transA=True)
        # ---- main: the dgrad chain
        if self.tc_dense:
            K.gemm_tc(self.dz[:B], wd, dpool, B, self.F, self.D, out_mode=1, block_n=32)
        else:
            K.gemm_simple(self.dz[:B], wd, dpool, B, self.F, self.D, transB=True)
        ev_dgrad = torch.cuda.Event()
        ev_dgrad.record(main)
        if not self.adasum:
            # dense bucket: gradient complete (side 1) and W no longer read (dgrad done)
            s1.wait_event(ev_dgrad)
            with torch.cuda.stream(s1):
                self._timed(self._comm, 0, self.comm_dense, par)
        ev_s1 = torch.cuda.Event()
        ev_s1.record(s1)
        gscale = 1.0 / (1.0 - self.feat_drop) if (self.dropout and self.feat_drop > 0) else 1.0
        K.unpool_relu_bwd(self.dpool[:B], self.a2[:B], self.am2[:B], self.dy2[:B], self.pv("conv2.bias", g), gscale=gscale)
        ev_dy = torch.cuda.Event()
        ev_dy.record(main)
        k2 = self.pv("conv2.kernel", w)
        s2.wait_event(ev_dy)
        with torch.cuda.stream(s2):
            if self.tc_conv:
                K.conv_tc_wgrad(self.a1[:B], self.dy2[:B], self.pv("conv2.kernel", g), None, pad=self.PAD2)
            else:
                K.conv_direct_wgrad(self.a1[:B], self.dy2[:B], self.pv("conv2.kernel", g), None, self.PAD2)
            ev_s2 = torch.cuda.Event()
            ev_s2.record(s2)
        if self.tc_conv:
            K.conv_tc_dgrad(self.dy2[:B], k2, self.da1[:B], pad=self.PAD2)
        else:
            K.conv_direct_dgrad(self.dy2[:B], k2, self.da1[:B], self.PAD2)
        if self.keras:
            K.conv1_valid_wgrad(self.x[:B], self.da1[:B], self.a1[:B], self.pv("conv1.kernel", g), self.pv("conv1.bias", g))
        else:
            K.conv1_wgrad(self.x[:B], self.da1[:B], self.a1[:B], self.am1[:B], self.pv("conv1.kernel", g), self.pv("conv1.bias", g))
        main.wait_event(ev_s2)
        main.wait_event(ev_s1)
        if self.adasum:
            self._adasum_update(par)
        else:
            self._timed(self._comm, 1, self.comm_rest, par)

    def _timed(self, fn, *a, **k) -> None:
        """Run a collective kernel launch; when this step samples timings, bracket it with events on its stream."""
        if not self._timing_now:
            fn(*a, **k)
            return
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn(*a, **k)
        e1.record()
        self._comm_events.append((e0, e1))

    def last_comm_ms(self) -> Optional[float]:
        """Device time of the collective kernels of the last sampled step (gather/allreduce+Adam kernels summed; the
        barrier waits inside them included), or None when no step was sampled yet."""
        if not self._comm_events:
            return None
        self._comm_events[-1][1].synchronize()
        return float(sum(a.elapsed_time(b) for a, b in self._comm_events))

    def _dense_wgrad_adam(self, B: int) -> None:
        """dense.kernel: gather (n > 1) -> wgrad GEMM over the global batch -> Adam in the epilogue -> bf16 weights to all."""
        n, T4 = self.n, self.T * 4
        dk = self.spec.param("dense.kernel")
        Fs = self.F // n
        if n > 1:
            self._timed(K.allgather_acts, self.arena.h, 3, 148, self.seg_act, self.ACT_A2, self.seg_act, self.ACT_DZ, self.seg_act,
                        self.ACT_A2G, self.ACT_DZG, B, self.F, self.D)
            A = self.arena.view(self.seg_act, self.ACT_A2G, n * B * Fs, bf16).view(n * B, Fs)
            Bm = self.arena.view(self.seg_act, self.ACT_DZG, n * B * self.D, bf16).view(n * B, self.D)
        else:
            A, Bm = self.a2[:B].view(B, self.F), self.dz[:B]
        tiles128 = -(-Fs // 128) * (self.D // 128) if self.D % 128 == 0 else 0
        bn = 128 if tiles128 >= 96 else 64
        K.gemm_tc_wgrad_adam(self.arena.h, A, Bm, Fs, self.D, n * B, Fs, self.D, state_seg=self.seg_state, master_off=0,
                             m_off=T4, v_off=2 * T4, w16_seg=self.seg_w16, w16_off=0, ctrl_seg=self.seg_ctrl,
                             step_off=CTRL_STEP, hyper_off=CTRL_HYPER, param_start=dk.offset, row0=self.rank * Fs,
                             gscale=self.scale, block_n=bn)

    def _backward_fused(self, B: int, par: int) -> None:
        """Backward with the fused dense path (see _plan_comm), three streams captured as graph branches:

            main : dense dgrad -> unpool (+conv2 bias grad) -> conv2 dgrad -> conv1 wgrad -> [join] -> comm(conv1): the only
                   collective whose latency the step sees
            side1: head wgrad ; after dgrad: gather dz/a2 over NVLink -> dense wgrad GEMM + Adam + weight push
            side2: conv2 wgrad ; comm(conv2, out, dense.bias) underneath conv2 dgrad / conv1 wgrad
        """
        w, g = self.w16, self.grads[par]
        main = torch.cuda.current_stream()
        s1, s2 = self.comm_stream, self.side_stream
        wd = self.pv("dense.kernel", w)
        dpool = self.dpool[:B].view(B, self.F)
        ev_head = torch.cuda.Event()
        ev_head.record(main)
        s1.wait_event(ev_head)
        with torch.cuda.stream(s1):
            K.head_wgrad(self.h[:B], self.dz[:B], self.dlogits[:B], self.pv("out.kernel", g), self.pv("out.bias", g),
                         self.pv("dense.bias", g))
            ev_hw = torch.cuda.Event()
            ev_hw.record(s1)
        K.gemm_tc(self.dz[:B], wd, dpool, B, self.F, self.D, out_mode=1, block_n=32)
        ev_dgrad = torch.cuda.Event()
        ev_dgrad.record(main)
        s1.wait_event(ev_dgrad)          # W is no longer read locally; arriving at the gather's barrier tells the peers so
        with torch.cuda.stream(s1):
            self._dense_wgrad_adam(B)
            ev_s1 = torch.cuda.Event()
            ev_s1.record(s1)
        gscale = 1.0 / (1.0 - self.feat_drop) if (self.dropout and self.feat_drop > 0) else 1.0
        K.unpool_relu_bwd(self.dpool[:B], self.a2[:B], self.am2[:B], self.dy2[:B], self.pv("conv2.bias", g), gscale=gscale)
        ev_dy = torch.cuda.Event()
        ev_dy.record(main)
        k2 = self.pv("conv2.kernel", w)
        # Alone (n == 1) the conv2 weight gradient runs behind conv1's on the main chain: beside the dgrad it only stretched
        # both (112.4 -> 110.5 us/step).  Across ranks it keeps its own branch, so the dense update that waits for SMs is
        # not pushed further back (serial at N=2: 120.0 -> 131.5 us).
        serial = _env_flag("B200DP_SERIAL_WGRAD", "1" if self.n == 1 else "0")
        s2.wait_event(ev_dy)
        if not serial:
            with torch.cuda.stream(s2):
                K.conv_tc_wgrad(self.a1[:B], self.dy2[:B], self.pv("conv2.kernel", g), None, pad=self.PAD2)
        K.conv_tc_dgrad(self.dy2[:B], k2, self.da1[:B], pad=self.PAD2)
        if serial:
            # alone: dgrad -> conv2 wgrad -> conv1 wgrad on one chain; the small-tensor update (which needs the conv2
            # gradient and rewrites the conv2 weights) then runs underneath conv1's weight gradient
            K.conv_tc_wgrad(self.a1[:B], self.dy2[:B], self.pv("conv2.kernel", g), None, pad=self.PAD2)
        ev_dg2 = torch.cuda.Event()
        ev_dg2.record(main)
        with torch.cuda.stream(s2):
            s2.wait_event(ev_hw)
            if self.comm_rest:
                s2.wait_event(ev_dg2)        # the update rewrites the conv2 weights the dgrad above is reading
                self._timed(self._comm, 1, self.comm_rest, par, nblocks=64, nthreads=256)
            ev_s2 = torch.cuda.Event()
            ev_s2.record(s2)
        if self.keras:
            K.conv1_valid_wgrad(self.x[:B], self.da1[:B], self.a1[:B], self.pv("conv1.kernel", g), self.pv("conv1.bias", g))
        else:
            K.conv1_wgrad(self.x[:B], self.da1[:B], self.a1[:B], self.am1[:B], self.pv("conv1.kernel", g), self.pv("conv1.bias", g))
        main.wait_event(ev_s2)
        main.wait_event(ev_s1)
        # every peer reaches this kernel's barrier after its own weight pushes: passing it means the new dense weights of
        # ALL owners have landed here, and that nobody still reads this rank's a2/dz
        self._timed(self._comm, 2, self.comm_tail, par, nblocks=1 if len(self.comm_tail) == 1 else 64, nthreads=256)
        if self.poison:
            self.a2.fill_(float("nan"))      # every peer has passed the barrier above: nobody may gather these any more
            self.dz.fill_(float("nan"))

    def _comm(self, channel: int, ranges, par: int, *, reduce: Optional[bool] = None, grad_off: Optional[int] = None,
              scale: Optional[float] = None, nblocks: Optional[int] = None, nthreads: int = 512) -> None:
        T4 = self.T * 4
        red = self.reduce if reduce is None else reduce
        sc = self.scale if scale is None else scale
        cur = par * T4 if grad_off is None else grad_off
        p2p = [r for r in ranges if r[2] != autoselect.NVLS or not red]
        for r in ranges:
            if r[2] == autoselect.NVLS and red:
                # in-switch reduction of the range (multimem.ld_reduce), Adam on the owned slice, multicast bf16 store
                K.check(K.lib().b200_fused_allreduce_adam_nvls(
                    self.arena.h, channel + 8, self.comm_blocks if r[1] >= (1 << 20) else 16, 512, self.seg_grad, cur,
                    (1 - par) * T4, self.seg_state, 0, T4, 2 * T4, self.seg_w16, 0, self.seg_ctrl, CTRL_STEP, CTRL_HYPER,
                    float(sc), r[0], r[1], r[3], ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)), "fused_nvls")
        if not p2p:
            return
        nb = nblocks or (self.comm_blocks if sum(r[1] for r in p2p) >= (1 << 20) else 32)
        self.arena.fused_allreduce_adam(
            channel=channel, nblocks=nb, nthreads=nthreads, reduce=red,
            grad_seg=self.seg_grad, grad_cur_off=cur, grad_other_off=(1 - par) * T4,
            state_seg=self.seg_state, master_off=0, m_off=T4, v_off=2 * T4, w16_seg=self.seg_w16, w16_off=0,
            ctrl_seg=self.seg_ctrl, step_off=CTRL_STEP, hyper_off=CTRL_HYPER, scale=sc, ranges=p2p)

    def _adasum_update(self, par: int) -> None:
        T4 = self.T * 4
        segs = self.spec.segments()
        res = self.arena.adasum_levels(self.seg_grad, par * T4, 2 * T4, 3 * T4, [s for s, _ in segs], [n for _, n in segs],
                                       nblocks=16)
        # Every rank has passed this step's first level barrier, hence finished reading the previous
        # step's buffers: the other-parity buffer can be cleared for the next accumulation now.
        self.grads[1 - par].zero_()
        ranges = [(r[0], r[1], 0, 0) for r in self.comm_dense + self.comm_rest]
        self._comm(0, ranges, par, reduce=False, grad_off=res, scale=1.0)

    def _train_graphable(self, B: int, par: int) -> None:
        c0 = K.LAUNCHES[0]
        nvtx = self._nvtx
        if nvtx:
            torch.cuda.nvtx.range_push("b200dp.forward")
        if self.poison:
            self.grads[par].zero_()          # the fused kernels left NaN here instead of zeros
        K.lib().b200_set_stable_weights(1)   # inside a step no kernel reads weights its stream predecessor wrote
        try:
            self._train_body(B, par)
        finally:
            K.lib().b200_set_stable_weights(0)
        self.launches_per_step = K.LAUNCHES[0] - c0

    def _train_body(self, B: int, par: int) -> None:
        nvtx = self._nvtx
        self._forward(B, True)
        self._head(B, True)
        if nvtx:
            torch.cuda.nvtx.range_pop()
            torch.cuda.nvtx.range_push("b200dp.backward+allreduce+adam")
        if self.dense_fused:
            self._backward_fused(B, par)
        else:
            self._backward_and_update(B, par)
        if nvtx:
            torch.cuda.nvtx.range_pop()

    # ------------------------------------------------------------------------------------ public
    def _stage_inputs(self, x, y, B: int, *, overlap: bool = False) -> None:
        """Copy a batch into the current input buffer.  ``overlap=True`` (training) issues the copy on the copy
        stream so it runs underneath the previous step; the main stream then waits on the copy's event."""
        par = self.global_step & 1
        dx, dy = self._xs[par], self._ys[par]
        main = torch.cuda.current_stream()
        st = self.copy_stream if overlap else main
        if overlap:
            st.wait_event(self._buf_free[par])          # the step that last read this buffer has finished
        with torch.cuda.stream(st):
            if isinstance(x, torch.Tensor) and x.is_cuda:
                dx[:B].copy_(x.reshape(B, 784), non_blocking=True)
                dy[:B].copy_(y.reshape(B).to(torch.int32), non_blocking=True)
            elif isinstance(x, torch.Tensor) and isinstance(y, torch.Tensor) and x.is_pinned() and y.is_pinned() \
                    and x.dtype == torch.float32 and y.dtype == torch.int32:
                # caller-owned pinned batches go straight over PCIe (caller keeps them alive until the step ran)
                dx[:B].copy_(x.reshape(B, 784), non_blocking=True)
                dy[:B].copy_(y.reshape(B), non_blocking=True)
            else:
                px, py, ev = self._pinned[self._pin_i]
                self._pin_i = (self._pin_i + 1) % len(self._pinned)
                ev.synchronize()
                xs = x if isinstance(x, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(x))
                ys = y if isinstance(y, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(y))
                px[:B].copy_(xs.reshape(B, 784))
                py[:B].copy_(ys.reshape(B))
                dx[:B].copy_(px[:B], non_blocking=True)
                dy[:B].copy_(py[:B], non_blocking=True)
                ev.record(st)
            if overlap:
                self._in_ready[par].record(st)
        if overlap:
            main.wait_event(self._in_ready[par])

    def stage_device_batch(self, x: torch.Tensor, y: torch.Tensor) -> None:
        """Device-resident batch ([B,784] fp32, [B] int32) into the next step's input buffer, copied on the copy stream
        underneath the step that is still running (the two input buffers alternate with the gradient-buffer parity)."""
        par = self.global_step & 1
        cs = self.copy_stream
        cs.wait_event(self._buf_free[par])
        with torch.cuda.stream(cs):
            self._xs[par][:x.shape[0]].copy_(x, non_blocking=True)
            self._ys[par][:x.shape[0]].copy_(y, non_blocking=True)
            self._in_ready[par].record(cs)
        torch.cuda.current_stream().wait_event(self._in_ready[par])

    def train_step(self, x, y) -> StepResult:
        if self.adam is None:
            raise RuntimeError("no optimizer set; pass optimizer=Adam(...) or compile() the model")
        B = int(x.shape[0])
        if B > self.cap:
            raise ValueError(f"batch {B} exceeds the engine's batch_size={self.cap}")
        self._stage_inputs(x, y, B, overlap=True)
        par = self.global_step & 1
        key = (B, par)
        if self.sample_comm_timing:
            # one eagerly launched step with events around its collectives (the log line's allreduce_ms)
            self.sample_comm_timing = False
            self._comm_events = []
            self._timing_now = True
            try:
                self._train_graphable(B, par)
            finally:
                self._timing_now = False
            self._warm.add(key)
        elif self.use_graph and key in self._warm:
            g = self._graphs.get(key)
            if g is None:
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=self.main_stream):
                    self._train_graphable(B, par)
                self._graphs[key] = g
                # capture does not execute: the step still has to run
            g.replay()
        else:
            self._train_graphable(B, par)
            self._warm.add(key)
        self._buf_free[par].record(torch.cuda.current_stream())
        self.global_step += 1
        self._shards_stale = bool(self._sharded_ranges)
        slot = self.global_step % RING
        if self._mirror:
            ev = self._buf_free[par]
            return StepResult(_HostScalar(self.host_loss, slot, ev), _HostScalar(self.host_correct, slot, ev), B)
        return StepResult(_LazyScalar(self._ring_read(slot, True)), _LazyScalar(self._ring_read(slot, False)), B)

    def _ring_read(self, slot: int, loss: bool) -> torch.Tensor:
        # a view: valid until the ring wraps (RING steps), which is far longer than any logging cadence
        return (self.loss_ring if loss else self.correct_ring)[slot]

    def train_step_device(self, B: Optional[int] = None) -> None:
        """Run a step on whatever is already in ``self.x`` / ``self.y`` (bench: kernel-only timing)."""
        B = B or self.cap
        par = self.global_step & 1
        key = (B, par)
        if self.use_graph and key in self._warm:
            g = self._graphs.get(key)
            if g is None:
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=self.main_stream):
                    self._train_graphable(B, par)
                self._graphs[key] = g
            g.replay()
        else:
            self._train_graphable(B, par)
            self._warm.add(key)
        self._buf_free[par].record(torch.cuda.current_stream())
        self.global_step += 1
        self._shards_stale = bool(self._sharded_ranges)

    @torch.no_grad()
    def eval_batch(self, x, y) -> StepResult:
        n = int(x.shape[0])
        self.eval_loss.zero_()
        self.eval_correct.zero_()
        for i in range(0, n, self.cap):
            B = min(self.cap, n - i)
            self._stage_inputs(x[i:i + B], y[i:i + B], B)
            self._forward(B, False)
            self._head(B, False)
        self._buf_free[self.global_step & 1].record(torch.cuda.current_stream())
        return StepResult(_LazyScalar(self.eval_loss.clone(), 1.0 / n), _LazyScalar(self.eval_correct.clone()), n)

    @torch.no_grad()
    def predict(self, x) -> torch.Tensor:
        n = int(x.shape[0])
        outs = []
        dummy = np.zeros(self.cap, np.int32)
        for i in range(0, n, self.cap):
            B = min(self.cap, n - i)
            self._stage_inputs(x[i:i + B], dummy[:B], B)
            self._forward(B, False)
            self._head(B, False)
            outs.append(self.logits[:B].clone())
        self._buf_free[self.global_step & 1].record(torch.cuda.current_stream())
        return torch.cat(outs)

    # ------------------------------------------------------------------------------------ state
    def _gather_sharded(self, sync: bool = True) -> None:
        """One-sided: pull every owner's slice of the sharded ranges into the local fp32 state (copies in stream order).
        Safe between steps: this rank's step has passed its last barrier, so every owner's update is complete, and no
        peer can start its next update before this rank — whose stream runs these copies first — joins the next barrier."""
        if self.n == 1 or not self._sharded_ranges or not self._shards_stale:
            return
        if sync:
            torch.cuda.synchronize()
        T = self.T
        for (st, ln) in self._sharded_ranges:
            chunk = (-(-ln // self.n) + 3) // 4 * 4
            for r in range(self.n):
                if r == self.rank:
                    continue
                lo, hi = min(r * chunk, ln), min((r + 1) * chunk, ln)
                if hi <= lo:
                    continue
                for base in (0, T, 2 * T):
                    self.arena.pull_from_peer(self.seg_state, (base + st + lo) * 4, (hi - lo) * 4, r)
        if sync:
            torch.cuda.synchronize()
        self._shards_stale = False

    def finalize(self) -> None:
        """Collective, at the end of training: every rank brings its copy of the owner-sharded fp32 state up to date
        (all-gather of the slices over NVLink), so whatever rank 0 saves or exports afterwards touches local memory only
        and no rank reads a peer that may already have shut down."""
        if self.n > 1 and self._sharded_ranges and self._shards_stale and _w.is_initialized():
            torch.cuda.synchronize()
            for (st, ln) in self._sharded_ranges:
                for base in (0, self.T, 2 * self.T):
                    self.arena.allgather_slices_(self.seg_state, (base + st) * 4, ln, nblocks=32)
            torch.cuda.synchronize()
            self._shards_stale = False
        if self.n > 1 and _w.is_initialized():
            _w.barrier()

    def state(self) -> Dict[str, Any]:
        self._gather_sharded()
        meta = self.meta()
        meta.update(global_step=self.global_step, seed=self.seed, engine="b200")
        if self.adam is not None:
            a = self.adam
            meta["optimizer"] = dict(lr=a.lr, beta1=a.beta1, beta2=a.beta2, epsilon=a.epsilon, eps_mode=a.eps_mode)
        return {"tensors": {"params": self.master.cpu().clone(), "adam_m": self.m.cpu().clone(), "adam_v": self.v.cpu().clone()},
                "meta": meta}

    def state_async(self) -> "_PendingState":
        """Checkpoint without stalling the job: the fp32 state is snapshotted device-to-device in stream order (~12 us for
        the MNIST CNN), training continues at once, and the snapshot drains to pinned host memory on the copy stream.
        ``result()`` of the returned handle (called by the checkpoint writer thread) waits for that copy only.  Peers
        therefore never sit in a device-side barrier while rank 0 writes a file."""
        self._gather_sharded(sync=False)
        T = self.T
        if getattr(self, "_snap", None) is None:
            self._snap = torch.empty(3 * T, device=self.device, dtype=torch.float32)
            self._snap_host = [torch.empty(3 * T, dtype=torch.float32).pin_memory() for _ in range(2)]
            self._snap_ev = [torch.cuda.Event() for _ in range(2)]
            self._snap_i = 0
            self.ckpt_stream = torch.cuda.Stream(device=self.device)      # not the input-copy stream: a 40 MB drain would delay batches
        i = self._snap_i
        self._snap_i ^= 1
        self._snap_ev[i].synchronize()                       # the writer that used this host buffer two saves ago is long done
        main = torch.cuda.current_stream()
        main.wait_stream(self.ckpt_stream)                   # an earlier drain must not still be reading the snapshot
        self._snap.copy_(self.arena.view(self.seg_state, 0, 3 * T, torch.float32), non_blocking=True)
        ready = torch.cuda.Event()
        ready.record(main)
        self.ckpt_stream.wait_event(ready)
        with torch.cuda.stream(self.ckpt_stream):
            self._snap_host[i].copy_(self._snap, non_blocking=True)
            self._snap_ev[i].record(self.ckpt_stream)
        meta = self.meta()
        meta.update(global_step=self.global_step, seed=self.seed, engine="b200")
        if self.adam is not None:
            a = self.adam
            meta["optimizer"] = dict(lr=a.lr, beta1=a.beta1, beta2=a.beta2, epsilon=a.epsilon, eps_mode=a.eps_mode)
        return _PendingState(self._snap_host[i], self._snap_ev[i], T, meta)

    def load_state(self, tensors: Dict[str, torch.Tensor], meta: Dict[str, Any]) -> None:
        if meta.get("arch", self.spec.arch) != self.spec.arch or tensors["params"].numel() != self.T:
            raise ValueError("checkpoint does not match this model (arch/width differ)")
        self.master.copy_(tensors["params"].to(self.device))
        if "adam_m" in tensors:
            self.m.copy_(tensors["adam_m"].to(self.device))
            self.v.copy_(tensors["adam_v"].to(self.device))
        K.cast_f32_to_bf16(self.master, self.w16)
        self.global_step = int(meta.get("global_step", meta.get("step", 0)))
        self._set_step_dev()
        if "seed" in meta:
            self.seed = int(meta["seed"])
            self._graphs.clear()
        for gb in self.grads:
            gb.zero_()
        self._shards_stale = False
        torch.cuda.synchronize()

    def broadcast_state(self, root_rank: int = 0) -> None:
        if not _w.is_initialized() or self.n == 1:
            return
        if self.rank == root_rank:
            self._gather_sharded()
        else:
            torch.cuda.synchronize()
        if (self.T * 3) % 4 == 0 and not self.arena.emulated:
            # master | m | v are one contiguous range of a segment every peer has mapped: the ranks pull it straight out of
            # the root's copy (scatter + all-gather of slices above 1 MB) in ONE kernel, with no staging buffer in between
            from ..parallel.arena import CH_BCAST
            K.check(K.lib().b200_broadcast(self.arena.h, CH_BCAST, 64, 512, root_rank, self.seg_state, 0, self.seg_state, 0,
                                           self.T * 3, ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)), "broadcast_state")
        else:
            for t in (self.master, self.m, self.v):
                C.broadcast_(t, root_rank)
        K.cast_f32_to_bf16(self.master, self.w16)
        step, seed = C.broadcast_object((self.global_step, self.seed), root_rank)
        if seed != self.seed or (step & 1) != (self.global_step & 1):
            self._graphs.clear()
        self.global_step, self.seed = step, seed
        self._set_step_dev()
        self._shards_stale = False
        torch.cuda.synchronize()
        _w.barrier()
