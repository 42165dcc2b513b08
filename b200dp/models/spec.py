"""Model specifications and the flat parameter layout shared by every engine.

The default ``mnist_cnn`` is the network *as coded* in both reference scripts
(``horovod/tensorflow_mnist.py:38-73``, ``horovod/tensorflow_mnist_gpu.py:40-88``):
conv5x5x32 SAME+ReLU -> maxpool2 -> conv5x5x64 SAME+ReLU -> maxpool2 ->
dense1024+ReLU -> dropout0.5 -> dense10  (3,274,634 parameters).
``keras128`` is the shape BASELINE.json's prose names
(conv3x3x32 -> conv3x3x64 -> maxpool -> dropout -> dense128 -> dense10; 1,199,882).

Parameters use TensorFlow's layouts (conv kernels HWIO, dense kernels [in, out])
so a checkpoint lines up tensor-for-tensor with the reference's variables, and
live in ONE flat fp32 buffer so the gradient allreduce and the fused Adam are
elementwise sweeps over a bucket with no per-tensor launches.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Dict, List, Tuple

import torch

ALIGN_ELEMS = 64          # every tensor starts on a 256-byte boundary in the fp32 arena


@dataclass(frozen=True)
class ParamSpec:
    name: str
    shape: Tuple[int, ...]
    offset: int               # in elements, into the flat buffer
    init: str                 # "glorot_uniform" | "zeros"
    fan_in: int = 0
    fan_out: int = 0

    @property
    def numel(self) -> int:
        return int(math.prod(self.shape))


@dataclass(frozen=True)
class LayerSpec:
    kind: str                 # conv | pool | flatten | dense | dropout
    args: dict = field(default_factory=dict)


@dataclass(frozen=True)
class ModelSpec:
    arch: str
    width: float
    layers: Tuple[LayerSpec, ...]
    params: Tuple[ParamSpec, ...]
    total: int                # padded flat length (elements)
    input_hw: Tuple[int, int] = (28, 28)
    num_classes: int = 10

    @property
    def num_params(self) -> int:
        return sum(p.numel for p in self.params)

    def param(self, name: str) -> ParamSpec:
        for p in self.params:
            if p.name == name:
                return p
        raise KeyError(name)

    def segments(self) -> List[Tuple[int, int]]:
        """(offset, numel) per tensor, in layout order."""
        return [(p.offset, p.numel) for p in self.params]

    def views(self, flat: torch.Tensor) -> Dict[str, torch.Tensor]:
        return {p.name: flat[p.offset:p.offset + p.numel].view(p.shape) for p in self.params}

    def fwd_flops_per_image(self) -> int:
        h, w = self.input_hw
        c = 1
        flops = 0
        feat = None
        for L in self.layers:
            a = L.args
            if L.kind == "conv":
                if a["padding"] == "VALID":
                    h, w = h - a["k"] + 1, w - a["k"] + 1
                flops += 2 * h * w * a["k"] * a["k"] * c * a["cout"]
                c = a["cout"]
            elif L.kind == "pool":
                h, w = (h + 1) // 2, (w + 1) // 2
            elif L.kind == "flatten":
                feat = h * w * c
            elif L.kind == "dense":
                flops += 2 * feat * a["units"]
                feat = a["units"]
        return flops


def _ch(base: int, width: float) -> int:
    return max(1, int(round(base * width)))


def _build(arch: str, width: float, layers: List[LayerSpec], num_classes: int = 10) -> ModelSpec:
    params: List[ParamSpec] = []
    off = 0
    h, w = 28, 28
    c = 1
    feat = None

    def add(name, shape, init, fi=0, fo=0):
        nonlocal off
        params.append(ParamSpec(name, tuple(shape), off, init, fi, fo))
        n = int(math.prod(shape))
        off += (n + ALIGN_ELEMS - 1) // ALIGN_ELEMS * ALIGN_ELEMS

    ci = di = 0
    for L in layers:
        a = L.args
        if L.kind == "conv":
            ci += 1
            k, co = a["k"], a["cout"]
            add(f"conv{ci}.kernel", (k, k, c, co), "glorot_uniform", k * k * c, k * k * co)
            add(f"conv{ci}.bias", (co,), "zeros")
            if a["padding"] == "VALID":
                h, w = h - k + 1, w - k + 1
            c = co
        elif L.kind == "pool":
            h, w = (h + 1) // 2, (w + 1) // 2
        elif L.kind == "flatten":
            feat = h * w * c
        elif L.kind == "dense":
            di += 1
            u = a["units"]
            nm = a.get("name", f"dense{di}")
            add(f"{nm}.kernel", (feat, u), "glorot_uniform", feat, u)
            add(f"{nm}.bias", (u,), "zeros")
            feat = u
    return ModelSpec(arch=arch, width=width, layers=tuple(layers), params=tuple(params), total=off, num_classes=num_classes)


def mnist_cnn(width: float = 1.0) -> ModelSpec:
    c1, c2, d = _ch(32, width), _ch(64, width), _ch(1024, width)
    return _build("mnist_cnn", width, [
        LayerSpec("conv", dict(k=5, cout=c1, padding="SAME", relu=True)),
        LayerSpec("pool"),
        LayerSpec("conv", dict(k=5, cout=c2, padding="SAME", relu=True)),
        LayerSpec("pool"),
        LayerSpec("flatten"),
        LayerSpec("dense", dict(units=d, relu=True, name="dense")),
        LayerSpec("dropout", dict(rate=0.5)),
        LayerSpec("dense", dict(units=10, relu=False, name="out")),
    ])


def keras128(width: float = 1.0) -> ModelSpec:
    c1, c2, d = _ch(32, width), _ch(64, width), _ch(128, width)
    return _build("keras128", width, [
        LayerSpec("conv", dict(k=3, cout=c1, padding="VALID", relu=True)),
        LayerSpec("conv", dict(k=3, cout=c2, padding="VALID", relu=True)),
        LayerSpec("pool"),
        LayerSpec("dropout", dict(rate=0.25)),
        LayerSpec("flatten"),
        LayerSpec("dense", dict(units=d, relu=True, name="dense")),
        LayerSpec("dropout", dict(rate=0.5)),
        LayerSpec("dense", dict(units=10, relu=False, name="out")),
    ])


def custom(name: str, layers: List[LayerSpec], width: float = 1.0) -> ModelSpec:
    """A user-defined 28x28 CNN from a layer list (``conv`` k/cout/padding/relu, ``pool``, ``flatten``, ``dense``
    units/relu/name, ``dropout`` rate).  It gets the same flat parameter arena, checkpoints, collectives and training
    loops as the built-in families; only ``mnist_cnn`` and ``keras128`` have hand-written sm_100a kernels, so a custom
    spec trains through :class:`b200dp.train.engine.TorchEngine` (on the GPU when one is present)."""
    if name in ARCHS:
        raise ValueError(f"{name!r} is a built-in architecture")
    kinds = {"conv", "pool", "flatten", "dense", "dropout"}
    seen_flat = False
    for L in layers:
        if L.kind not in kinds:
            raise ValueError(f"unknown layer kind {L.kind!r}; choose from {sorted(kinds)}")
        if L.kind == "flatten":
            seen_flat = True
        if L.kind == "dense" and not seen_flat:
            raise ValueError("a dense layer needs a flatten layer before it")
        if L.kind in ("conv", "pool") and seen_flat:
            raise ValueError(f"{L.kind} after flatten")
        if L.kind == "conv":
            L.args.setdefault("padding", "SAME")
            L.args.setdefault("relu", True)
            if L.args["padding"] not in ("SAME", "VALID"):
                raise ValueError("conv padding must be SAME or VALID")
        if L.kind == "dense":
            L.args.setdefault("relu", False)
    dense = [L for L in layers if L.kind == "dense"]
    if not dense:
        raise ValueError("the model needs a final dense layer producing the class logits")
    return _build(name, width, list(layers), num_classes=int(dense[-1].args["units"]))


ARCHS = {"mnist_cnn": mnist_cnn, "keras128": keras128}


def get_spec(arch: str = "mnist_cnn", width: float = 1.0) -> ModelSpec:
    try:
        return ARCHS[arch](width)
    except KeyError:
        raise ValueError(f"unknown arch {arch!r}; choose from {sorted(ARCHS)}") from None


def init_flat(spec: ModelSpec, seed: int = 0, device="cpu") -> torch.Tensor:
    """Random init (Glorot uniform kernels, zero biases — the TF/Keras defaults
    the reference gets from ``layers.conv2d`` / ``layers.dense``)."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    flat = torch.zeros(spec.total, dtype=torch.float32)
    for p in spec.params:
        if p.init == "glorot_uniform":
            lim = math.sqrt(6.0 / (p.fan_in + p.fan_out))
            flat[p.offset:p.offset + p.numel] = (torch.rand(p.numel, generator=g) * 2 - 1) * lim
    return flat.to(device)
