"""fp32 PyTorch restatement of the reference network — the oracle every
sm_100a kernel is tested against, and the model the CPU/gloo engine trains.

Mirrors ``conv_model`` (``horovod/tensorflow_mnist.py:38-73``) and
``MNISTModel.call`` (``horovod/tensorflow_mnist_gpu.py:68-88``): NHWC semantics,
TF weight layouts, SAME/VALID padding, max-pool 2x2/2, dropout only in training,
softmax cross-entropy averaged over the batch.
"""
from __future__ import annotations

from typing import Dict, Optional, Tuple

import torch
import torch.nn.functional as F

from ..utils.philox import dropout_keep_mask
from .spec import ModelSpec


def forward(spec: ModelSpec, params: Dict[str, torch.Tensor], x: torch.Tensor, *, training: bool,
            seed: int = 0, step: int = 0, dropout: bool = True,
            compute_dtype: Optional[torch.dtype] = None) -> torch.Tensor:
    """``x``: [B, 784] (or [B,28,28]) in [0,1].  Returns logits [B, 10] (fp32).

    ``compute_dtype=torch.bfloat16`` rounds activations and weights to bf16 at the
    same points the CUDA engine does (fp32 accumulate), for tight kernel tests.
    """
    def q(t):
        return t.to(compute_dtype).float() if compute_dtype is not None else t

    B = x.shape[0]
    h = x.reshape(B, 1, spec.input_hw[0], spec.input_hw[1]).float()
    ci = di = 0
    drop_id = 0
    flat = None
    for L in spec.layers:
        a = L.args
        if L.kind == "conv":
            ci += 1
            wk = q(params[f"conv{ci}.kernel"]).permute(3, 2, 0, 1)      # HWIO -> OIHW
            pad = a["k"] // 2 if a["padding"] == "SAME" else 0
            h = F.conv2d(q(h), wk, params[f"conv{ci}.bias"].float(), padding=pad)
            if a["relu"]:
                h = F.relu(h)
        elif L.kind == "pool":
            h = F.max_pool2d(h, 2, 2, ceil_mode=True)
        elif L.kind == "flatten":
            flat = h.permute(0, 2, 3, 1).reshape(B, -1)                   # NHWC flatten, as TF
        elif L.kind == "dense":
            di += 1
            nm = a.get("name", f"dense{di}")
            flat = q(flat) @ q(params[f"{nm}.kernel"]) + params[f"{nm}.bias"].float()
            if a["relu"]:
                flat = F.relu(flat)
        elif L.kind == "dropout":
            if training and dropout and a["rate"] > 0:
                tgt = flat if flat is not None else h
                if flat is None:
                    # conv-feature dropout (keras128): mask over the NHWC-flattened activation
                    nhwc = h.permute(0, 2, 3, 1)
                    keep = dropout_keep_mask(nhwc.numel(), a["rate"], seed, step, drop_id).view(nhwc.shape)
                    h = (nhwc * keep.to(h.device) / (1.0 - a["rate"])).permute(0, 3, 1, 2)
                else:
                    keep = dropout_keep_mask(tgt.numel(), a["rate"], seed, step, drop_id).view(tgt.shape)
                    flat = tgt * keep.to(tgt.device) / (1.0 - a["rate"])
            drop_id += 1
    return flat


def loss_and_metrics(logits: torch.Tensor, labels: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """Mean softmax cross-entropy and the number of correct argmax predictions."""
    labels = labels.long()
    loss = F.cross_entropy(logits.float(), labels)
    correct = (logits.argmax(dim=1) == labels).sum()
    return loss, correct


def adam_step_(param: torch.Tensor, grad: torch.Tensor, m: torch.Tensor, v: torch.Tensor, step: int, *,
               lr: float, beta1: float = 0.9, beta2: float = 0.999, eps: float = 1e-8,
               eps_mode: str = "tf") -> None:
    """One Adam update in place; ``step`` is the 1-based step being applied.

    ``eps_mode="tf"``: TF1/Keras form, ``lr_t = lr*sqrt(1-b2^t)/(1-b1^t)``,
    ``p -= lr_t * m / (sqrt(v) + eps)`` (``tf.train.AdamOptimizer``,
    ``horovod/tensorflow_mnist.py:130``).  ``"torch"``: epsilon added after the
    bias-corrected sqrt.
    """
    m.mul_(beta1).add_(grad, alpha=1 - beta1)
    v.mul_(beta2).addcmul_(grad, grad, value=1 - beta2)
    bc1 = 1 - beta1 ** step
    bc2 = 1 - beta2 ** step
    if eps_mode == "tf":
        lr_t = lr * (bc2 ** 0.5) / bc1
        param.addcdiv_(m, v.sqrt().add_(eps), value=-lr_t)
    elif eps_mode == "torch":
        param.addcdiv_(m, (v.sqrt() / (bc2 ** 0.5)).add_(eps), value=-lr / bc1)
    else:
        raise ValueError(f"eps_mode {eps_mode!r}")
