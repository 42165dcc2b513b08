#!/usr/bin/env python
"""Run the conv2 forward+dgrad kernels a few times eagerly (target for ncu -k regex:conv_halo)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from b200dp.ops import kernels as K
B = int(sys.argv[1]) if len(sys.argv) > 1 else 128
bf16 = torch.bfloat16
x = torch.randn(B, 14, 14, 32, device="cuda").to(bf16)
w = (torch.randn(5, 5, 32, 64, device="cuda") * 0.05).to(bf16)
b = torch.zeros(64, device="cuda", dtype=bf16)
out = torch.empty(B, 7, 7, 64, device="cuda", dtype=bf16)
am = torch.empty(B, 7, 7, 64, device="cuda", dtype=torch.uint8)
dy = torch.randn(B, 14, 14, 64, device="cuda").to(bf16)
din = torch.empty(B, 14, 14, 32, device="cuda", dtype=bf16)
for _ in range(3):
    K.conv_tc_fwd(x, w, b, out, am, relu=True, pool=True)
    K.conv_tc_dgrad(dy, w, din)
torch.cuda.synchronize()
print("ok", float(out.float().abs().mean()), float(din.float().abs().mean()))
