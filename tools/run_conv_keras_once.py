#!/usr/bin/env python
"""Run the keras128 second-layer convolutions (3x3 VALID, 26x26x32 -> 24x24x64) a few times eagerly: forward(+pool),
dgrad, wgrad.  Target for `ncu -k regex:conv_res_kernel`."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from b200dp.ops import kernels as K
B = int(sys.argv[1]) if len(sys.argv) > 1 else 128
bf16 = torch.bfloat16
x = torch.randn(B, 26, 26, 32, device="cuda").to(bf16)
w = (torch.randn(3, 3, 32, 64, device="cuda") * 0.05).to(bf16)
b = torch.zeros(64, device="cuda", dtype=bf16)
out = torch.empty(B, 12, 12, 64, device="cuda", dtype=bf16)
am = torch.empty(B, 12, 12, 64, device="cuda", dtype=torch.uint8)
dy = torch.randn(B, 24, 24, 64, device="cuda").to(bf16)
din = torch.empty(B, 26, 26, 32, device="cuda", dtype=bf16)
dw = torch.zeros(3, 3, 32, 64, device="cuda")
for _ in range(3):
    K.conv_tc_fwd(x, w, b, out, am, relu=True, pool=True, pad=0)
    K.conv_tc_dgrad(dy, w, din, pad=0)
    K.conv_tc_wgrad(x, dy, dw, None, pad=0)
torch.cuda.synchronize()
print("ok", float(out.float().abs().mean()), float(din.float().abs().mean()))
