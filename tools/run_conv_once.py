#!/usr/bin/env python
"""Run the conv2 forward/dgrad/wgrad and dense kernels a few times eagerly (target for ncu -k regex:...)."""
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
dw = torch.zeros(5, 5, 32, 64, device="cuda")
a2 = torch.randn(B, 3136, device="cuda").to(bf16); dz = torch.randn(B, 1024, device="cuda").to(bf16)
gd = torch.empty(3136, 1024, device="cuda")
for _ in range(3):
    K.conv_tc_fwd(x, w, b, out, am, relu=True, pool=True)
    K.conv_tc_dgrad(dy, w, din)
    K.conv_tc_wgrad(x, dy, dw, None)
    K.gemm_tc(a2, dz, gd, 3136, 1024, B, a_mn_major=True, b_mn_major=True, out_mode=0, block_n=128)
torch.cuda.synchronize()
print("ok", float(out.float().abs().mean()), float(din.float().abs().mean()))
