"""upconv2 of the decoder at nyu_b8 ([8,128,120,160] -> [8,64,240,320]), forward + backward, three times: the case
`ncu --set full -k regex:upconv_fwd_kernel|upconv_dgrad|upconv_wgrad --launch-skip 3 -c 3` profiles."""
import sys, torch
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bts_b200 import functional as Fn
x = torch.randn(8, 128, 120, 160, device="cuda", requires_grad=True)
wt = (torch.randn(64, 128, 3, 3, device="cuda") * 0.03).requires_grad_(True)
gy = torch.randn(8, 64, 240, 320, device="cuda")
for _ in range(3):
    x.grad = wt.grad = None
    Fn.upconv_forward(x, wt).backward(gy)
torch.cuda.synchronize()
