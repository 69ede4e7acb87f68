"""One forward + backward of each of the five decoder upconvs at nyu_b8 (the command `ncu --metrics gpu__time_duration.sum
-k regex:upconv` is run on for a per-kernel breakdown)."""
import sys, torch
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bts_b200 import functional as Fn
B, H, W = 8, 480, 640
for name, cin, cout, div in (("upconv5", 2208, 512, 32), ("upconv4", 512, 256, 16), ("upconv3", 256, 128, 8), ("upconv2", 128, 64, 4), ("upconv1", 64, 32, 2)):
    h, w = H // div, W // div
    x = torch.randn(B, cin, h, w, device="cuda", requires_grad=True)
    wt = (torch.randn(cout, cin, 3, 3, device="cuda") * (2.0 / (9 * cin)) ** 0.5).requires_grad_(True)
    gy = torch.randn(B, cout, 2 * h, 2 * w, device="cuda")
    for _ in range(2):
        x.grad = wt.grad = None
        Fn.upconv_forward(x, wt).backward(gy)
    torch.cuda.synchronize()
