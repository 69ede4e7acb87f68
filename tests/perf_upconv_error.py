"""Forward error of the upconv kernel against the fp64 oracle at three decoder widths (and of the CPU fp32 / cuDNN fp32
sequences, for scale): the figure behind the drained accumulators of upconv.cu."""
import sys, torch
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bts_b200 import functional as Fn
from oracle import torch_port as tp
for cin, cout, B, h, w in ((64, 32, 1, 24, 32), (512, 256, 1, 15, 20), (2208, 512, 1, 7, 10)):
    gen = torch.Generator().manual_seed(7 + cin)
    x = torch.randn(B, cin, h, w, generator=gen)
    wt = torch.randn(cout, cin, 3, 3, generator=gen) * (2.0 / (9 * cin)) ** 0.5
    ref = tp.upconv_forward(x.double(), wt.double())
    ref32 = tp.upconv_forward(x, wt).double()
    with torch.no_grad():
        y = Fn.upconv_forward(x.cuda(), wt.cuda()).cpu().double()
        torch.backends.cudnn.allow_tf32 = False
        yc = tp.upconv_forward(x.cuda(), wt.cuda()).cpu().double()
    pk = ref.abs().max()
    print("%4d->%3d  ours %.2e  (signed mean err / peak %.2e)  cpu fp32 %.2e  cudnn fp32 %.2e" % (
        cin, cout, (y - ref).abs().max() / pk, ((y - ref) * ref.sign()).mean() / pk, (ref32 - ref).abs().max() / pk, (yc - ref).abs().max() / pk))
