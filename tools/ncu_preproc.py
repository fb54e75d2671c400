"""ncu target: a few launches of the pre-processing kernels on a 64-frame batch (640x480x3 -> 432x368x3).
usage: python tools/ncu_preproc.py [resize_path]      (0 = automatic, 3 = the plain strip form, ...)"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tfpose_b200 as tp  # noqa: E402

d = tp.PoseDecoder(max_batch=1, max_upsample=(184, 216))
if len(sys.argv) > 1:
    d.set_option("resize_path", int(sys.argv[1]))
src = torch.randint(0, 256, (64, 480, 640, 3), dtype=torch.uint8, device="cuda")
x = torch.rand((64, 368, 432, 3), device="cuda") * 2 - 1
for _ in range(3):
    out = d.resize_cubic(src, (432, 368))
    q = d.quantize_img(x)
torch.cuda.synchronize()
print("ok", out.shape, q.shape)
d.close()
