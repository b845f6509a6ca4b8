"""One NHWC 3x3 convolution (forward, dgrad, wgrad) at a ResNet layer shape: target for `ncu --set full -k regex:conv_`."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ddl25spring_b200.ops import conv as K

N, H, Cin, Cout = [int(v) for v in (sys.argv[1:5] if len(sys.argv) > 4 else (64, 32, 128, 128))]
x = torch.randn(N, H, H, Cin, device="cuda")
w = torch.randn(Cout, 3, 3, Cin, device="cuda")
with torch.no_grad():
    for _ in range(3):
        y = K._fprop(x, None, w, 1, 1)
        gx = K._dgrad(y, w, 1, 1, H, H)
        gw = K._wgrad(y, x, None, tuple(w.shape), 1, 1)
torch.cuda.synchronize()
print("ok", float(y.abs().mean()), float(gx.abs().mean()), float(gw.abs().mean()))
