"""Three training steps of ResNet18NHWC (batch 256) and nothing else: run under `ncu --metrics gpu__time_duration.sum` to get the
per-kernel GPU time of one step (scripts/agg_launches.py summarises the CSV)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F
from ddl25spring_b200.ops.nn import kernel_classification_loss
from ddl25spring_b200.models.resnet_nhwc import ResNet18NHWC

torch.manual_seed(0)
B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
net = ResNet18NHWC(10).cuda()
x = torch.randn(B, 3, 32, 32, device="cuda"); t = torch.randint(0, 10, (B,), device="cuda")
for _ in range(3):
    for p in net.parameters():
        p.grad = None
    kernel_classification_loss(net(x), t).backward()
torch.cuda.synchronize()
print("ok")
