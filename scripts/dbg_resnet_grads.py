"""Per-parameter gradient error of ResNet18NHWC (tcgen05 TF32) and of the cuDNN-TF32 model, both against the fp32 model."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F
from ddl25spring_b200.models.resnet import ResNet18
from ddl25spring_b200.models.resnet_nhwc import ResNet18NHWC, from_torch_layout

torch.manual_seed(2)
B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
ref = ResNet18(10).cuda(); ref2 = ResNet18(10).cuda(); net = ResNet18NHWC(10).cuda()
ref2.load_state_dict(ref.state_dict()); net.load_state_dict(from_torch_layout(ref.state_dict()))
x = torch.randn(B, 3, 32, 32, device="cuda"); t = torch.randint(0, 10, (B,), device="cuda")
rel = lambda a, b: ((a - b).norm() / b.norm().clamp_min(1e-12)).item()
torch.backends.cudnn.allow_tf32 = False; torch.backends.cuda.matmul.allow_tf32 = False
y0 = ref(x); F.cross_entropy(y0, t).backward()
torch.backends.cudnn.allow_tf32 = True; torch.backends.cuda.matmul.allow_tf32 = True
y2 = ref2(x); F.cross_entropy(y2, t).backward()
y1 = net(x); F.cross_entropy(y1, t).backward()
print("logits rel: ours %.3e cudnn-tf32 %.3e" % (rel(y1, y0), rel(y2, y0)))
g0 = dict(ref.named_parameters()); g2 = dict(ref2.named_parameters())
for name, p in net.named_parameters():
    g = p.grad
    if g.dim() == 4:
        g = g.permute(0, 3, 1, 2)
        if name == "conv1.weight":
            g = g[:, :3]
    print("%-32s ours %.3e   cudnn-tf32 %.3e" % (name, rel(g, g0[name].grad), rel(g2[name].grad, g0[name].grad)))
