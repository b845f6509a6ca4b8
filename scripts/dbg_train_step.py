import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn.functional as F
from ddl25spring_b200.core.arena import FlatArena
from ddl25spring_b200.models.mnist_cnn import MnistCnn
from ddl25spring_b200.fl.packed import PackedMnistEngine
DEV = "cuda"
def mk(seed):
    torch.manual_seed(seed); m = MnistCnn(0.0, 0.0).to(DEV); return m, FlatArena.from_module(m)
def grads_of(m, x, y):
    m.train(); m.zero_grad(); F.nll_loss(m(x), y).backward(); return [p.grad.detach().clone() for p in m.parameters()]
G, B, bszs, lr = 3, 100, [100, 100, 37], 0.05
models, arenas = zip(*[mk(10 + g) for g in range(G)])
W = torch.stack([a.flat for a in arenas]).contiguous()
X = torch.randn(G, B, 784, device=DEV); Y = torch.randint(0, 10, (G, B), device=DEV)
perm = torch.stack([torch.randperm(B, device=DEV) for _ in range(G)]).int().contiguous()
eng = PackedMnistEngine(G, B, p_drop1=0.0, p_drop2=0.0); eng.prep_weights(W); W0 = W.clone()
eng.train_step(W, X, B * 784, Y, B, perm, B, 0, bszs, lr, seed=1); torch.cuda.synchronize()
for g in range(G):
    x = X[g][perm[g, :bszs[g]].long()].view(-1, 1, 28, 28); y = Y[g][perm[g, :bszs[g]].long()]
    torch.backends.cudnn.allow_tf32 = False
    g32 = grads_of(models[g], x, y)
    torch.backends.cudnn.allow_tf32 = True
    gtf = grads_of(models[g], x, y)
    got_flat = (W0[g] - W[g]) / lr
    for spec, a, b in zip(arenas[g].specs, g32, gtf):
        got = got_flat[spec.offset:spec.offset + spec.numel].view(spec.shape)
        mx = a.abs().max()
        print(g, spec.name, "ours-vs-fp32 %.2e" % float((got - a).abs().max() / mx), " cudnnTF32-vs-fp32 %.2e" % float((b - a).abs().max() / mx),
              " rel-fro ours %.2e cudnn %.2e" % (float((got - a).norm() / a.norm()), float((b - a).norm() / a.norm())))
