"""Three LLaMA (6x288, vocab 32000, batch 3 x 256) training steps on the kernel path: target for an ncu launch list."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ddl25spring_b200.models.llama import LLama, causalLLMLoss
from ddl25spring_b200.ops.nn import kernel_causal_lm_loss, use_kernels
from ddl25spring_b200.parallel.dp import DataParallel
arm = sys.argv[1] if len(sys.argv) > 1 else "kernels"
torch.manual_seed(0)
net = LLama(None, 32000, 288, 6, "cuda", 6, 256)
if arm == "kernels":
    use_kernels(net)
loss_fn = (lambda m, t: kernel_causal_lm_loss(m(t), t)) if arm == "kernels" else (lambda m, t: causalLLMLoss(m(t), t, 32000))
dp = DataParallel(net, optimizer="adam", lr=8e-4)
tok = torch.randint(0, 32000, (3, 256), device="cuda")
for _ in range(3):
    loss = dp.train_step(loss_fn, tok)
torch.cuda.synchronize()
print("ok", float(loss))
