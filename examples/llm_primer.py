"""Single-device LLaMA training loop (lab/tutorial_1b/primer/intro.py): dmodel 288, 6 heads, 6 layers, seq 256,
batch 3, Adam 8e-4 — on the synthetic token stream; the loss runs through the fused softmax-xent kernel on GPU."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ddl25spring_b200.core.data import SimpleTokenizer, SyntheticTokens
from ddl25spring_b200.models.llama import CausalLLama, LLama, causalLLMLoss


def main(iters=50, dmodel=288, num_heads=6, n_layers=6, seq_l=256, batch_size=3, vocab=32_000, device=None, log=print):
    device = device or ("cuda" if torch.cuda.is_available() else "cpu")
    torch.manual_seed(0)
    tok = SimpleTokenizer(vocab)
    net = LLama(CausalLLama, tok.vocab_size, dmodel=dmodel, num_heads=num_heads, device=device, n_layers=n_layers,
                ctx_size=seq_l, padding_idx=tok.pad_id)
    ds = iter(SyntheticTokens(tok.vocab_size, batch_size, seq_l))
    optim = torch.optim.Adam(net.parameters(), lr=8e-4)
    losses = []
    for itr in range(iters):
        optim.zero_grad()
        x = next(ds).to(device)
        loss = causalLLMLoss(net(x), x, tok.vocab_size)
        loss.backward()
        optim.step()
        losses.append(float(loss))
        if log:
            log(itr, losses[-1])
    return losses


if __name__ == "__main__":
    main(int(sys.argv[1]) if len(sys.argv) > 1 else 50)
