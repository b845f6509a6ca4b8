"""Vertical-FL tutorial + homework-2 exercises (lab/tutorial_2b/vfl.py, exercise_1/2/3.py) on the heart table.

  python examples/vfl_exercises.py [--csv path/to/heart.csv] [--epochs 300]"""
import argparse, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ddl25spring_b200.core.data import SyntheticHeart
from ddl25spring_b200.models.tabular import ClientDecoder, ClientEncoder, ServerVAE, VFLVAE
from ddl25spring_b200.parallel import vfl


def run(table, parts, epochs, seed=42, **kw):
    torch.manual_seed(seed)
    net = vfl.build_vfl(table, parts, **kw)
    xtr, ytr, xte, yte = table.x[:821], table.y[:821], table.x[821:], table.y[821:]      # vfl.py:151-152
    hist = net.train_with_settings(epochs, 64, len(parts), parts, xtr, ytr)
    acc, loss = net.test(xte, yte)
    return float(acc), float(loss), hist


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--csv", default=None)
    ap.add_argument("--epochs", type=int, default=300)
    ap.add_argument("--vae_epochs", type=int, default=1000)
    a = ap.parse_args(argv)
    table = (SyntheticHeart.load_csv(a.csv) if a.csv else SyntheticHeart()).encoded()
    out = {}
    acc, loss, _ = run(table, vfl.contiguous_raw_partition(table, 4), a.epochs)
    out["tutorial_4_parties"] = (acc, loss)
    print(f"tutorial (4 parties, contiguous raw columns): test acc {acc * 100:.2f}% loss {loss:.3f}")
    for seed in (0, 1, 2):                                                                # exercise 1
        acc, loss, _ = run(table, vfl.random_permutation_partition(table.columns, 4, seed), a.epochs)
        out[f"ex1_perm{seed}"] = (acc, loss)
        print(f"exercise 1, permutation {seed}: test acc {acc * 100:.2f}% loss {loss:.3f}")
    for n in (2, 4, 6, 8):                                                                # exercise 2
        acc, loss, _ = run(table, vfl.balanced_partition(table.columns, n), a.epochs)
        out[f"ex2_{n}_parties"] = (acc, loss)
        print(f"exercise 2, {n} parties: test acc {acc * 100:.2f}% loss {loss:.3f}")
    data = torch.cat([table.x, table.y[:, None].float()], 1)                              # exercise 3
    data = (data - data.mean(0)) / data.std(0).clamp_min(1e-6)
    counts = vfl.balanced_counts(data.shape[1], 4)
    model = VFLVAE([ClientEncoder(c, 8) for c in counts], ServerVAE(32), [ClientDecoder(8, c) for c in counts], 8)
    losses = vfl.train_vfl_vae(model, list(data.split(counts, dim=1)), epochs=a.vae_epochs)
    out["ex3_vae"] = (losses[0], losses[-1])
    print(f"exercise 3 (VFL-VAE): loss {losses[0]:.2f} -> {losses[-1]:.2f}")
    return out


if __name__ == "__main__":
    main()
