"""Pipeline-parallel LLaMA training (lab/tutorial_1b/PP/1F1B): sequential, GPipe micro-batches, 1F1B, and the
pipelines x stages hybrid the reference could not get to run (homework-1 cell 48).

  python -m torch.distributed.run --nproc-per-node 3 --master-addr 127.0.0.1 examples/pp_train.py --schedule gpipe --micro 3
  python -m torch.distributed.run --nproc-per-node 6 --master-addr 127.0.0.1 examples/pp_train.py --schedule 1f1b --pipelines 2"""
import argparse, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from ddl25spring_b200.core.data import SyntheticTokens
from ddl25spring_b200.models.llama import LLama, causalLLMLoss
from ddl25spring_b200.parallel.pp import Mesh, PipelineStage
from ddl25spring_b200.utils.trace import Watchdog


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--schedule", default="1f1b", choices=["sequential", "gpipe", "1f1b"])
    ap.add_argument("--micro", type=int, default=3)
    ap.add_argument("--pipelines", type=int, default=1)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--dmodel", type=int, default=288)
    ap.add_argument("--vocab", type=int, default=32_000)
    ap.add_argument("--kernels", action="store_true", help="run the stage modules on the hand-written kernels (GPU only)")
    a = ap.parse_args()
    cuda = torch.cuda.is_available()
    if cuda:
        torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    dist.init_process_group("nccl" if cuda else "gloo")
    dev = "cuda" if cuda else "cpu"
    world = dist.get_world_size()
    mesh = Mesh(world // a.pipelines, a.pipelines)
    torch.manual_seed(0)
    full = LLama(None, a.vocab, a.dmodel, 6, dev, 6, 256)
    stage = full.split_stages(mesh.n_stages)[mesh.stage_id]
    if a.kernels and cuda:
        from ddl25spring_b200.ops.nn import use_kernels
        use_kernels(stage)
    n_micro = 1 if a.schedule == "sequential" else a.micro
    st = PipelineStage(stage, mesh, (1, 256, a.dmodel), schedule=a.schedule, n_micro=n_micro,
                       loss_fn=lambda o, t: causalLLMLoss(o, t, a.vocab), device=dev)
    opt = torch.optim.Adam(stage.parameters(), lr=8e-4)
    ds = iter(SyntheticTokens(a.vocab, n_micro, 256, skip=mesh.pipeline_id * 5000))   # stage 0 and the last stage stay in lock-step
    wd = Watchdog(120, what="pipeline iteration").start()
    for itr in range(a.iters):
        opt.zero_grad()
        batch = next(ds)
        mbs = list(batch.split(1))
        losses = st.run(inputs=mbs, targets=mbs)
        st.allreduce_grads()
        opt.step()
        wd.beat()
        if st.is_last:
            print(itr, sum(losses), flush=True)
    wd.stop()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
