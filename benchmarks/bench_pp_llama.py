#!/usr/bin/env python
"""Pipeline-parallel LLaMA (lab/tutorial_1b/PP: 6 x 288, ctx 256, vocab 32 000, Adam) over NCCL p2p.

  torchrun --nproc-per-node S benchmarks/bench_pp_llama.py [--schedule 1f1b --micro 8 --kernels]

S stages, `--micro` micro-batches of one sequence per iteration; reports tokens/s (device-timed, max over ranks) for the
plain PyTorch stage modules, for the kernel path (`ops/nn.use_kernels`) and for the kernel path with every micro-batch op
replayed from a CUDA graph (`PipelineStage(graph=True)`: forward graph + recomputing forward-backward graph per stage)."""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--schedule", default="1f1b", choices=["gpipe", "1f1b"])
    ap.add_argument("--micro", type=int, default=8)
    ap.add_argument("--iters", type=int, default=10)
    a = ap.parse_args()
    from ddl25spring_b200 import ops
    from ddl25spring_b200.models.llama import LLama, causalLLMLoss
    from ddl25spring_b200.ops.nn import kernel_causal_lm_loss, use_kernels
    from ddl25spring_b200.parallel.pp import Mesh, PipelineStage
    from ddl25spring_b200.core.arena import FlatArena
    from ddl25spring_b200.utils.timing import max_over_ranks
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
    world, rank = dist.get_world_size(), dist.get_rank()
    V = 32_000
    tokens = torch.randint(0, V, (a.micro, 256), generator=torch.Generator().manual_seed(1)).cuda()
    mbs = list(tokens.split(1))
    out = {"stages": world, "schedule": a.schedule, "micro_batches": a.micro, "ms_per_iter": {}, "tokens_per_s": {}}
    for arm in ("torch", "kernels", "kernels_graphed"):
        torch.manual_seed(0)
        mesh = Mesh(world, 1)
        stage = LLama(None, V, 288, 6, "cuda", 6, 256).split_stages(world)[mesh.stage_id]
        if arm.startswith("kernels"):
            use_kernels(stage)
        loss_fn = (lambda o, t: kernel_causal_lm_loss(o, t)) if arm.startswith("kernels") else (lambda o, t: causalLLMLoss(o, t, V))
        st = PipelineStage(stage, mesh, (1, 256, 288), schedule=a.schedule, n_micro=a.micro, loss_fn=loss_fn, device="cuda",
                           graph=arm.endswith("graphed"))
        arena = st.arena                      # the stage keeps parameters / gradients in one flat arena each
        m, v = torch.zeros_like(arena.flat), torch.zeros_like(arena.flat)

        def it(step):
            arena.grad.zero_()
            st.run(inputs=mbs, targets=mbs)
            with torch.no_grad():
                ops.adam_step(arena.flat, arena.grad, m, v, step, 8e-4, 0.9, 0.999, 1e-8, 0.0)

        for s in range(3):
            it(s + 1)
        torch.cuda.synchronize(); dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for s in range(a.iters):
            it(s + 4)
        e1.record(); torch.cuda.synchronize()
        ms = max_over_ranks(e0.elapsed_time(e1) / a.iters, "cuda")
        out["ms_per_iter"][arm] = round(ms, 3)
        out["tokens_per_s"][arm] = round(a.micro * 256 / ms * 1e3, 1)
    if rank == 0:
        print(json.dumps(out))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
