"""Prefill tokens/s of a random-init Llama/Qwen-shaped decoder stack on the NVFP4+ARC hot path (BASELINE.json configs
3 and 5; reference: benchmarks/benchmark_e2e_arc.py).  One JSON line.

  python scripts/prefill_bench.py --model qwen2.5-7b --bsz 4 --seq 1024 [--layers N] [--steps 5]
  torchrun --nproc-per-node 8 scripts/prefill_bench.py --model qwen2.5-32b --bsz 8 --seq 2048      # tensor parallel
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from arcquant_b200 import agemm, prefill  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="qwen2.5-7b", choices=sorted(prefill.MODELS))
    ap.add_argument("--bsz", type=int, default=4)
    ap.add_argument("--seq", type=int, default=1024)
    ap.add_argument("--layers", type=int, default=None)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--unfused", action="store_true", help="the reference's launch structure: 7 GEMMs + torch glue per layer")
    args = ap.parse_args()
    world, rank, local = (int(os.environ.get(k, d)) for k, d in (("WORLD_SIZE", "1"), ("RANK", "0"), ("LOCAL_RANK", "0")))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    cfg = prefill.PrefillConfig.named(args.model, args.layers)
    model = prefill.ArcPrefillModel(cfg, dev, world=world, rank=rank, fused=not args.unfused)
    h = torch.randn(args.bsz, args.seq, cfg.hidden, device=dev, generator=torch.Generator(device=dev).manual_seed(1)).to(torch.bfloat16)
    for _ in range(args.warmup):
        model(h)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    n0 = agemm.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        model(h)
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item()) / args.steps
    tokens = args.bsz * args.seq
    if rank == 0:
        print(json.dumps({
            "metric": "prefill_tokens_per_s", "value": tokens / (ms * 1e-3), "unit": "tokens/s", "n_gpus": world,
            "ms_per_forward": ms, "linear_tflops": model.linear_flops(tokens) / (ms * 1e-3) / 1e12,
            "gpu_launches_per_forward": (agemm.launch_count() - n0) / args.steps, "data": "synthetic, random-init weights",
            "config": {"workload": "%s decoder stack prefill (speed-model shape, MHA), %d layers" % (args.model, cfg.layers),
                       "bsz": args.bsz, "seq": args.seq, "parallelism": "tp%d" % world if world > 1 else "single GPU",
                       "attention": "torch SDPA (not on the hot path)",
                       "launches": "fused q/k/v + SwiGLU-epilogue gate/up + in-kernel pre-scale" if not args.unfused else "reference structure (7 GEMMs per layer)"}}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
