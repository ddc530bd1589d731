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
    ap.add_argument("--no-sp", action="store_true", help="world > 1: replicated activations + NCCL all-reduce instead of the "
                                                         "sequence-parallel path over peer memory")
    ap.add_argument("--no-graph", action="store_true", help="world > 1, sequence parallel: launch every forward from Python")
    ap.add_argument("--check", action="store_true", help="world > 1: compare the sequence-parallel output with the all-reduce path")
    args = ap.parse_args()
    world, rank, local = (int(os.environ.get(k, d)) for k, d in (("WORLD_SIZE", "1"), ("RANK", "0"), ("LOCAL_RANK", "0")))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    cfg = prefill.PrefillConfig.named(args.model, args.layers)
    model = prefill.ArcPrefillModel(cfg, dev, world=world, rank=rank, fused=not args.unfused)
    h = torch.randn(args.bsz, args.seq, cfg.hidden, device=dev, generator=torch.Generator(device=dev).manual_seed(1)).to(torch.bfloat16)
    tokens = args.bsz * args.seq
    sp, note = False, None
    if world > 1 and not args.no_sp and not args.unfused:
        ok = 1
        try:
            model.enable_seq_parallel(dev, tokens)
        except Exception as e:  # noqa: BLE001
            ok, note = 0, "sequence-parallel path unavailable (%s): all-reduce path" % str(e).splitlines()[0][:160]
        flag = torch.tensor([ok], dtype=torch.int32, device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        sp = bool(int(flag.item()))

    def run():
        return model.forward_seq_parallel(h, gather_output=False) if sp else model(h)

    check = None
    if args.check and sp:
        ref = model(h).reshape(tokens, cfg.hidden)
        got = model.forward_seq_parallel(h).reshape(tokens, cfg.hidden)
        torch.cuda.synchronize()
        model.xc.group.status()
        check = {"rel_fro_vs_allreduce_path": float((got.float() - ref.float()).norm() / ref.float().norm()),
                 "finite": bool(torch.isfinite(got.float()).all())}
    for _ in range(args.warmup):
        run()
    torch.cuda.synchronize()
    graph = None
    if (sp or world == 1) and not args.no_graph:   # a shard's kernels are short: replay the whole forward (only this library's kernels + torch glue)
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=side, capture_error_mode="thread_local"):
                run()
        torch.cuda.current_stream().wait_stream(side)
        graph.replay()
        torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    n0 = agemm.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        if graph is not None:
            graph.replay()
        else:
            run()
    e1.record()
    torch.cuda.synchronize()
    launches = (agemm.launch_count() - n0) / args.steps
    if sp:
        model.xc.group.status()
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item()) / args.steps
    if rank == 0:
        par = "single GPU, %s" % ("CUDA-graph replay" if graph is not None else "launched from Python")
        if world > 1:
            par = ("tp%d, sequence parallel over peer memory (scatter-epilogue GEMMs + slot sums, codes gathered by peer stores, "
                   "residual adds fused into the norms), %s" % (world, "CUDA-graph replay" if graph is not None else "launched from Python")
                   ) if sp else "tp%d, replicated activations + NCCL all-reduce" % world
        out = {
            "metric": "prefill_tokens_per_s", "value": tokens / (ms * 1e-3), "unit": "tokens/s", "n_gpus": world,
            "ms_per_forward": ms, "linear_tflops": model.linear_flops(tokens) / (ms * 1e-3) / 1e12,
            "gpu_launches_per_forward": launches if graph is None else None, "data": "synthetic, random-init weights",
            "config": {"workload": "%s decoder stack prefill (speed-model shape, MHA), %d layers" % (args.model, cfg.layers),
                       "bsz": args.bsz, "seq": args.seq, "parallelism": par,
                       "attention": "torch SDPA (not on the hot path)",
                       "launches": "fused q/k/v (+ bias) + SwiGLU-epilogue gate/up + in-kernel pre-scale + fused residual add / norm / quantise"
                       if not args.unfused else "reference structure (7 GEMMs per layer)"}}
        if check is not None:
            out["check"] = check
        if note:
            out["note"] = note
        print(json.dumps(out))
    sys.stdout.flush()
    if world > 1:
        os._exit(0)   # no NCCL teardown after graph replays (see bench.py::_leave)


if __name__ == "__main__":
    main()
