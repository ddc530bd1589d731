"""Paths that need MORE THAN ONE GPU (skipped on a one-GPU box; run with `gpurun --gpus 2`):

* the tensor-parallel data plane the scaling bench times -- reduce-scatter of bf16 partials, all-gather of codes + sliced
  scale blocks, token slabs on a communication stream, CUDA-graph replay -- checked end to end by bench.py's own oracle
  check (`parity_check`), at world size 2, for every exchange mode;
* two devices driven from ONE process: the per-device caches of cudaFuncSetAttribute / cluster occupancy.
"""
import json
import os
import socket
import subprocess
import sys

import pytest
import torch

from oracle import torchref as T

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _need2():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")


@pytest.mark.parametrize("args", [
    ["--parallelism", "tp"],                                   # default: exchange over peer memory (CUDA IPC), one slab, from Python
    ["--parallelism", "tp", "--graph"],                        # the same, replayed from a CUDA graph (device-resident barrier epochs)
    ["--parallelism", "tp", "--no-peer"],                      # NCCL reduce-scatter + all-gather
    ["--parallelism", "tp", "--slabs", "2", "--graph"],        # micro-batch slabs on a communication stream, CUDA-graph replay
    ["--parallelism", "tp", "--slabs", "4", "--unfused"],
    ["--parallelism", "tp-ar"],
    ["--parallelism", "tp", "--workload", "llama3-70b"],
])
def test_two_rank_step_parity(args):
    _need2()
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(_free_port()), os.path.join(ROOT, "bench.py"), "--gpus", "2", "--steps", "3", "--warmup", "1",
           "--tokens", "2048", "--e2e-steps", "1", "--no-cpu-baseline", "--no-prefill"] + args
    # own process group + hard kill on timeout: a hung rank must not keep the GPUs (and the box) busy
    p = subprocess.Popen(cmd, cwd=ROOT, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, start_new_session=True)
    try:
        out, err = p.communicate(timeout=240)
    except subprocess.TimeoutExpired:
        import signal
        os.killpg(p.pid, signal.SIGKILL)
        out, err = p.communicate()
        pytest.fail("2-rank bench timed out: " + err[-1500:])
    assert p.returncode == 0, err[-2000:]
    line = [l for l in out.splitlines() if l.startswith("{")][-1]
    out = json.loads(line)
    pc = out["parity_check"]
    assert pc["ranks"] == 2 and pc["ok"], pc
    assert pc["quant_bytes_differ"] == 0 and pc["tp_gather_bytes_differ"] == 0
    assert out["n_gpus"] == 2 and out["gpu_launches"] > 0 and out["scaling"] == "strong"
    peer_expected = "--no-peer" not in args and "--slabs" not in args and "tp-ar" not in args
    assert out["exchange"] == ("peer" if peer_expected else "nccl"), out.get("exchange_note")
    if peer_expected:
        assert pc["peer_barrier_timeouts"] == 0


def test_two_devices_from_one_process():
    """First launch of every kernel family on cuda:1 AFTER cuda:0 used it (per-device opt-in of > 48 KB dynamic shared
    memory, per-device cluster occupancy), and a call on a non-current device."""
    _need2()
    from arcquant_b200 import agemm
    res = []
    for d in (0, 1, 0, 1):
        dev = torch.device("cuda", d)
        g = torch.Generator(device=dev).manual_seed(7)
        M, KQ, KE, N = 2304, 4096, 256, 8192 + 256
        x = (torch.randn(M, KQ, device=dev, generator=g) * 30).to(torch.bfloat16)
        idx = torch.randperm(KQ, device=dev, generator=g).to(torch.int16)
        w = (torch.randn(N, KQ, device=dev, generator=g) * 30).to(torch.bfloat16)
        # deliberately NOT inside `with torch.cuda.device(d)`: the binding switches device itself
        qx, sfx = agemm.reorder_quantize_x(x, idx, KE)
        qw, sfw = agemm.reorder_quantize_w(w, idx, KE)
        y = agemm.matmul(qx, qw, sfx, sfw, 1e-4)                 # >= 74 pair tiles: cta_group::2 kernel
        y_small = agemm.matmul(qx[:64], qw, sfx, sfw, 1e-4)      # split-K cluster kernel
        y_mid = agemm.matmul(qx[:300], qw[:384], sfx, sfw, 1e-4)  # single-CTA kernel
        qn, sfn = agemm.rmsnorm_quantize_x(x, torch.ones(KQ, dtype=torch.bfloat16, device=dev), 1e-6, idx, KE)
        torch.cuda.synchronize(dev)
        assert y.device == dev and qn.device == dev
        worst, rel = T.check_gemm(y, qx, sfx, qw, sfw, 1e-4)
        assert worst <= 1.0 and rel < 2e-3
        w2, r2 = T.check_gemm(y_small, qx[:64], sfx, qw, sfw, 1e-4, torch.arange(64))
        w3, r3 = T.check_gemm(y_mid, qx[:300], sfx, qw[:384], sfw, 1e-4, torch.arange(300))
        assert max(w2, w3) <= 1.0
        res.append((qx.cpu(), y.cpu()))
    assert torch.equal(res[0][0], res[2][0]) and torch.equal(res[1][0], res[3][0])
    assert torch.equal(res[0][1], res[2][1]) and torch.equal(res[1][1], res[3][1])


def test_two_rank_sequence_parallel_prefill_matches_allreduce_path():
    """scripts/prefill_bench.py on 2 ranks: the sequence-parallel stack over peer memory (scatter-epilogue GEMMs, slot sums,
    gathered codes, fused residual add / norm / quantise, CUDA-graph replay) against the replicated-activation + NCCL
    all-reduce path on the SAME shards: only the reduction's rounding differs (fp32 slot sum vs bf16 ring)."""
    _need2()
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(_free_port()), os.path.join(ROOT, "scripts", "prefill_bench.py"), "--model", "tiny", "--bsz", "2",
           "--seq", "512", "--steps", "2", "--warmup", "1", "--check"]
    p = subprocess.Popen(cmd, cwd=ROOT, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, start_new_session=True)
    try:
        out, err = p.communicate(timeout=240)
    except subprocess.TimeoutExpired:
        import signal
        os.killpg(p.pid, signal.SIGKILL)
        out, err = p.communicate()
        pytest.fail("2-rank prefill timed out: " + err[-1500:])
    assert p.returncode == 0, err[-2000:]
    res = json.loads([l for l in out.splitlines() if l.startswith("{")][-1])
    assert "sequence parallel over peer memory" in res["config"]["parallelism"], res
    assert res["check"]["finite"] and res["check"]["rel_fro_vs_allreduce_path"] < 2e-2, res["check"]
