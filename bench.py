#!/usr/bin/env python
"""bench.py -- NVFP4+ARC hot path on B200: quantise (reorder + NVFP4 + ARC) + block-scaled GEMM.

A "step" is one pass of the hot path over one batch of M synthetic tokens through the seven linears of one decoder
layer (Llama-3-8B by default; --workload llama3-70b = BASELINE.json config 4), i.e. what the reference's layers ask of
`agemm` per layer (model/qLlamaLayer.py:296-303,364-367,429-443):

  default      4 quantise + 4 GEMM launches: q/k/v run as ONE launch over the concatenated weight with per-column scales
                and gate/up as ONE launch whose epilogue writes bf16(silu(gate)) * up (§8 f2);
  --no-swiglu  gate and up stay two plain launches (4 + 5);
  --unfused    the reference's launch structure, 4 quantise + 7 GEMM launches (A/B).

`value` = useful GEMM FLOPs (2*M*N*(KQ+KE), ARC columns included; identical for both structures) of the step / device
time of the whole step (quantise kernels included), inputs resident in HBM.  `e2e` = same metric through the public
`agemm` API with HOST (pinned) activations copied in and results copied out inside the timed region.  After the timed
regions (never inside them) `--check` (on by default) verifies the outputs of the last step against the oracle:
quantised bytes of sampled rows bit for bit (CPU oracle), every GEMM output on sampled rows against float64
dequantise-then-matmul, and under tensor parallelism the gathered codes and the reduce-scattered sums across ranks.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--tokens M] [--impl native|reference] [--unfused] [--no-check]
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

# name -> (hidden, intermediate, q_out, kv_out, KE of the hidden-size inputs, KE of the intermediate-size input)
WORKLOADS = {
    "llama3-8b": (4096, 14336, 4096, 1024, 256, 512),     # SURVEY.md App. B; BASELINE.json config 2
    "llama3-70b": (8192, 28672, 8192, 1024, 512, 1024),   # BASELINE.json config 4
}
ROW_PARALLEL = ("o", "down")  # K-split + reduce; the others are N-split


def layer_items(workload):
    """(name, KQ, KE, [N of each GEMM that shares this quantised activation])"""
    H, I, q, kv, ke_h, ke_i = WORKLOADS[workload]
    return [("qkv", H, ke_h, [q, kv, kv]), ("o", q, ke_h, [H]), ("gate_up", H, ke_h, [I, I]), ("down", I, ke_i, [H])]


def quant_bytes(M, KQ, KE):
    """Algorithmic bytes of one quantise launch: bf16 row in, codes + scales out, index once."""
    return M * (2 * KQ + (KQ + KE) // 2 + (KQ + KE) // 16) + 2 * KQ


def gemm_bytes(M, N, K, out_cols=None):
    return 0.5625 * K * (M + N) + 2.0 * M * (N if out_cols is None else out_cols)


def load_json(*parts):
    p = os.path.join(ROOT, *parts)
    if os.path.exists(p):
        try:
            with open(p) as f:
                return json.load(f)
        except Exception:
            return {}
    return {}


def load_peaks():
    d = load_json("MEASURED_PEAKS.json")
    if d:
        return {"hbm_gbs": d["hbm_gbs"], "bf16_tflops": d["bf16_tflops"],
                "bf16_tflops_sustained": d.get("bf16_tflops_sustained", d["bf16_tflops"]), "source": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx = gpu_index
        self.proc, self.lines, self.marks, self.thread = None, [], {}, None

    def start(self):
        """Started EARLY (before the layer is built): nvidia-smi needs ~0.1-0.3 s before its first line, and the driver's
        timed region can be as short as 20 x 0.9 ms.  Lines are stamped on arrival; stop() picks by time window."""
        import threading
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "20"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True, bufsize=1)
        except Exception:
            self.proc = None
            return

        def reader():
            try:
                for line in self.proc.stdout:
                    self.lines.append((time.perf_counter(), line))
            except Exception:
                pass
        self.thread = threading.Thread(target=reader, daemon=True)
        self.thread.start()

    def mark(self, name):
        self.marks[name] = time.perf_counter()

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.08)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        if self.thread is not None:
            self.thread.join(timeout=2)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        rows = []
        for t, line in list(self.lines):
            f = [c.strip() for c in line.split(",")]
            if len(f) < 9:
                continue
            try:
                rows.append((t, float(f[1]), float(f[2]), float(f[3]), [n for n, v in zip(names, f[5:9]) if v.lower().startswith("active")]))
            except ValueError:
                continue
        t0, t1 = self.marks.get("t0"), self.marks.get("t1")
        # a line describes the ~20 ms before it arrives; the driver's timed region can be as short as 20 x 0.9 ms, so the
        # window is the timed region plus one sampling period, and -- if nvidia-smi's pace put no line there -- the first
        # lines after it (the clocks the GPU left the timed loop with).  No extra load is run for the sampler's sake: 100
        # more steps in front of the timed ones were measured to move the step from burst (0.878 ms) to power-capped (0.998 ms)
        timed = [r for r in rows if t0 is not None and t1 is not None and t0 <= r[0] <= t1 + 0.025]
        use, window = timed, "timed region (+ one 20 ms sampling period)"
        if not use and t1 is not None:
            use, window = [r for r in rows if r[0] > t1][:3], "first samples after the timed region (none fell inside it)"
        if not use:
            use, window = rows[-3:], "last samples taken"
        reasons = sorted({n for r in use for n in r[4]})
        return {"sm_mhz": statistics.median([r[1] for r in use]) if use else None, "sm_max_mhz": max(r[2] for r in use) if use else None,
                "power_w_max": max(r[3] for r in use) if use else None, "samples": len(use), "samples_in_timed_region": len(timed),
                "window": window, "reasons": reasons}


# ---- the reference's CPU implementation of the path (fake-quant), bounded sample -----------------------------------

CPU_SAMPLE = ("fake-quant linear of BASELINE.json config 1: one 4096x4096 linear, M=%d, KE=256, fp32 on CPU, weight quantised once "
              "(fake_reorder_quantize_w outside the step, like the GPU arm's weights), step = fake_reorder_quantize_x + F.linear "
              "over K+KE (model/quantize.py path, restated in oracle/fakequant.py because /root/reference does not travel)")


def make_cpu_state(torch, M):
    from oracle import fakequant as fq
    from tests._util import make_acts
    KQ, KE, N = 4096, 256, 4096
    x, idx = make_acts(M, KQ, KE, seed=0)
    g = torch.Generator().manual_seed(1)
    w = (torch.randn(N, KQ, generator=g) * 0.02).to(torch.bfloat16).float()
    qw, _, sw = fq.fake_reorder_quantize_w(w, idx.long(), KE)   # once, outside the timed steps
    return (x.float(), qw, sw, idx.long(), KE), (KQ, KE, N)


def cpu_fakequant_step(torch, state):
    from oracle import fakequant as fq
    import torch.nn.functional as F
    x, qw, sw, idx, KE = state
    t0 = time.perf_counter()
    qx, _, _ = fq.fake_reorder_quantize_x(x, idx, KE)
    y = F.linear(qx, qw) * sw
    return time.perf_counter() - t0, y


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path on all host cores (rank 0 only)."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    import torch
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    M = args.cpu_tokens
    state, (KQ, KE, N) = make_cpu_state(torch, M)
    for _ in range(max(1, min(args.warmup, 2))):
        cpu_fakequant_step(torch, state)
    ts = [cpu_fakequant_step(torch, state)[0] for _ in range(args.steps)]
    total = sum(ts)
    val = 2.0 * M * N * (KQ + KE) * len(ts) / total / 1e12
    sample = CPU_SAMPLE % M
    print(json.dumps({
        "impl": "reference", "metric": "nvfp4_arc_linear_tflops", "value": val, "unit": "TFLOP/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / len(ts), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "nvfp4 (e2m1 x e2m1, ue4m3 block scales, fp32 accumulate)",
        "data": "synthetic",
        "config": {"workload": "llama3-8b decoder-layer linears, NVFP4+ARC quantize + GEMM", "sample": sample},
        "cpu_baseline": {"value": val, "unit": "TFLOP/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": "TFLOP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


# ---- the device-resident layer ----------------------------------------------------------------------------------------

class Layer:
    """Device-resident state of one decoder layer's linears: quantised weights + activation buffers.
    world > 1: this rank's tensor-parallel shard (arcquant_b200/tp.py): qkv and gate/up split N, o and down split K by
    channel ownership with their ARC columns.  Exchange steps:
      mode "tp"    (default) sequence-parallel edges: the row-parallel partials are REDUCE-SCATTERED over tokens, and the
                   column-parallel inputs are quantised on the rank's own M/world tokens and ALL-GATHERED as 4.5-bit
                   codes + scales (0.5625 B/element on the wire instead of 2 B of bf16).  With slabs > 1 the tokens are
                   cut into slabs and the collective of slab i runs on a communication stream under the GEMM of slab i+1;
      mode "tp-ar" plain all-reduce of the bf16 partials after o and down (activations replicated)."""

    def __init__(self, torch, agemm, M, dev, workload="llama3-8b", seed=0, world=1, rank=0, sp=True, fused=True, slabs=1, swiglu=False,
                 dist=None, peer=False):
        import numpy as np
        from tests._util import make_acts
        from arcquant_b200 import tp
        self.torch, self.agemm, self.M, self.world, self.rank, self.dev = torch, agemm, M, world, rank, dev
        self.sp = sp and world > 1
        self.fused = fused
        self.slabs = slabs if self.sp else 1
        if self.sp and M % (128 * world * self.slabs):
            raise SystemExit("bench.py: tokens must be a multiple of 128 * world * slabs for the sequence-parallel edges")
        self.items = []
        self.dist = dist
        # exchange over peer memory (arcquant_b200/peer.py, csrc/peer.cu): the row-parallel GEMM's epilogue scatters into the
        # owners' receive slots and the codes are gathered by peer stores -- no NCCL kernel on the data path
        self.peer, self.peer_note = None, None
        if peer and self.sp and self.slabs == 1 and dist is not None:
            self.peer, self.peer_note = self._peer_setup(dist, dev, world)
        g = torch.Generator().manual_seed(seed)
        for name, KQ, KE, Ns in layer_items(workload):
            xs, idx = make_acts(min(M, 256), KQ, KE, seed=seed + KQ)  # distribution A, pre-scaled
            row = name in ROW_PARALLEL and world > 1
            gather = self.sp and not row
            if row:
                lo, hi = tp.split_rows(KQ, world, rank, 64)
                idx, KE = tp.shard_reorder_index(idx, KE, world, rank)
                xs, KQ = xs[:, lo:hi].contiguous(), hi - lo
            elif world > 1:
                Ns = [n // world for n in Ns]
            Mx = M // world if gather else M          # rows this rank quantises
            My = M // world if (row and self.sp) else M  # rows of the result this rank keeps
            reps = (Mx + xs.shape[0] - 1) // xs.shape[0]
            x_host = xs.repeat(reps, 1)[:Mx].contiguous()
            # de-correlate repeated blocks cheaply: flip signs per row block
            sign = (torch.randint(0, 2, (Mx, 1), generator=g) * 2 - 1).to(torch.bfloat16)
            x_host = (x_host * sign).contiguous().pin_memory()
            idx_d = idx.to(dev)
            ws = []
            for N in Ns:
                w = (torch.randn(N, KQ, generator=g) * 0.02).to(torch.bfloat16)
                sw = w.max().float() / (448.0 * 6.0)
                qw, sfw = agemm.reorder_quantize_w((w / sw).to(torch.bfloat16).to(dev), idx_d, KE)
                ws.append((N, qw, sfw, float(sw)))
            K = KQ + KE
            # the activations are pre-scaled to +-2688 like the callers' x / scale_x (qLlamaLayer.py:73-77); alpha carries the
            # scale back: scale_x of a tensor whose largest magnitude was 8.0, so that the outputs have realistic O(1) size
            sx = float(np.float32(8.0) / np.float32(2688.0))
            it = {"name": name, "KQ": KQ, "KE": KE, "idx": idx_d, "x_host": x_host, "reduce": row, "gather": gather,
                  "x": x_host.to(dev), "w_sep": ws, "Ns": Ns}
            # launch list: (label, N, qw, sfw, alpha, col_scale, epilogue, out_cols)
            if swiglu and name == "gate_up" and Ns[0] % 128 == 0:
                qw, sfw, cs = agemm.interleave_gate_up(tuple(ws[0][1:]), tuple(ws[1][1:]))
                it["launches"] = [("gate_up+swiglu", 2 * Ns[0], qw, sfw, sx, cs, agemm.EPILOGUE_SWIGLU, Ns[0])]
            elif fused and name == "qkv" and all(n % 128 == 0 for n in Ns):
                qw, sfw, cs = agemm.fuse_weights([w[1:] for w in ws])
                it["launches"] = [("+".join(["q", "k", "v"][:len(ws)]), sum(Ns), qw, sfw, sx, cs, agemm.EPILOGUE_NONE, sum(Ns))]
            else:
                it["launches"] = [(name if len(ws) == 1 else "%s[%d]" % (name, j), N, qw, sfw, float(np.float32(sx) * np.float32(sw)), None,
                                   agemm.EPILOGUE_NONE, N) for j, (N, qw, sfw, sw) in enumerate(ws)]
            it["y"] = [torch.empty((M, l[7]), dtype=torch.bfloat16, device=dev) for l in it["launches"]]
            it["y_host"] = [torch.empty((My, l[7]), dtype=torch.bfloat16).pin_memory() for l in it["launches"]]
            if gather:
                it["qx_full"] = torch.empty((M, K // 2), dtype=torch.uint8, device=dev)
                it["sfx_full"] = torch.zeros((agemm.sf_bytes(M, K),), dtype=torch.uint8, device=dev)
                it["sf_valid"] = (M * K // 16, Mx * K // 16)
            if row and self.sp:
                it["y_rs"] = [torch.empty((My, l[7]), dtype=torch.bfloat16, device=dev) for l in it["launches"]]
            self.items.append(it)
        if self.peer is not None:
            self.peer, self.peer_note = self._peer_buffers(dist, dev)
        if self.slabs > 1:
            self.comm = torch.cuda.Stream()
        torch.cuda.synchronize()

    # -- peer-memory exchange: set-up (collective; every rank ends up on the same path) --
    def _agree(self, dist, dev, ok):
        flag = self.torch.tensor([1 if ok else 0], dtype=self.torch.int32, device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        return bool(int(flag.item()))

    def _peer_setup(self, dist, dev, world):
        from arcquant_b200 import peer as P
        grp, note = None, None
        try:
            if world > P.MAX_PEERS:
                raise RuntimeError("more than %d ranks" % P.MAX_PEERS)
            grp = P.PeerGroup(dist, dev).init_flags()
        except Exception as e:  # noqa: BLE001
            grp, note = None, "peer memory unavailable (%s): NCCL exchange" % str(e).splitlines()[0][:160]
        if not self._agree(dist, dev, grp is not None):
            return None, note or "peer memory unavailable on another rank: NCCL exchange"
        return grp, None

    def _peer_buffers(self, dist, dev):
        """Gathered code / scale buffers and receive slots as peer-mapped allocations (replacing the torch ones)."""
        from arcquant_b200 import peer as P
        torch, W = self.torch, self.world
        note = None
        try:
            for it in self.items:
                K = it["KQ"] + it["KE"]
                if it["gather"]:
                    it["qx_pb"] = P.PeerBuffer(self.peer, it["qx_full"].numel())
                    it["sfx_pb"] = P.PeerBuffer(self.peer, it["sfx_full"].numel())
                    it["qx_full"] = it["qx_pb"].view(torch.uint8, it["qx_full"].shape)
                    it["sfx_full"] = it["sfx_pb"].view(torch.uint8, it["sfx_full"].shape)
                if it["reduce"]:
                    My = self.M // W
                    it["recv"] = [P.PeerBuffer(self.peer, P.world_bytes(W, My, l[7])) for l in it["launches"]]
        except Exception as e:  # noqa: BLE001
            note = "peer buffers unavailable (%s): NCCL exchange" % str(e).splitlines()[0][:160]
        if not self._agree(dist, dev, note is None):
            return None, note or "peer buffers unavailable on another rank: NCCL exchange"
        return self.peer, None

    # -- one quantise launch + its GEMMs (+ the exchange steps of the tensor-parallel modes) --
    def _gemms(self, it, qx, sfx, rows=None, events=None):
        ag, torch = self.agemm, self.torch
        outs = []
        for j, (label, N, qw, sfw, alpha, cs, epi, out_cols) in enumerate(it["launches"]):
            y = it["y"][j] if rows is None else it["y"][j][rows[0]:rows[1]]
            if events is not None:
                e0 = torch.cuda.Event(enable_timing=True); e0.record()
            ag.matmul(qx, qw, sfx, sfw, alpha, col_scale=cs, epilogue=epi, out=y)
            if events is not None:
                e1 = torch.cuda.Event(enable_timing=True); e1.record()
                events.append(("gemm", label, it["KQ"], it["KE"], N, qx.shape[0], out_cols, e0, e1))
            outs.append(y)
        return outs

    def _item(self, it, x, events=None):
        ag, torch, dist = self.agemm, self.torch, getattr(self, "dist", None)
        if self.slabs > 1 and (it["gather"] or it["reduce"]):
            return self._item_slabs(it, x)
        if events is not None:
            e0 = torch.cuda.Event(enable_timing=True); e0.record()
        qx, sfx = ag.reorder_quantize_x(x, it["idx"], it["KE"])
        if events is not None:
            e1 = torch.cuda.Event(enable_timing=True); e1.record()
            events.append(("quant", it["name"], it["KQ"], it["KE"], 0, x.shape[0], 0, e0, e1))
        if it["gather"]:  # every rank quantised its own tokens: gather the codes and the 128-row scale blocks
            full, mine = it["sf_valid"]
            if self.peer is not None:   # one kernel: peer stores of both segments + barrier (csrc/peer.cu)
                if events is not None:
                    e0 = torch.cuda.Event(enable_timing=True); e0.record()
                self.peer.allgather(qx, it["qx_pb"], self.rank * qx.numel(), sfx[:mine], it["sfx_pb"], self.rank * mine, channel=1)
                if events is not None:
                    e1 = torch.cuda.Event(enable_timing=True); e1.record()
                    events.append(("xchg", "peer_allgather " + it["name"], it["KQ"], it["KE"], 0, x.shape[0], (qx.numel() + mine) * self.world, e0, e1))
            else:
                dist.all_gather_into_tensor(it["qx_full"], qx)
                dist.all_gather_into_tensor(it["sfx_full"][:full], sfx[:mine])
            qx, sfx = it["qx_full"], it["sfx_full"]
        it["last_q"] = (qx, sfx)
        if it["reduce"] and self.peer is not None:
            # row-parallel GEMM whose epilogue stores each 128-row block into its owner's receive slot over NVLink, then one
            # barrier + local fp32 sum of the W slots: the whole reduce-scatter, without a collective kernel
            outs, My = [], self.M // self.world
            for j, (label, N, qw, sfw, alpha, cs, epi, out_cols) in enumerate(it["launches"]):
                if events is not None:
                    e0 = torch.cuda.Event(enable_timing=True); e0.record()
                self.peer.gemm_scatter(qx, qw, sfx, sfw, alpha, it["recv"][j], My, col_scale=cs)
                if events is not None:
                    e1 = torch.cuda.Event(enable_timing=True); e1.record()
                    events.append(("gemm", label + "+scatter", it["KQ"], it["KE"], N, qx.shape[0], out_cols, e0, e1))
                self.peer.reduce(it["recv"][j], My, N, it["y_rs"][j], channel=0)
                if events is not None:
                    e2 = torch.cuda.Event(enable_timing=True); e2.record()
                    events.append(("xchg", "peer_reduce " + label, it["KQ"], it["KE"], N, My, (self.world + 1) * My * N * 2, e1, e2))
                outs.append(it["y_rs"][j])
            return outs
        outs = self._gemms(it, qx, sfx, None, events)
        if it["reduce"]:
            for j, y in enumerate(outs):
                if self.sp:
                    dist.reduce_scatter_tensor(it["y_rs"][j], y)
                    outs[j] = it["y_rs"][j]
                else:
                    dist.all_reduce(y)
        return outs

    def _item_slabs(self, it, x):
        """Tokens cut into S slabs (micro-batches) that flow through the layer independently: slab s of a linear depends only
        on slab s of the linear before it.  Collectives run on a communication stream; the compute stream meanwhile works on
        the other slabs, across linears (reduce chunk i under the GEMM of chunk i+1).  self.dep[s] = event
        after the last collective that produced slab s (the next linear's slab s waits for it, nothing else does)."""
        ag, torch, dist = self.agemm, self.torch, self.dist
        cur, comm, S, W = torch.cuda.current_stream(), self.comm, self.slabs, self.world
        K = it["KQ"] + it["KE"]
        capturing = torch.cuda.is_current_stream_capturing()
        if it["reduce"]:
            ms = self.M // S                      # rows of a slab; each rank keeps ms / W of them
            qs = []
            for s in range(S):
                if self.dep[s] is not None:
                    cur.wait_event(self.dep[s])
                qs.append(ag.reorder_quantize_x(x[s * ms:(s + 1) * ms], it["idx"], it["KE"]))
            it["last_q"] = qs
            for s in range(S):
                outs = self._gemms(it, qs[s][0], qs[s][1], (s * ms, (s + 1) * ms))
                ev = torch.cuda.Event(); ev.record(cur)
                with torch.cuda.stream(comm):
                    comm.wait_event(ev)
                    for j, y in enumerate(outs):
                        # rank r keeps rows [r ms / W, (r + 1) ms / W) of slab s: y_rs is laid out [slab][ms / W rows]
                        dist.reduce_scatter_tensor(it["y_rs"][j][s * ms // W:(s + 1) * ms // W], y)
                    self.dep[s] = torch.cuda.Event(); self.dep[s].record(comm)
            return it["y_rs"]
        # column-parallel: quantise slab s of the rank's own tokens, gather it on the communication stream, GEMM when it lands
        mx = x.shape[0] // S                      # own rows per slab
        mg = mx * W                               # gathered rows per slab
        evs = []
        for s in range(S):
            if self.dep[s] is not None:
                cur.wait_event(self.dep[s])
            qx, sfx = ag.reorder_quantize_x(x[s * mx:(s + 1) * mx], it["idx"], it["KE"])
            ev = torch.cuda.Event(); ev.record(cur)
            with torch.cuda.stream(comm):
                comm.wait_event(ev)
                dist.all_gather_into_tensor(it["qx_full"][s * mg:(s + 1) * mg], qx)
                dist.all_gather_into_tensor(it["sfx_full"][s * mg * K // 16:(s + 1) * mg * K // 16], sfx[:mx * K // 16])
                e2 = torch.cuda.Event(); e2.record(comm)
            if not capturing:
                qx.record_stream(comm); sfx.record_stream(comm)
            evs.append(e2)
        it["last_q"] = (it["qx_full"], it["sfx_full"])
        for s in range(S):
            cur.wait_event(evs[s])
            self._gemms(it, it["qx_full"][s * mg:(s + 1) * mg], it["sfx_full"][s * mg * K // 16:], (s * mg, (s + 1) * mg))
            self.dep[s] = None                    # the next linear's slab s follows on the compute stream itself
        return it["y"]

    def step(self, events=None):
        self.dep = [None] * max(1, self.slabs)
        for it in self.items:
            it["outs"] = self._item(it, it["x"], events)
        if self.slabs > 1:
            self.torch.cuda.current_stream().wait_stream(self.comm)

    def step_e2e(self):
        """Same step through the public API with HOST buffers: H2D of every activation (pinned -> device) and D2H of
        every result inside the step.  Copies run on their own streams and are ordered against the kernels with events,
        so PCIe traffic of one linear overlaps the kernels of another; nothing is skipped or cached."""
        torch = self.torch
        cur = torch.cuda.current_stream()
        if not hasattr(self, "_h2d"):
            self._h2d, self._d2h = torch.cuda.Stream(), torch.cuda.Stream()
            for it in self.items:
                it["x_in"] = torch.empty_like(it["x"])
                it["ev_in"], it["ev_read"], it["ev_y"], it["ev_out"] = (torch.cuda.Event() for _ in range(4))
                it["ev_read"].record(cur)
                it["ev_out"].record(self._d2h)
        with torch.cuda.stream(self._h2d):
            for it in self.items:
                self._h2d.wait_event(it["ev_read"])  # the previous step's quantise has consumed x_in
                it["x_in"].copy_(it["x_host"], non_blocking=True)
                it["ev_in"].record(self._h2d)
        self.dep = [None] * max(1, self.slabs)
        for it in self.items:
            cur.wait_event(it["ev_in"])
            cur.wait_event(it["ev_out"])  # the previous step's D2H of this item's result buffers is done
            outs = self._item(it, it["x_in"])
            if self.slabs > 1:
                cur.wait_stream(self.comm)  # the reduce-scattered result is produced on the communication stream
            it["ev_read"].record(cur)
            it["ev_y"].record(cur)
            with torch.cuda.stream(self._d2h):
                self._d2h.wait_event(it["ev_y"])
                for y, yh in zip(outs, it["y_host"]):
                    yh.copy_(y, non_blocking=True)
                it["ev_out"].record(self._d2h)

    def e2e_join(self):
        cur = self.torch.cuda.current_stream()
        if hasattr(self, "_d2h"):
            cur.wait_stream(self._d2h)
            cur.wait_stream(self._h2d)

    def flops(self):
        return sum(2.0 * self.M * N * (it["KQ"] + it["KE"]) for it in self.items for N in it["Ns"])

    def exchange_bytes(self):
        """Bytes this rank hands to NCCL per step (collective inputs)."""
        b = sum(y.numel() * 2 for it in self.items if it["reduce"] for y in it["y"])
        b += sum(it["qx_full"].numel() + it["sf_valid"][0] for it in self.items if it["gather"]) // max(self.world, 1)
        return b

    def e2e_bytes(self):
        h2d = sum(it["x_host"].numel() * 2 for it in self.items)
        d2h = sum(y.numel() * 2 for it in self.items for y in it["y_host"])
        return h2d, d2h

    # -- parity of the last step's outputs (outside every timed region) --
    def check(self):
        """Oracle check of what the last step() left in the buffers.  quantise: bytes of sampled rows against the CPU
        oracle; GEMM: sampled rows of every output against float64 dequantise-then-matmul (tolerance of
        tests/test_gpu_gemm.py); SwiGLU output: relative error of silu(g) * u recomputed in float64; tensor-parallel:
        gathered codes against every rank's own, reduce-scattered sums against the float64 sum of the partials."""
        import numpy as np
        import oracle
        from oracle import torchref as T
        from tests._util import bits
        torch, dist, W = self.torch, getattr(self, "dist", None), self.world
        O = oracle.lib()
        res = {"gemm_max_err_over_tol": 0.0, "gemm_rel_fro_max": 0.0, "swiglu_rel_err_max": 0.0, "quant_bytes_differ": 0,
               "quant_rows_checked": 0, "gemm_outputs_checked": 0, "tp_gather_bytes_differ": 0, "tp_reduce_max_err_over_tol": 0.0,
               "ranks": W}
        torch.cuda.synchronize()
        for it in self.items:
            K = it["KQ"] + it["KE"]
            if isinstance(it["last_q"], list):   # row-parallel linear quantised slab by slab: the slabs are contiguous row ranges
                qx = torch.cat([q for q, _ in it["last_q"]])
                nsl = qx.shape[0] // len(it["last_q"]) * K // 16
                sfx = torch.cat([sf[:nsl] for _, sf in it["last_q"]] + [torch.zeros(128 * K // 16, dtype=torch.uint8, device=self.dev)])
            else:
                qx, sfx = it["last_q"]
            # (1) this rank's own quantise output: rebuild it from x and compare sampled rows with the CPU oracle
            q_own, sf_own = self.agemm.reorder_quantize_x(it["x"], it["idx"], it["KE"])
            rows = T.sample_rows(it["x"].shape[0], per_block=1, seed=it["KQ"])[:48]
            xr = it["x"][rows.to(self.dev)].cpu()
            q_o, sf_o = O.quantize_x(bits(xr), it["idx"].cpu().numpy(), it["KE"])
            res["quant_bytes_differ"] += int((q_own[rows.to(self.dev)].cpu().numpy() != q_o).sum())
            sf_rows = sf_own[T.sf_offsets(rows.to(self.dev), K)].cpu().numpy()
            res["quant_bytes_differ"] += int((sf_rows != sf_o[O.sf_valid_offsets(rows.numel(), K)]).sum())
            res["quant_rows_checked"] += int(rows.numel())
            if it["gather"]:
                # (2) the gathered codes / scale blocks: rank r's slice must be rank r's own output, byte for byte
                mx, S = it["x"].shape[0], self.slabs
                own = torch.stack([q_own[s * (mx // S):(s + 1) * (mx // S)] for s in range(S)])         # [S, mx/S, K/2]
                allq = [torch.empty_like(own) for _ in range(W)]
                dist.all_gather(allq, own)
                want = torch.cat([torch.cat([allq[r][s] for r in range(W)]) for s in range(S)])
                res["tp_gather_bytes_differ"] += int((want != qx).sum())
                nsf = mx * K // 16
                allsf = [torch.empty(nsf, dtype=torch.uint8, device=self.dev) for _ in range(W)]
                dist.all_gather(allsf, sf_own[:nsf].contiguous())
                per = nsf // S
                want_sf = torch.cat([torch.cat([allsf[r][s * per:(s + 1) * per] for r in range(W)]) for s in range(S)])
                res["tp_gather_bytes_differ"] += int((want_sf != sfx[:want_sf.numel()]).sum())
            # (3) every GEMM output of the step on sampled rows
            grows = T.sample_rows(self.M, per_block=1, seed=it["KQ"] + 1)
            if grows.numel() > 80:
                grows = grows[torch.linspace(0, grows.numel() - 1, 80).long()]
            grows = grows.to(self.dev)
            for j, (label, N, qw, sfw, alpha, cs, epi, out_cols) in enumerate(it["launches"]):
                ref, mag = T.gemm_rows(qx, sfx, qw, sfw, grows, alpha, cs)
                y = it["y"][j]
                peer_rs = it["reduce"] and self.peer is not None   # the partial never existed locally: it sits in the owners' slots
                if epi == self.agemm.EPILOGUE_SWIGLU:
                    nb = N // 256
                    r3 = ref.view(-1, nb, 2, 128)
                    want = (torch.nn.functional.silu(r3[:, :, 0]) * r3[:, :, 1]).reshape(-1, N // 2)
                    got = y[grows].double()
                    res["swiglu_rel_err_max"] = max(res["swiglu_rel_err_max"], float((got - want).norm() / want.norm().clamp_min(1e-30)))
                else:
                    tol = 2.0 ** -8 * ref.abs() + 2.0 ** -18 * mag + 1e-30
                    if peer_rs:   # this rank's partial of the rows it owns = its own slot of its own receive buffer
                        mine = self._rs_rows(grows)
                        if mine is not None:
                            sel, local = mine
                            My = self.M // W
                            slot = it["recv"][j].view(torch.bfloat16, (W, My, N))[self.rank]
                            got_p = slot[local].double()
                            res["gemm_max_err_over_tol"] = max(res["gemm_max_err_over_tol"], float(((got_p - ref[sel]).abs() / tol[sel]).max()))
                            res["gemm_rel_fro_max"] = max(res["gemm_rel_fro_max"], float((got_p - ref[sel]).norm() / ref[sel].norm().clamp_min(1e-30)))
                    else:
                        got = y[grows].double()
                    in_place_sum = it["reduce"] and not self.sp   # tp-ar: y already holds the all-reduced sum, not the partial
                    if not in_place_sum and not peer_rs:
                        res["gemm_max_err_over_tol"] = max(res["gemm_max_err_over_tol"], float(((got - ref).abs() / tol).max()))
                        res["gemm_rel_fro_max"] = max(res["gemm_rel_fro_max"], float((got - ref).norm() / ref.norm().clamp_min(1e-30)))
                    if it["reduce"]:
                        # (4) reduced result = sum over ranks of the partials: float64 sum of every rank's reference
                        tot, tol_sum = ref.clone(), tol.clone() + 2.0 ** -8 * ref.abs()
                        dist.all_reduce(tot); dist.all_reduce(tol_sum)
                        if self.sp:
                            mine = self._rs_rows(grows)
                            if mine is not None:
                                sel, local = mine
                                got_r = it["y_rs"][j][local].double()
                                e = (got_r - tot[sel]).abs() / (tol_sum[sel] + 2.0 ** -7 * tot[sel].abs())
                                res["tp_reduce_max_err_over_tol"] = max(res["tp_reduce_max_err_over_tol"], float(e.max()))
                        else:
                            e = (got - tot).abs() / (tol_sum + 2.0 ** -7 * tot.abs())
                            res["tp_reduce_max_err_over_tol"] = max(res["tp_reduce_max_err_over_tol"], float(e.max()))
                res["gemm_outputs_checked"] += 1
        if W > 1:
            t = torch.tensor([res["gemm_max_err_over_tol"], res["gemm_rel_fro_max"], res["swiglu_rel_err_max"],
                              res["tp_reduce_max_err_over_tol"]], dtype=torch.float64, device=self.dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            c = torch.tensor([res["quant_bytes_differ"], res["tp_gather_bytes_differ"]], dtype=torch.int64, device=self.dev)
            dist.all_reduce(c)
            res["gemm_max_err_over_tol"], res["gemm_rel_fro_max"], res["swiglu_rel_err_max"], res["tp_reduce_max_err_over_tol"] = t.tolist()
            res["quant_bytes_differ"], res["tp_gather_bytes_differ"] = (int(v) for v in c.tolist())
        if self.peer is not None:
            try:
                self.peer.status()
                res["peer_barrier_timeouts"] = 0
            except RuntimeError:
                res["peer_barrier_timeouts"] = 1
        res["ok"] = bool(res.get("peer_barrier_timeouts", 0) == 0 and res["quant_bytes_differ"] == 0 and res["tp_gather_bytes_differ"] == 0 and res["gemm_max_err_over_tol"] <= 1.0
                         and res["gemm_rel_fro_max"] < 2e-3 and res["swiglu_rel_err_max"] < 1e-2 and res["tp_reduce_max_err_over_tol"] <= 1.0)
        return res

    def _rs_rows(self, grows):
        """Which of the sampled global rows this rank keeps after the reduce-scatter, and where (y_rs is [slab][ms/W rows])."""
        S, W, r = self.slabs, self.world, self.rank
        ms = self.M // S
        slab, off = grows // ms, grows % ms
        keep = (off // (ms // W)) == r
        if not bool(keep.any()):
            return None
        local = slab[keep] * (ms // W) + off[keep] % (ms // W)
        return keep.nonzero().flatten(), local


def run_native(args):
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
            os.environ["NCCL_DEBUG"] = "WARN"  # keep stdout to the one JSON line (NCCL prints its version there)
        dist.init_process_group("nccl", device_id=dev)
    from arcquant_b200 import agemm
    for kv in args.option:
        agemm.set_option(*kv.split("=", 1))
    peaks = load_peaks()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    M = args.tokens
    tp_mode = world > 1 and args.parallelism in ("tp", "tp-ar")
    slabs = args.slabs if (tp_mode and args.parallelism == "tp") else 1
    layer = Layer(torch, agemm, M, dev, workload=args.workload, world=world if tp_mode else 1, rank=rank if tp_mode else 0,
                  sp=args.parallelism == "tp", fused=not args.unfused, slabs=slabs, swiglu=args.swiglu and not args.unfused,
                  dist=dist if tp_mode else None, peer=args.peer and tp_mode and args.parallelism == "tp" and slabs == 1)
    layer.dist = dist

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    if layer.peer is not None:
        # first step on the peer-memory path: if any barrier timed out (a rank that cannot see its peers' stores), every rank
        # drops to the NCCL exchange together instead of timing a broken step
        layer.step()
        torch.cuda.synchronize()
        try:
            layer.peer.status()
            ok = True
        except RuntimeError as e:
            ok, layer.peer_note = False, "peer barrier timed out in the first step (%s): NCCL exchange" % str(e)[:120]
        if not layer._agree(dist, dev, ok):
            layer.peer, layer.peer_note = None, layer.peer_note or "peer barrier timed out on another rank: NCCL exchange"
    for _ in range(max(args.warmup, 3)):
        layer.step()
    barrier()
    n0 = agemm.launch_count()
    layer.step()
    launches_per_step = agemm.launch_count() - n0   # this library's kernels in one step (a graph replay re-issues exactly these)
    barrier()
    # tensor-parallel shards are launch-bound (a few hundred microseconds of kernels against ~20 host calls): capture the
    # step (kernels + NCCL collectives, both streams) in a CUDA graph and replay it
    graph = None
    # default with the peer-memory exchange: the captured step then holds only this library's kernels (no NCCL node), and at
    # N >= 4 a shard's kernels are shorter than the ~20 host calls that launch them
    if tp_mode and (args.graph or layer.peer is not None) and not args.no_graph:
        ok = 1
        try:
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                layer.step()
                torch.cuda.synchronize()
                graph = torch.cuda.CUDAGraph()
                # thread_local: NCCL's watchdog thread polls events while we capture; that must not invalidate the capture
                with torch.cuda.graph(graph, stream=side, capture_error_mode="thread_local"):
                    layer.step()
            torch.cuda.current_stream().wait_stream(side)
        except Exception as e:  # noqa: BLE001
            sys.stderr.write("bench.py[rank %d]: CUDA-graph capture of the tensor-parallel step failed (%s)\n" % (rank, str(e).splitlines()[0][:200]))
            ok, graph = 0, None
        # every rank must take the same path: one eager rank against replaying ranks would deadlock in the first collective
        torch.cuda.synchronize()
        flag = torch.tensor([ok], dtype=torch.int32, device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if int(flag.item()) == 0:
            graph = None
        else:
            for _ in range(3):
                graph.replay()
        barrier()

    def run_step():
        if graph is not None:
            graph.replay()
        else:
            layer.step()

    n0 = agemm.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    sampler.mark("t0")
    e0.record()
    for _ in range(args.steps):
        run_step()
    e1.record()
    barrier()
    sampler.mark("t1")
    ms = e0.elapsed_time(e1)
    launches = agemm.launch_count() - n0
    clocks = sampler.stop() if rank == 0 else None
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max = float(t.item())
    ft = torch.tensor([layer.flops()], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(ft)  # whole-job FLOPs of one step: sum of the ranks' shards (TP) or replicas
    flops_job = float(ft.item())
    value = flops_job * args.steps / (ms_max * 1e-3) / 1e12

    # ---- e2e: host activations in, results out, inside the timed region ----
    for _ in range(2):
        layer.step_e2e()
    layer.e2e_join()
    barrier()
    e0.record()
    for _ in range(args.e2e_steps):
        layer.step_e2e()
    layer.e2e_join()
    e1.record()
    barrier()
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_val = flops_job * args.e2e_steps / (float(t.item()) * 1e-3) / 1e12
    h2d, d2h = layer.e2e_bytes()

    # ---- per-kernel averages (instrumented pass: CUDA events around every launch, same stream; slabs off) ----
    events = []
    saved_slabs, layer.slabs = layer.slabs, 1
    for _ in range(args.steps):
        layer.step(events)
    barrier()
    layer.slabs = saved_slabs

    # ---- parity of the step's outputs (outside the timed regions) ----
    parity = None
    if not args.no_check:
        layer.step()
        barrier()
        parity = layer.check()

    if rank != 0:
        _leave(world)
        return
    per, meta = {}, {}
    for kind, label, KQ, KE, N, rows, out_cols, a, b in events:
        key = (kind, label, KQ, KE, N)
        per.setdefault(key, []).append(a.elapsed_time(b))
        meta[key] = (rows, out_cols)
    kernels = []
    total_ms = 0.0
    refk = {(r["M"], r["KQ"], r["KE"]): r for r in load_json("profiles", "r2_nvfp4_peak.json").get("quant", [])}
    for key, v in sorted(per.items()):
        kind, label, KQ, KE, N = key
        rows, out_cols = meta[key]
        avg = sum(v) / len(v)
        calls_per_step = len(v) / args.steps
        total_ms += avg * calls_per_step
        if kind == "quant":
            by = quant_bytes(rows, KQ, KE)
            k = {"kernel": "quant", "input_of": label, "KQ": KQ, "KE": KE, "rows": rows, "us": 1e3 * avg, "calls_per_step": calls_per_step,
                 "GBps": by / (avg * 1e-3) / 1e9, "frac_hbm": by / (avg * 1e-3) / 1e9 / peaks["hbm_gbs"]}
            r = refk.get((rows, KQ, KE))
            if r and "ref_kernel_us_best" in r:  # the reference's own kernel on this shape, measured by scripts/measure_nvfp4_peak.py
                k["ref_kernel_us"] = r["ref_kernel_us_best"]
                k["speedup_vs_ref_kernel"] = r["ref_kernel_us_best"] / (1e3 * avg)
            kernels.append(k)
        elif kind == "xchg":   # barrier wait (rank skew) included; bytes = what the kernel reads + writes
            kernels.append({"kernel": label.split()[0], "of": label.split()[1], "rows": rows, "us": 1e3 * avg, "calls_per_step": calls_per_step,
                            "bytes": out_cols, "GBps": out_cols / (avg * 1e-3) / 1e9})
        else:
            fl = 2.0 * rows * N * (KQ + KE)
            kernels.append({"kernel": "gemm", "launch": label, "N": N, "K": KQ + KE, "rows": rows, "us": 1e3 * avg,
                            "calls_per_step": calls_per_step, "TFLOPs": fl / (avg * 1e-3) / 1e12})
    ncu = load_json("profiles", "ncu_summary.json") if (M == 8192 and world == 1) else {}
    gemms = [k for k in kernels if k["kernel"] == "gemm"]
    dom = max(gemms, key=lambda k: k["us"] * k["calls_per_step"])
    # denominator: the measured cuBLASLt NVFP4 figure on this pool (profiles/r2_nvfp4_peak.json, scripts/measure_nvfp4_peak.py)
    # when it exists, else 4 x the measured cuBLAS bf16 figure; BURST basis: the timed region is milliseconds long
    pk = load_json("profiles", "r2_nvfp4_peak.json")
    if pk.get("nvfp4_peak_tflops_burst"):
        nvfp4_peak, basis = pk["nvfp4_peak_tflops_burst"], ("measured cuBLASLt NVFP4 (torch._scaled_mm) burst, best shape, "
                                                           "profiles/r2_nvfp4_peak.json; sustained %.0f" % pk.get("nvfp4_peak_tflops_sustained", 0))
    else:
        nvfp4_peak, basis = 4.0 * peaks["bf16_tflops"], "4 x %s cuBLAS bf16 burst (%.1f TF/s): no measured NVFP4 peak on file" % (
            peaks["source"], peaks["bf16_tflops"])
    dk = "gemm_N%d_K%d" % (dom["N"], dom["K"])
    roofline = {"bound": "tensor", "kernel": "gemm_nvfp4 %s N=%d K=%d M=%d" % (dom["launch"], dom["N"], dom["K"], dom["rows"]),
                "achieved": dom["TFLOPs"], "peak": nvfp4_peak, "unit": "TFLOP/s", "frac": dom["TFLOPs"] / nvfp4_peak,
                "peak_basis": basis, "frac_of_4x_bf16_burst": dom["TFLOPs"] / (4.0 * peaks["bf16_tflops"]),
                "frac_of_nominal_9000": dom["TFLOPs"] / 9000.0,
                "traffic": ncu.get(dk, {}).get("dram_bytes"),
                "ncu_tensor_pipe_active_pct": ncu.get(dk, {}).get("tensor_pipe_active_pct"),
                "algorithmic_bytes": gemm_bytes(dom["rows"], dom["N"], dom["K"]),
                "share_of_step": dom["us"] * dom["calls_per_step"] / (1e3 * total_ms)}
    qk = max([k for k in kernels if k["kernel"] == "quant"], key=lambda k: k["us"])
    roofline_quant = {"bound": "hbm", "kernel": "quant_tma_kernel KQ=%d KE=%d M=%d" % (qk["KQ"], qk["KE"], qk["rows"]),
                      "achieved": qk["GBps"], "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": qk["frac_hbm"],
                      "peak_basis": "%s copy bandwidth" % peaks["source"],
                      "traffic": ncu.get("quant_KQ%d" % qk["KQ"], {}).get("dram_bytes"),
                      "algorithmic_bytes": quant_bytes(qk["rows"], qk["KQ"], qk["KE"])}

    # ---- CPU baseline on this box's host cores (bounded sample) ----
    cpu = None
    if not args.no_cpu_baseline:
        cores = os.cpu_count() or 1
        torch.set_num_threads(cores)
        Mc = args.cpu_tokens
        state, (KQ, KE, N) = make_cpu_state(torch, Mc)
        cpu_fakequant_step(torch, state)
        ts = [cpu_fakequant_step(torch, state)[0] for _ in range(args.cpu_reps)]
        cpu = {"value": 2.0 * Mc * N * (KQ + KE) * len(ts) / sum(ts) / 1e12, "unit": "TFLOP/s", "cores": cores, "kind": "port",
               "sample": "%d x %s; %.1f s of CPU work" % (len(ts), CPU_SAMPLE % Mc, sum(ts))}

    # ---- BASELINE.json config 3 beside it (driver-run): Qwen2.5-7B speed-model prefill, bs 4 x seq 1024, one GPU ----
    prefill_res = None
    if world == 1 and not args.no_prefill:
        try:
            prefill_res = run_prefill(torch, agemm, dev)
        except Exception as e:  # noqa: BLE001
            prefill_res = {"error": str(e).splitlines()[0][:200]}

    layer_swiglu = args.swiglu and not args.unfused
    structure = "%d x reorder+NVFP4+ARC quantize + %d x block-scaled GEMM per step" % (
        len(layer.items), sum(len(it["launches"]) for it in layer.items))
    out = {
        "metric": "nvfp4_arc_linear_tflops", "value": value, "unit": "TFLOP/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms_max / args.steps, "higher_is_better": True,
        "scaling": "weak" if (world > 1 and not tp_mode) else "strong",
        "vs_baseline": None, "dtype": "nvfp4 (e2m1 x e2m1, ue4m3 block scales, fp32 accumulate, bf16 out)",
        "data": "synthetic",
        "config": {"workload": "%s decoder-layer linears (qkv, o, gate/up, down): %s%s" % (
                       args.workload, structure, "" if args.unfused else (" (q/k/v fused, gate/up fused with SwiGLU epilogue)" if layer_swiglu else " (q/k/v fused)")),
                   "tokens_global": M if (tp_mode or world == 1) else M * world,
                   "tokens_per_gpu": (M // world if (tp_mode and layer.sp) else M),
                   "select_num": {n: ke for n, _, ke, _ in layer_items(args.workload)},
                   "l2": "inputs larger than L2 (per-step working set > 1 GB vs 126 MB L2); no flush needed",
                   "scaling_note": "strong: the same %d tokens at every N under tensor parallelism" % M if tp_mode or world == 1
                                   else "weak: %d tokens per GPU (replicas)" % M,
                   "parallelism": ("tp%d: qkv/gate/up column-parallel, o/down row-parallel (ARC columns with their shard); %s; "
                                   "%d slab(s), collectives on a communication stream under the next slab's kernels; step %s; "
                                   "%d bytes per rank per step %s" % (
                                       world, "reduce-scatter of bf16 partials + all-gather of NVFP4 codes/scales (sequence-parallel edges)"
                                       if layer.sp else "all-reduce of bf16 partials", layer.slabs,
                                       "replayed from a CUDA graph" if graph is not None else "launched from Python",
                                       layer.exchange_bytes(),
                                       "over PEER MEMORY (NVLink P2P stores from the GEMM epilogue / gather kernel + flag barriers; no NCCL kernel "
                                       "on the data path)" if layer.peer is not None else "into NCCL")) if tp_mode
                   else ("replicas (token-parallel), no collective" if world > 1 else "single GPU")},
        "tokens_per_s_linear_path": (M if (tp_mode or world == 1) else world * M) * args.steps / (ms_max * 1e-3),
        "e2e": {"value": e2e_val, "unit": "TFLOP/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "steps": args.e2e_steps},
        "gpu_launches": int(launches) if graph is None else int(args.steps * launches_per_step),
        "roofline": roofline, "roofline_quant": roofline_quant, "kernels": kernels,
        "parity_check": parity, "cpu_baseline": cpu, "clocks": clocks, "prefill_config3": prefill_res,
        "binding": agemm.BACKEND,
    }
    if args.option:
        out["options"] = args.option
    if tp_mode:
        out["exchange"] = "peer" if layer.peer is not None else "nccl"
        if layer.peer_note:
            out["exchange_note"] = layer.peer_note
    print(json.dumps(out))
    _leave(world)


def _leave(world):
    """Multi-rank runs end here without tearing NCCL down: every collective of the run is complete (the ranks have
    synchronised after the last one), and destroying a communicator that CUDA graphs or a second stream still reference
    was seen to hang on this pool (torchrun then waits for ever).  Flush and exit the process with status 0."""
    sys.stdout.flush()
    sys.stderr.flush()
    if world > 1:
        os._exit(0)


def run_prefill(torch, agemm, dev, name="qwen2.5-7b", bsz=4, seq=1024, steps=3):
    """BASELINE.json config 3 (benchmarks/benchmark_e2e_arc.py shape): random-init Qwen2.5-7B speed-model stack."""
    from arcquant_b200 import prefill
    cfg = prefill.PrefillConfig.named(name)
    model = prefill.ArcPrefillModel(cfg, dev)
    h = torch.randn(bsz, seq, cfg.hidden, device=dev, generator=torch.Generator(device=dev).manual_seed(1)).to(torch.bfloat16)
    for _ in range(2):
        model(h)
    torch.cuda.synchronize()
    graph = None
    try:   # the whole forward (this library's kernels + torch glue) replayed from one CUDA graph: no host launch time in the number
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=side):
                model(h)
        torch.cuda.current_stream().wait_stream(side)
        graph.replay()
        torch.cuda.synchronize()
    except Exception:  # noqa: BLE001 -- fall back to launching from Python
        graph = None
        torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        if graph is not None:
            graph.replay()
        else:
            model(h)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    launch = "CUDA-graph replay" if graph is not None else "from Python"
    del graph
    del model
    torch.cuda.empty_cache()
    return {"metric": "prefill_tokens_per_s", "value": bsz * seq / (ms * 1e-3), "unit": "tokens/s", "ms_per_forward": ms,
            "config": {"workload": "%s speed-model prefill, %d layers, bs %d x seq %d, random-init, torch SDPA attention" % (
                name, cfg.layers, bsz, seq), "launch": launch}}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--tokens", type=int, default=8192, help="M: rows (tokens) of the step (global under tensor parallelism)")
    ap.add_argument("--workload", default="llama3-8b", choices=sorted(WORKLOADS))
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--parallelism", default="tp", choices=["tp", "tp-ar", "replicas"],
                    help="N>1: tensor-parallel shards of one layer with reduce-scatter + quantised all-gather (default), "
                         "with plain all-reduce (tp-ar), or independent token replicas")
    ap.add_argument("--slabs", type=int, default=1,
                    help="tp: token slabs (micro-batches) whose collectives run on a communication stream under the other slabs' "
                         "kernels; measured SLOWER than 1 at N=2 (0.97 vs 0.83 ms: split GEMMs + NCCL kernels taking SMs from the "
                         "persistent GEMM grid), so off by default")
    ap.add_argument("--graph", action="store_true", help="tp: capture the step in a CUDA graph and replay it (default with the peer-memory "
                                                         "exchange; with --no-peer the graph also holds the NCCL collectives)")
    ap.add_argument("--unfused", action="store_true", help="7 GEMM launches per step (the reference's launch structure)")
    ap.add_argument("--no-swiglu", dest="swiglu", action="store_false",
                    help="keep gate and up as two plain launches (default: ONE launch whose epilogue writes bf16(silu(gate)) * up, "
                         "which also replaces the callers' torch SiLU*up pass and halves the bytes written)")
    ap.set_defaults(swiglu=True)
    ap.add_argument("--no-graph", action="store_true", help="tp: launch every step from Python")
    ap.add_argument("--no-peer", dest="peer", action="store_false",
                    help="tp: exchange through NCCL (reduce_scatter_tensor + all_gather_into_tensor) instead of the peer-memory path "
                         "(GEMM epilogue scatters into the owners' receive slots over NVLink + barrier/fp32-sum kernel; codes gathered by "
                         "peer stores), which is the default when the ranks can map each other's memory")
    ap.set_defaults(peer=True)
    ap.add_argument("--no-check", action="store_true", help="skip the oracle check of the last step's outputs")
    ap.add_argument("--no-prefill", action="store_true", help="skip the config-3 prefill line (N=1)")
    ap.add_argument("--option", action="append", default=[], metavar="NAME=VALUE",
                    help="arcq_set_option override for A/B runs (every choice computes the same result), e.g. gemm_raster=8")
    ap.add_argument("--e2e-steps", type=int, default=5)
    ap.add_argument("--cpu-tokens", type=int, default=2048)
    ap.add_argument("--cpu-reps", type=int, default=5)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_native(args)


if __name__ == "__main__":
    main()
