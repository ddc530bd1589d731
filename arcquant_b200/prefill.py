"""Prefill harness: a Llama/Qwen-shaped decoder stack whose linears run on the NVFP4+ARC hot path.

Mirrors how the reference's speed model drives ``agemm`` (benchmarks/modeling_arc.py: ``LlamaRMSNorm.forward`` :220-227
-> ``rmsnorm_quantize_x``; ``QLlamaAttention.forward`` :156-209 -> q/k/v ``matmul`` on the shared quantised
activation, attention, ``reorder_quantize_x`` + o_proj; ``QLlamaMLP.forward`` :108-110 -> gate/up ``matmul``,
SiLU*up, ``reorder_quantize_x`` + down_proj; ``LlamaDecoderLayer.forward`` :268-300 residual adds).  Differences:
random-init weights (no checkpoints here), attention through torch SDPA instead of flashinfer (attention is not on
the hot path), per-tensor scales kept on the device (no host sync), and optional tensor parallelism
(arcquant_b200/tp.py: q/k/v/gate/up column-parallel, o/down row-parallel + all-reduce).

Only the linears and the quantisers are this repo's kernels; everything else is plain torch glue.
"""
from __future__ import annotations

from dataclasses import dataclass

import torch
import torch.distributed as dist
import torch.nn.functional as F

from . import tp

MODELS = {
    # name: hidden, intermediate, heads, layers, qkv bias  (speed-model shapes: MHA, benchmarks/modeling_arc.py:139-154)
    "qwen2.5-7b": (3584, 18944, 28, 28, True),
    "llama-3-8b": (4096, 14336, 32, 32, False),
    "qwen2.5-32b": (5120, 27648, 40, 64, True),
    "llama-3-70b": (8192, 28672, 64, 80, False),
    "tiny": (2048, 4096, 16, 2, True),
}


@dataclass
class PrefillConfig:
    hidden: int
    intermediate: int
    heads: int
    layers: int
    qkv_bias: bool
    ke_hidden: int = 256       # select_num of the hidden-size inputs (multiple of 64, utilize.py:475)
    ke_intermediate: int = 512
    eps: float = 1e-6

    @staticmethod
    def named(name: str, layers: int | None = None) -> "PrefillConfig":
        h, i, nh, nl, b = MODELS[name]
        return PrefillConfig(h, i, nh, layers if layers is not None else nl, b)


def _rand_weight(n, k, dev, gen):
    return (torch.randn(n, k, device=dev, generator=gen) * 0.02).to(torch.bfloat16)


def _rand_index(k, dev, gen):
    return torch.randperm(k, device=dev, generator=gen).to(torch.int16)


def _rope(q, k):
    """LLaMA rotary embedding on [B, H, S, D] (reference: flashinfer pos_encoding_mode="ROPE_LLAMA")."""
    b, h, s, d = q.shape
    inv = 1.0 / (10000 ** (torch.arange(0, d, 2, device=q.device, dtype=torch.float32) / d))
    f = torch.outer(torch.arange(s, device=q.device, dtype=torch.float32), inv)
    cos, sin = torch.cat((f, f), -1).cos().to(q.dtype), torch.cat((f, f), -1).sin().to(q.dtype)

    def rot(x):
        return torch.cat((-x[..., d // 2:], x[..., : d // 2]), -1)

    return q * cos + rot(q) * sin, k * cos + rot(k) * sin


class ArcDecoderLayer(torch.nn.Module):
    """fused=True (default): 4 quantise + 4 GEMM launches per layer -- q/k/v as ONE GEMM over the concatenated weight with
    per-column scales, gate/up as ONE GEMM whose epilogue writes bf16(silu(gate)) * up, per-tensor pre-scales inside the
    quantise kernels.  fused=False: the reference's launch structure (7 GEMMs, SiLU*up and the pre-scales in torch)."""

    def __init__(self, cfg: PrefillConfig, dev, gen, world=1, rank=0, group=None, keep_ref=False, fused=True, calib=None,
                 layer_idx=0):
        """calib: (reorder_index, select_nums) dicts in the reference's on-disk schema (arcquant_b200/calib.py); without it
        every quantise input gets a random permutation and the configured select_num."""
        super().__init__()
        from . import agemm
        self.ref = {} if keep_ref else None  # bf16 weights / biases kept only for the parity test
        self.agemm, self.cfg, self.world, self.rank, self.group = agemm, cfg, world, rank, group
        H, I = cfg.hidden, cfg.intermediate
        if cfg.heads % world or I % (64 * world):
            raise ValueError("heads=%d / intermediate=%d do not shard over %d ranks" % (cfg.heads, I, world))
        self.local_heads = cfg.heads // world
        self.head_dim = H // cfg.heads
        self.fused = bool(fused) and (H // world) % 128 == 0 and (I // world) % 128 == 0
        self.idx_h1, self.idx_h2 = _rand_index(H, dev, gen), _rand_index(H, dev, gen)  # inputs of qkv and of gate/up
        idx_o, idx_d = _rand_index(H, dev, gen), _rand_index(I, dev, gen)
        ke_qkv = ke_gu = ke_o = cfg.ke_hidden
        ke_d = cfg.ke_intermediate
        if calib is not None:
            from . import calib as C
            e = C.layer_entries(calib[0], calib[1], layer_idx, dev)
            (self.idx_h1, ke_qkv), (idx_o, ke_o), (self.idx_h2, ke_gu), (idx_d, ke_d) = e["qkv"], e["o"], e["gate_up"], e["down"]
        self.ke_qkv, self.ke_gu = ke_qkv, ke_gu
        self.norm1 = torch.ones(H, dtype=torch.bfloat16, device=dev)
        self.norm2 = torch.ones(H, dtype=torch.bfloat16, device=dev)

        def bias(n):
            return (torch.randn(n, device=dev, generator=gen) * 0.02).to(torch.bfloat16) if cfg.qkv_bias else None

        def weight(name, n, k, b=None):
            w = _rand_weight(n, k, dev, gen)
            if self.ref is not None:
                self.ref[name] = (w, b)
            return w

        # rmsnorm_quantize_x is always the 16-wide flavour (bindings.cpp:243): the weights it feeds must be too
        F16 = agemm.FLAVOUR_16

        def col(name, n, k, idx, ke, b=None):
            return tp.ColumnParallelQLinear(weight(name, n, k, b), idx, ke, world, rank, b, flavour=F16)

        self.q = col("q", H, H, self.idx_h1, ke_qkv, bias(H))
        self.k = col("k", H, H, self.idx_h1, ke_qkv, bias(H))
        self.v = col("v", H, H, self.idx_h1, ke_qkv, bias(H))
        self.gate = col("gate", I, H, self.idx_h2, ke_gu)
        self.up = col("up", I, H, self.idx_h2, ke_gu)
        self.o = tp.RowParallelQLinear(weight("o", H, H), idx_o, ke_o, world, rank, group=group)
        self.down = tp.RowParallelQLinear(weight("down", H, I), idx_d, ke_d, world, rank, group=group)
        if self.fused:
            self.qkv_w = agemm.fuse_weights([(l.W, l.scale_w, l.scale) for l in (self.q, self.k, self.v)])
            self.qkv_bias = torch.cat([l.bias for l in (self.q, self.k, self.v)]) if cfg.qkv_bias else None
            self.gu_w = agemm.interleave_gate_up((self.gate.W, self.gate.scale_w, self.gate.scale),
                                                 (self.up.W, self.up.scale_w, self.up.scale))
            for l in (self.q, self.k, self.v, self.gate, self.up):  # the fused copies are the ones that run
                l.W = l.scale_w = None
            self.one = torch.ones(1, dtype=torch.float32, device=dev)

    def _norm_linears(self, x2d, w, idx, ke, layers):
        """One fused rmsnorm+quantise launch feeding several column-parallel GEMMs (modeling_arc.py:220-227)."""
        qx, sfx = self.agemm.rmsnorm_quantize_x(x2d, w, self.cfg.eps, idx, ke)
        return [l.forward_quantized(qx, sfx, 1.0) for l in layers]

    @torch.no_grad()
    def forward(self, h):
        ag = self.agemm
        B, S, H = h.shape
        x2d = h.reshape(B * S, H)
        if self.fused:
            qx, sfx = ag.rmsnorm_quantize_x(x2d, self.norm1, self.cfg.eps, self.idx_h1, self.ke_qkv)
            qkv = ag.matmul(qx, self.qkv_w[0], sfx, self.qkv_w[1], self.one, col_scale=self.qkv_w[2], bias=self.qkv_bias)
            q, k, v = qkv.split(H // self.world, dim=1)
        else:
            q, k, v = self._norm_linears(x2d, self.norm1, self.idx_h1, self.ke_qkv, (self.q, self.k, self.v))
        shp = (B, S, self.local_heads, self.head_dim)
        q, k, v = (t.reshape(shp).transpose(1, 2) for t in (q, k, v))
        q, k = _rope(q, k)
        a = F.scaled_dot_product_attention(q, k, v, is_causal=True).transpose(1, 2).reshape(B * S, -1).contiguous()
        o = self._row(self.o, a)                                # quantise (local permutation + ARC) -> GEMM -> all-reduce
        if self.fused and 2048 <= H <= 8192 and H % 64 == 0 and self.ke_gu % 64 == 0:
            # residual add + norm + quantise of the MLP block in ONE launch (modeling_arc.py:294-301): same bits as h + o
            qx, sfx, hs = ag.add_rmsnorm_quantize_x(o, x2d, self.norm2, self.cfg.eps, self.idx_h2, self.ke_gu)
            h = hs.view(B, S, H)
            m = ag.matmul(qx, self.gu_w[0], sfx, self.gu_w[1], self.one, col_scale=self.gu_w[2], epilogue=ag.EPILOGUE_SWIGLU)
            return h + self._row(self.down, m).view(B, S, H)
        h = h + o.view(B, S, H)
        x2d = h.reshape(B * S, H)
        if self.fused:
            qx, sfx = ag.rmsnorm_quantize_x(x2d, self.norm2, self.cfg.eps, self.idx_h2, self.ke_gu)
            m = ag.matmul(qx, self.gu_w[0], sfx, self.gu_w[1], self.one, col_scale=self.gu_w[2], epilogue=ag.EPILOGUE_SWIGLU)
        else:
            g, u = self._norm_linears(x2d, self.norm2, self.idx_h2, self.ke_gu, (self.gate, self.up))
            m = (F.silu(g) * u).contiguous()
        return h + self._row(self.down, m).view(B, S, H)

    def _row(self, layer, x_local):
        y = layer.forward_quantized(*layer.quantize_input(x_local, fused=self.fused))
        if self.world > 1:
            dist.all_reduce(y, op=dist.ReduceOp.SUM, group=self.group)
        return y

    # ---- sequence-parallel form over peer memory (SeqParallelExchange below) ----
    @torch.no_grad()
    def forward_sp(self, xc: "SeqParallelExchange", h_loc, pending, B, S):
        """xc = None: the whole token range on this rank (one GPU, or replicated activations with `_row`'s all-reduce) -- the
        same chain without the exchange steps.  Otherwise this rank owns tokens [rank * M / W, +M / W) of the residual
        stream (h_loc [M/W, H]); `pending` is the previous layer's (reduce-scattered) down_proj output, not yet added.  Per layer: 2 x (norm + quantise of the own tokens ->
        gather of codes + scale blocks by peer stores -> column-parallel GEMM) and 2 x (row-parallel GEMM whose epilogue
        scatters into the owners' slots -> barrier + slot sum); both residual adds ride in front of the norms
        (add_rmsnorm_quantize_x).  No NCCL call, no torch pass over the residual stream.  Returns (h_loc, pending)."""
        ag, H, W, eps = self.agemm, self.cfg.hidden, self.world, self.cfg.eps
        M = B * S
        if pending is None:
            qx, sfx = ag.rmsnorm_quantize_x(h_loc, self.norm1, eps, self.idx_h1, self.ke_qkv)
        else:
            qx, sfx, h_loc = ag.add_rmsnorm_quantize_x(pending, h_loc, self.norm1, eps, self.idx_h1, self.ke_qkv)
        qf, sff = (qx, sfx) if xc is None else xc.gather(qx, sfx, H + self.ke_qkv)
        qkv = ag.matmul(qf, self.qkv_w[0], sff, self.qkv_w[1], self.one, col_scale=self.qkv_w[2], bias=self.qkv_bias)
        q, k, v = qkv.split(H // W, dim=1)
        shp = (B, S, self.local_heads, self.head_dim)
        q, k, v = (t.reshape(shp).transpose(1, 2) for t in (q, k, v))
        q, k = _rope(q, k)
        a = F.scaled_dot_product_attention(q, k, v, is_causal=True).transpose(1, 2).reshape(M, -1).contiguous()
        o_loc = self._row(self.o, a) if xc is None else xc.row_linear(self.o, a)
        qx, sfx, h_loc = ag.add_rmsnorm_quantize_x(o_loc, h_loc, self.norm2, eps, self.idx_h2, self.ke_gu)
        qf, sff = (qx, sfx) if xc is None else xc.gather(qx, sfx, H + self.ke_gu)
        m = ag.matmul(qf, self.gu_w[0], sff, self.gu_w[1], self.one, col_scale=self.gu_w[2], epilogue=ag.EPILOGUE_SWIGLU)
        return h_loc, (self._row(self.down, m) if xc is None else xc.row_linear(self.down, m))


class SeqParallelExchange:
    """Peer-memory buffers of the sequence-parallel prefill (one set per model: every layer has the same shapes) and the
    two exchange steps on them (arcquant_b200/peer.py).  Reuse is safe without double buffering because the calls strictly
    alternate: a peer reaches its next store into a buffer only after a barrier this rank entered after its last read."""

    def __init__(self, cfg: PrefillConfig, dev, tokens: int, ke_max: int, agemm):
        from . import peer as P
        self.P, self.agemm = P, agemm
        self.group = P.PeerGroup(dist, dev).init_flags()
        W = self.group.world
        if tokens % (128 * W):
            raise ValueError("sequence-parallel prefill needs tokens %% (128 * world) == 0")
        self.M, self.Mloc, self.H = tokens, tokens // W, cfg.hidden
        K = cfg.hidden + ke_max
        self.qbuf = P.PeerBuffer(self.group, tokens * K // 2)
        self.sbuf = P.PeerBuffer(self.group, agemm.sf_bytes(tokens, K))
        self.recv = P.PeerBuffer(self.group, P.world_bytes(W, self.Mloc, cfg.hidden))
        self.out = [torch.empty((self.Mloc, cfg.hidden), dtype=torch.bfloat16, device=dev) for _ in range(2)]
        self.flip = 0

    def gather(self, qx, sfx, K):
        """codes [M/W, K/2] + scale blocks of the own tokens -> every rank's [M, K/2] / swizzled scale buffer"""
        g, r = self.group, self.group.rank
        nsf = self.Mloc * K // 16
        g.allgather(qx, self.qbuf, r * qx.numel(), sfx[:nsf], self.sbuf, r * nsf, channel=1)
        return self.qbuf.view(torch.uint8, (self.M, K // 2)), self.sbuf.view(torch.uint8, (self.agemm.sf_bytes(self.M, K),))

    def row_linear(self, layer, x_local):
        """row-parallel linear: quantise the local K shard, GEMM with the scatter epilogue, barrier + slot sum -> [M/W, H]"""
        qx, sfx, sc = layer.quantize_input(x_local, fused=True)
        self.group.gemm_scatter(qx, layer.W, sfx, layer.scale_w, sc * layer.scale, self.recv, self.Mloc)
        out = self.out[self.flip]
        self.flip ^= 1
        return self.group.reduce(self.recv, self.Mloc, self.H, out, channel=0)


def reference_layer_forward(layer: "ArcDecoderLayer", h: torch.Tensor) -> torch.Tensor:
    """The same decoder layer in plain fp32 torch with the UNQUANTISED weights (needs keep_ref=True): test oracle."""
    c, r = layer.cfg, layer.ref
    lin = lambda name, x: F.linear(x, r[name][0].float(), None if r[name][1] is None else r[name][1].float())

    def rms(x, w):
        return x * torch.rsqrt(x.pow(2).mean(-1, keepdim=True) + c.eps) * w.float()

    B, S, H = h.shape
    h = h.float()
    x = rms(h, layer.norm1)
    shp = (B, S, c.heads, H // c.heads)
    q, k, v = (lin(n, x).view(shp).transpose(1, 2) for n in ("q", "k", "v"))
    q, k = _rope(q, k)
    a = F.scaled_dot_product_attention(q, k, v, is_causal=True).transpose(1, 2).reshape(B, S, H)
    h = h + lin("o", a)
    x = rms(h, layer.norm2)
    return h + lin("down", F.silu(lin("gate", x)) * lin("up", x))


class ArcPrefillModel(torch.nn.Module):
    def __init__(self, cfg: PrefillConfig, device, seed=0, world=1, rank=0, group=None, keep_ref=False, fused=True, calib=None):
        super().__init__()
        gen = torch.Generator(device=device).manual_seed(seed)  # same seed on every rank: shards of the same weights
        self.cfg = cfg
        self.layers = torch.nn.ModuleList([ArcDecoderLayer(cfg, device, gen, world, rank, group, keep_ref, fused, calib, i)
                                           for i in range(cfg.layers)])

    @torch.no_grad()
    def forward(self, h):
        c = self.cfg
        if all(l.fused for l in self.layers) and 2048 <= c.hidden <= 8192 and c.hidden % 64 == 0 and \
                all(l.ke_qkv % 64 == 0 and l.ke_gu % 64 == 0 for l in self.layers):
            # every residual add rides in front of the next norm (add_rmsnorm_quantize_x): no torch pass over the stream
            B, S, H = h.shape
            x, pending = h.reshape(B * S, H), None
            for layer in self.layers:
                x, pending = layer.forward_sp(None, x, pending, B, S)
            return (x + pending).view(B, S, H)
        for layer in self.layers:
            h = layer(h)
        return h

    def enable_seq_parallel(self, dev, tokens: int):
        """Collective: set up the peer-memory exchange for `tokens` = B * S tokens per forward (tensor parallel, world > 1)."""
        l0 = self.layers[0]
        if not all(l.fused for l in self.layers):
            raise ValueError("the sequence-parallel path runs the fused launches")
        self.xc = SeqParallelExchange(self.cfg, dev, tokens, max(max(l.ke_qkv, l.ke_gu) for l in self.layers), l0.agemm)
        return self.xc

    @torch.no_grad()
    def forward_seq_parallel(self, h, gather_output: bool = True):
        """h [B, S, H] replicated on every rank (as forward takes it); the residual stream is kept token-sharded through
        the whole stack.  Returns the full [B, S, H] (one all-gather at the very end) or this rank's [M/W, H] rows."""
        B, S, H = h.shape
        xc = self.xc
        r = xc.group.rank
        h_loc = h.reshape(B * S, H)[r * xc.Mloc:(r + 1) * xc.Mloc].contiguous()
        pending = None
        for layer in self.layers:
            h_loc, pending = layer.forward_sp(xc, h_loc, pending, B, S)
        h_loc = h_loc + pending
        if not gather_output:
            return h_loc
        full = torch.empty((B * S, H), dtype=h_loc.dtype, device=h_loc.device)
        dist.all_gather_into_tensor(full, h_loc)
        return full.view(B, S, H)

    def linear_flops(self, tokens: int) -> float:
        c = self.cfg
        per = 2.0 * tokens * (4 * c.hidden * (c.hidden + c.ke_hidden) + 2 * c.intermediate * (c.hidden + c.ke_hidden)
                              + c.hidden * (c.intermediate + c.ke_intermediate))
        return per * c.layers
