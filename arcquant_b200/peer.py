"""Host side of the peer-memory exchange (include/arcq.h "Tensor-parallel exchange over peer memory", csrc/peer.cu).

The reference has no distributed code; this is the B200 answer to SURVEY.md §8 f4 for the sharding the
north star names: row-parallel GEMMs whose epilogue stores straight into the owner rank's receive slots over NVLink,
finished by one barrier + local fp32 sum (``PeerGroup.reduce``), and the all-gather of NVFP4 codes + scale blocks as plain
peer stores (``PeerGroup.allgather``).  No NCCL kernel runs on this data path; ``torch.distributed`` is used ONCE, at
set-up, to carry the 64-byte CUDA IPC handles between the ranks of the node.

Everything here goes through the C ABI (ctypes): a handful of calls per layer, none of them in an inner loop.
"""
from __future__ import annotations

import ctypes

import torch

from . import _lib

MAX_PEERS, FLAG_BYTES, CHANNELS = 8, 4096, 32


class _Raw:
    """Exposes a raw device allocation to torch through __cuda_array_interface__ (no copy, no ownership)."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3, "strides": None}


class PeerBuffer:
    """One allocation per rank, mapped into every rank: ``ptrs[r]`` = rank r's copy as seen from this process."""

    def __init__(self, group: "PeerGroup", nbytes: int):
        L = _lib.load()
        self.group, self.nbytes, self.local = group, int(nbytes), 0
        err = None
        try:
            p = ctypes.c_void_p()
            with torch.cuda.device(group.device):
                _lib.check(L.arcq_peer_alloc(self.nbytes, ctypes.byref(p)), "arcq_peer_alloc")
            self.local = int(p.value)
        except Exception as e:  # noqa: BLE001 -- still take part in the exchange below, so that no rank is left waiting
            err = e
        self.ptrs = group._exchange(self.local, err)
        self._arr = (ctypes.c_void_p * group.world)(*self.ptrs)
        self.tensor = torch.as_tensor(_Raw(self.local, self.nbytes), device=group.device)   # uint8 view of the LOCAL copy

    def array(self):
        """ctypes array of `world` device pointers (what the `peer_*` arguments of the C ABI take)."""
        return self._arr

    def close(self):
        """Unmap the peers' copies and free the local one.  NOT collective, but every rank must have finished using the
        buffer (a barrier of the caller's choice): a peer that still stores into freed memory faults.  Idempotent."""
        L = _lib.load()
        local, self.local = self.local, 0
        if not local:
            return
        self.tensor = None
        with torch.cuda.device(self.group.device):
            if getattr(self.group, "_local", None) is None:          # multi-process: the other entries are IPC mappings
                for r, p in enumerate(self.ptrs):
                    if r != self.group.rank and p:
                        L.arcq_peer_close(ctypes.c_void_p(p))
            L.arcq_peer_free(ctypes.c_void_p(local))
        self.ptrs = []

    def view(self, dtype, shape, byte_offset: int = 0):
        n = 1
        for s in shape:
            n *= int(s)
        nb = n * torch.empty((), dtype=dtype).element_size()
        return self.tensor[byte_offset:byte_offset + nb].view(dtype).view(*shape)


class PeerGroup:
    """The ranks of ONE node that map each other's memory.

    ``PeerGroup(dist)``: one process per GPU (torchrun); handles travel through ``dist.all_gather_object`` once per buffer.
    ``PeerGroup.single_process(devices)``: one process driving several GPUs (tests): a list of groups, one per device.
    """

    def __init__(self, dist=None, device=None, _local=None):
        self._local = _local
        if _local is not None:
            self.world, self.rank, self.device = _local["world"], _local["rank"], _local["device"]
        else:
            self.world, self.rank = dist.get_world_size(), dist.get_rank()
            self.device = device if device is not None else torch.device("cuda", torch.cuda.current_device())
            self.dist = dist
        if self.world > MAX_PEERS:
            raise RuntimeError("peer exchange supports at most %d ranks, got %d" % (MAX_PEERS, self.world))
        self.flags = None

    def init_flags(self):
        """Allocate the flag block (collective: every rank calls it once, in the same order as its other buffers)."""
        self.flags = PeerBuffer(self, FLAG_BYTES)
        return self

    @classmethod
    def single_process(cls, devices):
        shared = {"pending": {}, "n": len(devices)}
        groups = []
        L = _lib.load()
        for r, d in enumerate(devices):
            dev = torch.device("cuda", d) if isinstance(d, int) else d
            with torch.cuda.device(dev):
                for o in devices:
                    oi = o if isinstance(o, int) else o.index
                    _lib.check(L.arcq_peer_enable(oi), "arcq_peer_enable")
            groups.append(cls(_local={"world": len(devices), "rank": r, "device": dev, "shared": shared}))
        return groups

    @staticmethod
    def alloc_all(groups, nbytes):
        """single-process mode: allocate one PeerBuffer per group (the pointers ARE valid on every device)."""
        L = _lib.load()
        locs = []
        for g in groups:
            p = ctypes.c_void_p()
            with torch.cuda.device(g.device):
                _lib.check(L.arcq_peer_alloc(int(nbytes), ctypes.byref(p)), "arcq_peer_alloc")
            locs.append(int(p.value))
        bufs = []
        for g in groups:
            b = PeerBuffer.__new__(PeerBuffer)
            b.group, b.nbytes, b.local, b.ptrs = g, int(nbytes), locs[g.rank], list(locs)
            b._arr = (ctypes.c_void_p * g.world)(*locs)
            b.tensor = torch.as_tensor(_Raw(b.local, b.nbytes), device=g.device)
            bufs.append(b)
        return bufs

    def _exchange(self, local_ptr: int, err=None):
        """Collective: every rank contributes the IPC handle of its allocation (or its failure) and maps the others'.
        Two object all-gathers, so that a failure on ANY rank raises on EVERY rank and nobody is left in a collective."""
        if self._local is not None:
            raise RuntimeError("single-process groups allocate with PeerGroup.alloc_all")
        L = _lib.load()
        mine = None
        if err is None:
            try:
                h = ctypes.create_string_buffer(64)
                _lib.check(L.arcq_peer_export(ctypes.c_void_p(local_ptr), h), "arcq_peer_export")
                mine = bytes(h.raw)
            except Exception as e:  # noqa: BLE001
                err = e
        handles = [None] * self.world
        self.dist.all_gather_object(handles, mine)
        ptrs = []
        if all(h is not None for h in handles):
            try:
                with torch.cuda.device(self.device):
                    for r, hb in enumerate(handles):
                        if r == self.rank:
                            ptrs.append(local_ptr)
                            continue
                        p = ctypes.c_void_p()
                        _lib.check(L.arcq_peer_open(ctypes.create_string_buffer(hb, 64), ctypes.byref(p)), "arcq_peer_open")
                        ptrs.append(int(p.value))
            except Exception as e:  # noqa: BLE001
                err = e
        elif err is None:
            err = RuntimeError("a peer rank could not allocate or export its buffer")
        oks = [None] * self.world
        self.dist.all_gather_object(oks, None if err is None else str(err).splitlines()[0][:200])
        bad = [(r, m) for r, m in enumerate(oks) if m is not None]
        if bad:
            raise RuntimeError("peer memory set-up failed on rank %d: %s" % bad[0])
        return ptrs

    # ---- the three data-path calls ---------------------------------------------------------------------------------
    def _stream(self):
        return torch.cuda.current_stream(self.device).cuda_stream

    def gemm_scatter(self, A, B, SFA, SFB, scale, recv: PeerBuffer, rows_per_rank: int, col_scale=None, ldd: int = 0):
        """Row-parallel GEMM: this rank's partial of rows [o * rows_per_rank, +rows_per_rank) lands in rank o's `recv`,
        slot `rank` ([world][rows_per_rank][N] bf16)."""
        M, N, K = A.shape[0], B.shape[0], A.shape[1] * 2
        if world_bytes(self.world, rows_per_rank, ldd or N) > recv.nbytes:
            raise RuntimeError("receive buffer too small")
        alpha, alpha_dev = (0.0, scale.data_ptr()) if isinstance(scale, torch.Tensor) else (float(scale), None)
        with torch.cuda.device(self.device):
            _lib.check(_lib.load().arcq_gemm_nvfp4_scatter(
                A.data_ptr(), B.data_ptr(), SFA.data_ptr(), SFB.data_ptr(), M, N, K, alpha, alpha_dev,
                col_scale.data_ptr() if col_scale is not None else None, recv.array(), self.world, self.rank, rows_per_rank, ldd,
                self._stream()), "arcq_gemm_nvfp4_scatter")

    def reduce(self, recv: PeerBuffer, rows: int, cols: int, out, channel: int = 0):
        """Barrier, then out[rows, cols] = sum over the `world` slots of the local receive buffer (fp32, rank order)."""
        with torch.cuda.device(self.device):
            _lib.check(_lib.load().arcq_peer_reduce(recv.local, self.flags.array(), self.world, self.rank, channel, rows, cols, 0,
                                                    out.data_ptr(), out.stride(0), self._stream()), "arcq_peer_reduce")
        return out

    def allgather(self, src0, dst0: PeerBuffer, off0: int, src1=None, dst1: PeerBuffer = None, off1: int = 0, channel: int = 1):
        """Copies the local byte tensors src0 (and src1) to byte offset off0 (off1) of every rank's dst0 (dst1), then barrier."""
        b1 = src1.numel() * src1.element_size() if src1 is not None else 0
        with torch.cuda.device(self.device):
            _lib.check(_lib.load().arcq_peer_allgather(
                src0.data_ptr(), src0.numel() * src0.element_size(), dst0.array(), off0,
                src1.data_ptr() if src1 is not None else None, b1, dst1.array() if dst1 is not None else None, off1,
                self.flags.array(), self.world, self.rank, channel, self._stream()), "arcq_peer_allgather")

    def status(self):
        """Raises if any barrier of this group timed out (synchronises the current stream)."""
        with torch.cuda.device(self.device):
            _lib.check(_lib.load().arcq_peer_status(self.flags.local, self._stream()), "arcq_peer_status")


def world_bytes(world: int, rows_per_rank: int, ld: int) -> int:
    """Bytes of one receive buffer: bf16 [world][rows_per_rank][ld]."""
    return world * rows_per_rank * ld * 2
