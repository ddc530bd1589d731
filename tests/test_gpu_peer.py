"""The peer-memory exchange (include/arcq.h "Tensor-parallel exchange over peer memory", csrc/peer.cu) through the
C ABI: the GEMM whose epilogue scatters 128-row blocks into the owners' receive slots, the barrier + fp32 slot sum that
finishes the reduce-scatter, and the all-gather by peer stores.

Runs on ONE GPU: the "ranks" are emulated by several communicator members on the same device (their peer pointers are
then ordinary device pointers, the barrier protocol, the addressing and the kernels are exactly those of the multi-GPU
run); with two GPUs visible the same tests also run with one rank per device over real peer mappings, and
tests/test_gpu_multi.py runs the multi-process (CUDA IPC) path through bench.py.  Oracle: the plain GEMM entry point
(itself checked against float64 in test_gpu_gemm*.py) -- the scattered partial must be BIT-identical to it, the reduced
sum must equal the fp32 sum of the bf16 partials in rank order rounded once, bit for bit.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _rand_operand(rows, K, seed, dev, lo=0x2c, hi=0x44):
    from arcquant_b200 import agemm
    g = torch.Generator(device=dev).manual_seed(seed)
    q = torch.randint(0, 256, (rows, K // 2), dtype=torch.uint8, device=dev, generator=g)
    sf = torch.randint(lo, hi, (agemm.sf_bytes(rows, K),), dtype=torch.uint8, device=dev, generator=g)
    return q, sf


def _devices(world):
    n = torch.cuda.device_count()
    return [[0] * world] + ([list(range(world))] if n >= world and world > 1 else [])


def _groups(devs):
    from arcquant_b200 import peer as P
    groups = P.PeerGroup.single_process(devs)
    flags = P.PeerGroup.alloc_all(groups, P.FLAG_BYTES)
    for g, f in zip(groups, flags):
        g.flags = f
    return groups


@pytest.mark.parametrize("world,M,N,K", [(1, 256, 512, 320), (2, 512, 768, 4352), (4, 1024, 256, 576), (2, 768 - 40, 512 + 8, 1088)])
def test_scatter_gemm_and_reduce(arcq_option, world, M, N, K):
    from arcquant_b200 import agemm, peer as P
    arcq_option("gemm_pair", 1)                              # the plain launches below use the kernel the scatter entry point uses
    rpr = -(-M // (128 * world)) * 128                       # rows per rank, a multiple of 128
    for devs in _devices(world):
        groups = _groups(devs)
        ld = N
        recv = P.PeerGroup.alloc_all(groups, P.world_bytes(world, rpr, ld))
        streams = [torch.cuda.Stream(device=g.device) for g in groups]
        parts, outs, ops = [], [], []
        for r, g in enumerate(groups):
            qa, sfa = _rand_operand(M, K, 10 + r, g.device)
            qb, sfb = _rand_operand(N, K, 20 + r, g.device)
            ops.append((qa, sfa, qb, sfb))
            parts.append(agemm.matmul(qa, qb, sfa, sfb, 1e-3 * (r + 1)))           # the plain entry point: the oracle here
            outs.append(torch.full((rpr, N), float("nan"), dtype=torch.bfloat16, device=g.device))
        for d in set(devs):
            torch.cuda.synchronize(d)
        for rep in range(2):                                 # twice: the second barrier runs on epoch 2 of the same flags
            for r, g in enumerate(groups):
                with torch.cuda.stream(streams[r]):
                    qa, sfa, qb, sfb = ops[r]
                    g.gemm_scatter(qa, qb, sfa, sfb, 1e-3 * (r + 1), recv[r], rpr)
            for d in set(devs):
                torch.cuda.synchronize(d)                    # one device emulating several ranks: no spinning kernel may hold the SMs a GEMM needs
            for r, g in enumerate(groups):
                with torch.cuda.stream(streams[r]):
                    g.reduce(recv[r], rpr, N, outs[r], channel=3)
            for d in set(devs):
                torch.cuda.synchronize(d)
            for g in groups:
                g.status()
            for o, g in enumerate(groups):                   # owner o: rows [o * rpr, (o + 1) * rpr) of the global result
                lo, hi = o * rpr, min((o + 1) * rpr, M)
                if hi <= lo:
                    continue
                slots = recv[o].view(torch.bfloat16, (world, rpr, ld))
                acc = torch.zeros((hi - lo, N), dtype=torch.float32, device=g.device)
                for s in range(world):
                    want = parts[s][lo:hi].to(g.device)
                    assert torch.equal(slots[s, :hi - lo, :N], want), "slot %d of owner %d differs from the plain GEMM (rep %d)" % (s, o, rep)
                    acc += want.float()
                assert torch.equal(outs[o][:hi - lo], acc.to(torch.bfloat16)), "reduced rows of owner %d (rep %d)" % (o, rep)


@pytest.mark.parametrize("world", [1, 2, 4])
def test_allgather_by_peer_stores(world):
    from arcquant_b200 import peer as P
    n0, n1 = 128 * 2176, 128 * 272                            # one 128-row slab of codes / scale blocks at K = 4352
    for devs in _devices(world):
        groups = _groups(devs)
        d0 = P.PeerGroup.alloc_all(groups, world * n0)
        d1 = P.PeerGroup.alloc_all(groups, world * n1)
        streams = [torch.cuda.Stream(device=g.device) for g in groups]
        for rep in range(3):
            srcs = []
            for r, g in enumerate(groups):
                gen = torch.Generator(device=g.device).manual_seed(100 * rep + r)
                srcs.append((torch.randint(0, 256, (n0,), dtype=torch.uint8, device=g.device, generator=gen),
                             torch.randint(0, 256, (n1,), dtype=torch.uint8, device=g.device, generator=gen)))
            for d in set(devs):
                torch.cuda.synchronize(d)
            for r, g in enumerate(groups):
                with torch.cuda.stream(streams[r]):
                    if rep == 2:   # single-segment form
                        g.allgather(srcs[r][0], d0[r], r * n0, channel=5)
                    else:
                        g.allgather(srcs[r][0], d0[r], r * n0, srcs[r][1], d1[r], r * n1, channel=5)
            for d in set(devs):
                torch.cuda.synchronize(d)
            for g in groups:
                g.status()
            for r, g in enumerate(groups):
                want0 = torch.cat([s[0].to(g.device) for s in srcs])
                assert torch.equal(d0[r].tensor, want0), "codes gathered on rank %d (rep %d)" % (r, rep)
                if rep != 2:
                    want1 = torch.cat([s[1].to(g.device) for s in srcs])
                    assert torch.equal(d1[r].tensor, want1), "scale blocks gathered on rank %d (rep %d)" % (r, rep)


def test_peer_argument_errors():
    from arcquant_b200 import _lib, peer as P
    import ctypes
    L = _lib.load()
    groups = _groups([0])
    g = groups[0]
    recv = P.PeerGroup.alloc_all(groups, 4096)[0]
    x = torch.zeros(4096, dtype=torch.uint8, device="cuda")
    s = torch.cuda.current_stream().cuda_stream
    # rows_per_rank not a multiple of 128
    rc = L.arcq_gemm_nvfp4_scatter(x.data_ptr(), x.data_ptr(), x.data_ptr(), x.data_ptr(), 128, 128, 64, 1.0, None, None, recv.array(), 1, 0, 100, 0, s)
    assert rc == -1 and "rows_per_rank" in _lib.last_error()
    # more than ARCQ_MAX_PEERS ranks
    arr = (ctypes.c_void_p * 9)(*([recv.local] * 9))
    rc = L.arcq_peer_reduce(recv.local, arr, 9, 0, 0, 8, 8, 0, x.data_ptr(), 0, s)
    assert rc == -1 and "peers" in _lib.last_error()
    # channel out of range, odd sizes
    rc = L.arcq_peer_reduce(recv.local, g.flags.array(), 1, 0, 99, 8, 8, 0, x.data_ptr(), 0, s)
    assert rc == -1 and "channel" in _lib.last_error()
    rc = L.arcq_peer_allgather(x.data_ptr(), 24, recv.array(), 0, None, 0, None, 0, g.flags.array(), 1, 0, 0, s)
    assert rc == -1 and "16" in _lib.last_error()
    g.status()
