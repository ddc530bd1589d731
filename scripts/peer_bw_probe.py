"""What can plain 16-byte peer stores reach on this box?  One process, two GPUs: arcq_peer_allgather (every SM streaming
uint4 stores to the local and the peer copy) with a large segment, both directions at once, beside cudaMemcpyPeer (copy
engine).  The scatter GEMM's NVLink rate (profiles/r2_bench_tp*_peer*.json) is to be read against these numbers."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from arcquant_b200 import peer as P  # noqa: E402


def main():
    nb = 256 << 20
    groups = P.PeerGroup.single_process([0, 1])
    flags = P.PeerGroup.alloc_all(groups, P.FLAG_BYTES)
    for g, f in zip(groups, flags):
        g.flags = f
    dst = P.PeerGroup.alloc_all(groups, 2 * nb)
    srcs = [torch.randint(0, 255, (nb,), dtype=torch.uint8, device=g.device) for g in groups]
    streams = [torch.cuda.Stream(device=g.device) for g in groups]
    out = {}
    for both in (True, False):
        ev = []
        for rep in range(4):
            for r, g in enumerate(groups):
                with torch.cuda.stream(streams[r]):
                    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
                    e0.record()
                    # `both` False: rank 1 sends 16 bytes only (it still takes part in the barrier)
                    n = nb if (both or r == 0) else 16
                    g.allgather(srcs[r][:n], dst[r], r * nb, channel=2)
                    e1.record()
                    ev.append((rep, r, e0, e1))
            for d in (0, 1):
                torch.cuda.synchronize(d)
        best = min(e0.elapsed_time(e1) for rep, r, e0, e1 in ev if rep > 0 and r == 0)
        out["peer_stores_%s_GBps" % ("bidirectional" if both else "one_direction")] = nb / best / 1e6
    for g in groups:
        g.status()
    a = torch.empty(nb, dtype=torch.uint8, device="cuda:0")
    b = torch.empty(nb, dtype=torch.uint8, device="cuda:1")
    for _ in range(2):
        b.copy_(a)
    torch.cuda.synchronize(0); torch.cuda.synchronize(1)
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    with torch.cuda.device(0):
        e0.record()
        for _ in range(4):
            b.copy_(a)
        e1.record()
    torch.cuda.synchronize(0); torch.cuda.synchronize(1)
    out["cudaMemcpyPeer_GBps"] = 4 * nb / e0.elapsed_time(e1) / 1e6
    print(json.dumps(out))


if __name__ == "__main__":
    main()
