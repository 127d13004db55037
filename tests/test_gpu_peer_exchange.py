"""GPU: the NVLink peer-memory all-gather (lzb_peer_allgather / sharding.PeerExchange) against plain concatenation.
One GPU: `world` ranks inside one process, each on its own stream (same kernel, same flag protocol).  Two or more GPUs
(skipped otherwise): one process per GPU under torchrun, buffers shared through CUDA IPC, checked against NCCL."""
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("world,rows,A", [(2, 64, 18), (4, 1000, 9), (8, 33, 5), (1, 17, 3)])
def test_in_process_ranks_exchange_bit_exact(world, rows, A):
    import torch
    from lightzero_b200.sharding import PeerExchange
    dev = torch.device("cuda")
    g = torch.Generator(device=dev).manual_seed(world * 1000 + rows)
    ranks = PeerExchange.local_group(world, rows, A + 1 + A)
    streams = [torch.cuda.Stream() for _ in range(world)]
    try:
        for epoch in range(5):  # more than two epochs: both halves of the double buffer are reused
            visits = [torch.randint(0, 50, (rows, A), generator=g, device=dev, dtype=torch.int32) for _ in range(world)]
            values = [torch.randn(rows, generator=g, device=dev) for _ in range(world)]
            pol = [torch.rand(rows, A, generator=g, device=dev) for _ in range(world)]
            torch.cuda.synchronize()
            # all ranks live in this one process: enqueue every rank's kernel before anything that could make the host
            # wait for a running kernel (a first-use module load while rank 0 spins for rank 1 would stall both)
            pending, outs = [], []
            for r in range(world):
                with torch.cuda.stream(streams[r]):
                    pending.append(ranks[r].launch([visits[r], values[r], pol[r]]))
            for r in range(world):
                with torch.cuda.stream(streams[r]):
                    outs.append(ranks[r].collect(pending[r]))
            torch.cuda.synchronize()
            for r in range(world):
                assert ranks[r].error() == 0
                v, x, p = outs[r]
                assert v.dtype == torch.int32 and x.dtype == torch.float32 and p.shape == (world * rows, A)
                assert torch.equal(v, torch.cat(visits)) and torch.equal(x, torch.cat(values)) and torch.equal(p, torch.cat(pol))
    finally:
        for x in ranks:
            x.close()


def test_a_missing_peer_times_out_instead_of_hanging():
    import torch
    from lightzero_b200.sharding import PeerExchange
    from lightzero_b200.sharding import PeerTimeout
    ranks = PeerExchange.local_group(2, 8, 4)
    try:
        mine = torch.arange(32, dtype=torch.int32, device="cuda").reshape(8, 4)
        with pytest.raises(PeerTimeout, match="rank 1"):  # rank 1 never calls: the library path raises ...
            ranks[0].gather([mine], timeout_ms=50)
        assert ranks[0].error() == 0  # ... and a reported error is cleared
        (out,) = ranks[0].gather([mine], timeout_ms=50, check=False)  # the caller-checked flavour
        assert ranks[0].error() == 2  # 1 + the rank that did not arrive
        assert torch.equal(out[:8], mine)
        assert (out[8:] == -2 ** 31).all(), "the missing rank's rows must be poisoned, never an earlier epoch's data"
    finally:
        for x in ranks:
            x.close()


def test_argument_errors():
    import torch
    from lightzero_b200.sharding import PeerExchange
    ranks = PeerExchange.local_group(1, 4, 3)
    try:
        with pytest.raises(ValueError):
            ranks[0].gather([torch.zeros(4, 2, dtype=torch.int32, device="cuda")])  # wrong row width
        with pytest.raises(ValueError):
            ranks[0].gather([torch.zeros(4, 3, dtype=torch.int64, device="cuda")])  # not 32-bit
    finally:
        ranks[0].close()


_WORKER = r"""
import os, sys, torch, torch.distributed as dist
sys.path.insert(0, %r)
from lightzero_b200.sharding import PeerExchange, gather_search_outputs
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl", device_id=dev)
rows, A = 512, 18
peer = PeerExchange(rows, A + 1)
g = torch.Generator(device=dev).manual_seed(7 + rank)
for epoch in range(6):
    visits = torch.randint(0, 50, (rows, A), generator=g, device=dev, dtype=torch.int32)
    values = torch.randn(rows, generator=g, device=dev)
    a = gather_search_outputs(visits, values, [rows] * world, peer=peer)
    b = gather_search_outputs(visits, values, [rows] * world)          # the NCCL route
    assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1]), (rank, epoch)
assert peer.error() == 0
dist.barrier()
peer.close()
dist.destroy_process_group()
print("rank", rank, "ok")
"""


def test_multi_process_exchange_matches_nccl(tmp_path):
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs two GPUs on one box")
    n = min(n, 8)
    script = tmp_path / "worker.py"
    script.write_text(_WORKER % ROOT)
    res = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=%d" % n,
                          "--master-addr", "127.0.0.1", "--master-port", "29641", str(script)],
                         capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert res.stdout.count("ok") == n
