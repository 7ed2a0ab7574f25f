"""Developer probe (torchrun, 2 GPUs): what one stage-to-stage hop of the decode ring costs on the device.

Rank 0 and rank 1 bounce an 8 KB hidden state: [store + flag into the peer's inbox] -> the peer's [wait for the flag +
copy] -> (optionally a small captured graph, as a stage's layers would be) -> store + flag back.  Everything is
enqueued ahead by the host; the time per round trip on rank 0's stream is 2 x (send + wait + what lies between them).

    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/hop_probe.py
"""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200stage  # noqa: E402,F401
from b200stage import ops  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
dev = torch.device(f"cuda:{local}")
torch.cuda.set_device(dev)
dist.init_process_group("nccl", device_id=dev)
ops._abi.init_device()
import torch.distributed._symmetric_memory as symm  # noqa: E402

NB = 8192 + 256
buf = symm.empty(NB, dtype=torch.uint8, device=dev)
hdl = symm.rendezvous(buf, dist.group.WORLD)
peer = hdl.get_buffer(1 - rank, (NB,), torch.uint8)
buf.zero_()
torch.cuda.synchronize()
dist.barrier()
err = torch.zeros((1,), dtype=torch.int32, device=dev)
x = torch.randn((1, 4096), dtype=torch.float16, device=dev)
y = torch.empty_like(x)
my_slot, my_flag = buf[:8192], buf[8192:8196].view(torch.int32)
peer_slot, peer_flag = peer[:8192], peer[8192:8196].view(torch.int32)


def small_graph(n_kernels):
    if n_kernels == 0:
        return None
    g = torch.cuda.CUDAGraph()
    a = torch.zeros((1024,), device=dev)
    s = torch.cuda.Stream(dev)
    with torch.cuda.graph(g, stream=s):
        for _ in range(n_kernels):
            a.add_(1.0)
    return g


def run(n_iter, base, graph):
    """rank 0: send, wait; rank 1: wait, send.  seq numbers base+1.."""
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(n_iter):
        seq = base + i + 1
        if rank == 0:
            ops.p2p_send(x, peer_slot, peer_flag, seq)
            ops.p2p_wait(my_flag, seq, src=my_slot, dst=y, err=err)
            if graph is not None:
                graph.replay()
        else:
            ops.p2p_wait(my_flag, seq, src=my_slot, dst=y, err=err)
            if graph is not None:
                graph.replay()
            ops.p2p_send(y, peer_slot, peer_flag, seq)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / n_iter


base = 0
for nk in (0, 1, 8, 24):
    g = small_graph(nk)
    run(50, base, g)
    base += 50
    t = run(500, base, g)
    base += 500
    if rank == 0:
        print(f"graph of {nk:2d} tiny kernels between wait and send: round trip {t:7.2f} us  (= 2 hops{' + 2 graphs' if nk else ''})", flush=True)
# the same kernels with both ends on this GPU's own memory: launch + execution cost without NVLink
if rank == 0:
    loc = torch.zeros((NB,), dtype=torch.uint8, device=dev)
    ls, lf = loc[:8192], loc[8192:8196].view(torch.int32)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(500):
        ops.p2p_send(x, ls, lf, i + 1)
        ops.p2p_wait(lf, i + 1, src=ls, dst=y, err=err)
    e1.record()
    torch.cuda.synchronize()
    print(f"send + wait on local memory: {e0.elapsed_time(e1) * 1e3 / 500:7.2f} us per pair; err {int(err.item())}", flush=True)
dist.barrier()
dist.destroy_process_group()
