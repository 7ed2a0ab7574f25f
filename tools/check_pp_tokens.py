"""Multi-GPU check (developer tool, run under torchrun): the pipelined multi-process decode stream
produces exactly the tokens of a single-process, single-stage run of the same seeded model.

    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/check_pp_tokens.py
"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200stage  # noqa: E402
from b200stage.pipeline import DistPipeline, Pipeline  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
dev = torch.device(f"cuda:{local}")
torch.cuda.set_device(dev)
dist.init_process_group("nccl", device_id=dev)
model = sys.argv[1] if len(sys.argv) > 1 else "tiny-llama-test"
cfg = b200stage.get_config(model)
P, NEW, NP = 40, 24, world + 1               # more prompts than stages
pipe = DistPipeline.load(cfg, rank, world, dev, seed=3, fast_init=True, max_rows=NP, max_seq=P + NEW + 8)
pipe._bcast_len = lambda T: P
if rank == 0:
    print(f"handoff: {pipe.p2p['how'] if getattr(pipe, 'p2p', None) else 'NCCL send/recv'}; shared lm_head: {pipe.split_head}", flush=True)
rng = np.random.default_rng(11)
prompts = [torch.from_numpy(rng.integers(0, cfg.vocab, P)).to(dev) for _ in range(NP)]
for rep in range(2):                           # second pass re-uses the captured graphs
    lgs = pipe.prefill_stream(prompts if pipe.first else [None] * NP, 16)
    toks = []
    for p in range(NP):
        t = None
        if pipe.last:
            t = lgs[p].view(1, -1).argmax(-1)
        if pipe.first and not pipe.last:
            t = torch.empty((1,), dtype=torch.int64, device=dev)
        toks.append(pipe.return_tokens(t))
    out = pipe.decode_stream(toks, [[P] for _ in range(NP)], [[p] for p in range(NP)], NEW - 1)
    if rank == 0:
        got = torch.stack([torch.cat([toks[p]] + out[p]) for p in range(NP)]).cpu().numpy()
        ref = Pipeline.load(cfg, 1, devices=[dev], seed=3, fast_init=True)
        want = np.stack([ref.generate(prompts[p], max_new_tokens=NEW, chunk=16).cpu().numpy()[P:] for p in range(NP)])
        assert np.array_equal(got, want), (got, want)
        print(f"pass {rep}: {NP} prompts x {NEW} tokens identical across {world} stages (lm_head: {getattr(pipe, 'head_path', '?')})", flush=True)
pipe.drain()
dist.barrier()
dist.destroy_process_group()
