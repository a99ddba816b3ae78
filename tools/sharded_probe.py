"""Runs only bench.py's codebook-sharded leg (BASELINE configs[4], K = 262 144) under torchrun; rank 0 prints its record.
    B200VQ_TRACE=1 python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/sharded_probe.py
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

import bench
import video_tokenizer_b200 as V

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
V.load_library()
rec = bench.codebook_sharded_leg(V, dev, world, rank)
if rank == 0:
    print(json.dumps(rec))
dist.barrier()
dist.destroy_process_group()
