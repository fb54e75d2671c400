"""Sharded decode on real GPUs: BASELINE.json configs[2] -- CMU 656x368 (46x82x57 maps), batch 128,
split by image over WORLD_SIZE GPUs, human tables gathered on rank 0 over NCCL -- checked against the same batch
decoded on rank 0's GPU alone.  Launch: python -m torch.distributed.run --nproc-per-node N tools/multi_gpu_check.py"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tfpose_b200 as tp
from tfpose_b200 import shard, synth

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
N, h, w = 128, 46, 82
H, W = 4 * h, 4 * w
lo, hi = shard.shard_bounds(N, rank, world)
x = np.concatenate([synth.synth_batch(8, h, w, config_id=3, first_frame=f) for f in range(lo, hi, 8)])[:hi - lo]
dec = tp.PoseDecoder(max_batch=N, max_upsample=(H, W), device=local)
res = dec.decode(torch.from_numpy(x).to(dev), (H, W))
counts, recs = res.compact()
out = shard.gather_results(counts, recs, dst=0, device=dev)
if rank == 0:
    all_counts, all_recs = out
    xa = np.concatenate([synth.synth_batch(8, h, w, config_id=3, first_frame=f) for f in range(0, N, 8)])[:N]
    ra = dec.decode(torch.from_numpy(xa).to(dev), (H, W))
    c1, r1 = ra.compact()
    assert np.array_equal(all_counts, c1), "per-frame human counts differ between the sharded and the single-GPU decode"
    assert all_recs.tobytes() == r1.tobytes(), "human records differ"
    print("multi-GPU check OK: %d frames over %d GPUs, %d humans, gathered tables identical to the single-GPU decode"
          % (N, world, int(c1.sum())))
dec.close()
dist.barrier()
dist.destroy_process_group()
