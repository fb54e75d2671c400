"""Host->device copy calibration under the bench's launch: what the box's host side gives N concurrent ranks.

The end-to-end figure of bench.py (`e2e`) moves 36.2 MB of page-locked input per 64-frame step and rank; at 8 GPUs it
reached 0.46 of 8x the single-GPU rate in round 1 while the device-resident figure scaled at 0.97.  This tool takes
the decode out of the picture: every rank copies the same 36.2 MB pinned buffer to its GPU back to back (a bare
cudaMemcpyAsync loop, CUDA-event timed), first ALONE (the other ranks idle behind a barrier), then ALL ranks at once.
If the concurrent per-rank bandwidth drops the way `e2e` does, the limiter is the host's PCIe / memory side (VM root
ports, one NUMA node), not the library's copy pattern.

usage (like bench.py):  python tools/h2d_calib.py
                        python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
                            --master-port 29533 tools/h2d_calib.py
Rank 0 prints one JSON line.
"""
import json
import os
import sys

import torch

BYTES = 64 * 46 * 54 * 57 * 4          # one c2 step's input
REPS = 40


def main():
    # stdout carries exactly one JSON line: NCCL prints its version banner on file descriptor 1 from C code
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist = None
    if world > 1:
        import torch.distributed as dist_
        dist_.init_process_group("nccl", device_id=dev)
        dist = dist_
    src = [torch.empty(BYTES, dtype=torch.uint8).pin_memory() for _ in range(4)]      # rotate like the bench's inputs
    for s in src:
        s.fill_(rank + 1)
    dst = torch.empty(BYTES, dtype=torch.uint8, device=dev)
    back = torch.empty(755200, dtype=torch.uint8).pin_memory()                          # the table copy of a step
    dsmall = torch.empty(755200, dtype=torch.uint8, device=dev)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def run(with_d2h):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for i in range(3):
            dst.copy_(src[i % 4], non_blocking=True)
        torch.cuda.synchronize(dev)
        e0.record()
        for i in range(REPS):
            dst.copy_(src[i % 4], non_blocking=True)
            if with_d2h:
                back.copy_(dsmall, non_blocking=True)
        e1.record()
        torch.cuda.synchronize(dev)
        return BYTES * REPS / (e0.elapsed_time(e1) * 1e-3) / 1e9      # GB/s

    alone = []
    for r in range(world):                 # one rank at a time, the others wait
        barrier()
        alone.append(run(False) if r == rank else 0.0)
    barrier()
    together = run(False)
    barrier()
    together_d2h = run(True)
    barrier()
    mine = torch.tensor([alone[rank], together, together_d2h], device=dev, dtype=torch.float64)
    if dist is not None:
        allv = torch.zeros((world, 3), device=dev, dtype=torch.float64)
        dist.all_gather_into_tensor(allv, mine[None])
    else:
        allv = mine[None]
    if rank == 0:
        v = allv.cpu().numpy()
        line = {"tool": "h2d_calib", "n_gpus": world, "bytes_per_copy": BYTES, "copies": REPS, "unit": "GB/s",
                "per_rank_alone": [float(x) for x in v[:, 0]], "per_rank_concurrent": [float(x) for x in v[:, 1]],
                "per_rank_concurrent_with_table_d2h": [float(x) for x in v[:, 2]],
                "sum_alone": float(v[:, 0].sum()), "sum_concurrent": float(v[:, 1].sum()),
                "concurrent_over_alone": float(v[:, 1].sum() / max(v[:, 0].sum(), 1e-9)),
                "frames_per_s_bound_concurrent": float(v[:, 1].sum() * 1e9 / (BYTES / 64)),
                "nproc": os.cpu_count()}
        os.write(real_stdout, (json.dumps(line) + "\n").encode())
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
