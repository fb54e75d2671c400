"""Multi-GPU sharding of a decode batch: frames are independent, so rank g of G
decodes the contiguous slice [g*N/G, (g+1)*N/G) on its own GPU with its own context.  No
collective touches the data path; the only communication is the final gather of the per-frame
human tables (a few hundred bytes per frame) to rank 0, done with ``torch.distributed``
(NCCL on the GPU box, gloo in the CPU tests)."""
import ctypes
import os
import platform

import numpy as np

from . import _lib


def shard_bounds(n_frames, rank, world):
    """[lo, hi) of rank's contiguous share; the first n_frames % world ranks get one more frame."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("bad rank/world %r/%r" % (rank, world))
    base, extra = divmod(int(n_frames), int(world))
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def pack_result(counts, records):
    """bytes-like uint8 array: [int32 n_frames][int32 counts[n_frames]][tfpd_human records]."""
    counts = np.ascontiguousarray(counts, np.int32)
    records = np.ascontiguousarray(records, _lib.HUMAN_DTYPE)
    assert int(counts.sum()) == len(records)
    head = np.array([len(counts)], np.int32)
    return np.concatenate([head.view(np.uint8), counts.view(np.uint8), records.view(np.uint8).reshape(-1)])


def unpack_result(buf):
    buf = np.ascontiguousarray(buf, np.uint8)
    n = int(buf[:4].view(np.int32)[0])
    counts = buf[4:4 + 4 * n].view(np.int32).copy()
    rec_bytes = buf[4 + 4 * n:4 + 4 * n + int(counts.sum()) * _lib.HUMAN_DTYPE.itemsize]
    return counts, rec_bytes.view(_lib.HUMAN_DTYPE).copy()


def gather_results(counts, records, group=None, dst=0, device=None):
    """Gather every rank's (counts, records) on ``dst``; returns there the concatenation in rank
    (== frame) order as (counts[N_total], records[sum]) and None elsewhere."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    dev = torch.device(device) if device is not None else torch.device("cpu")
    payload = torch.from_numpy(pack_result(counts, records)).to(dev)
    sizes = [torch.zeros(1, dtype=torch.int64, device=dev) for _ in range(world)]
    dist.all_gather(sizes, torch.tensor([payload.numel()], dtype=torch.int64, device=dev), group=group)
    cap = int(max(int(s.item()) for s in sizes))
    padded = torch.zeros(cap, dtype=torch.uint8, device=dev)
    padded[:payload.numel()] = payload
    bufs = [torch.zeros(cap, dtype=torch.uint8, device=dev) for _ in range(world)] if rank == dst else None
    dist.gather(padded, bufs, dst=dst, group=group)
    if rank != dst:
        return None
    all_counts, all_recs = [], []
    for r in range(world):
        c, rec = unpack_result(bufs[r][:int(sizes[r].item())].cpu().numpy())
        all_counts.append(c)
        all_recs.append(rec)
    return np.concatenate(all_counts), np.concatenate(all_recs)


def split_records(counts, records):
    """list (per frame) of record slices."""
    out, o = [], 0
    for c in counts:
        out.append(records[o:o + int(c)])
        o += int(c)
    return out


def _parse_cpulist(text):
    cpus = set()
    for part in text.strip().split(","):
        if not part:
            continue
        lo, _, hi = part.partition("-")
        cpus.update(range(int(lo), int(hi or lo) + 1))
    return cpus


def gpu_numa_node(pci_domain, pci_bus, pci_device, sysfs="/sys"):
    """NUMA node the GPU's PCIe slot hangs off (``/sys/bus/pci/devices/<id>/numa_node``), or -1 if unknown."""
    path = os.path.join(sysfs, "bus/pci/devices/%04x:%02x:%02x.0/numa_node" % (pci_domain, pci_bus, pci_device))
    try:
        with open(path) as fh:
            return int(fh.read().strip())
    except (OSError, ValueError):
        return -1


def bind_host_to_gpu_node(device_index, sysfs="/sys"):
    """One process per GPU on a multi-socket host: page-locked input buffers that live on the other socket reach
    the GPU through the inter-socket link, and with eight ranks reading one socket's memory the host side, not
    PCIe, bounds the host-to-device copies.  Call this first thing in a rank: it prefers the GPU's NUMA node for
    every later host allocation of this process (``set_mempolicy(MPOL_PREFERRED)``, so the pinned buffers land
    there) and moves the process onto that node's CPUs where the cpuset allows.  Best effort: returns a dict saying
    what was done (``node``, ``mempolicy``, ``cpus``), never raises.  ``unbind_host`` undoes it."""
    done = {"node": -1, "mempolicy": False, "cpus": 0, "prev_cpus": None}
    try:
        import torch
        p = torch.cuda.get_device_properties(device_index)
        node = gpu_numa_node(p.pci_domain_id, p.pci_bus_id, p.pci_device_id, sysfs)
    except Exception:
        return done
    done["node"] = node
    if node < 0:
        return done
    try:
        with open(os.path.join(sysfs, "devices/system/node/node%d/cpulist" % node)) as fh:
            local = _parse_cpulist(fh.read()) & os.sched_getaffinity(0)
        if local:
            done["prev_cpus"] = sorted(os.sched_getaffinity(0))
            os.sched_setaffinity(0, local)
            done["cpus"] = len(local)
    except (OSError, ValueError):
        pass
    nr = {"x86_64": 238, "aarch64": 237}.get(platform.machine())
    if nr is not None and node < 1024:
        try:
            libc = ctypes.CDLL(None, use_errno=True)
            mask = (ctypes.c_ulong * 16)()
            mask[node // 64] = 1 << (node % 64)
            MPOL_PREFERRED = 1
            done["mempolicy"] = libc.syscall(nr, MPOL_PREFERRED, mask, 1025) == 0
        except Exception:
            pass
    return done


def unbind_host(done):
    """Back to the default memory policy and the CPU set the process had before ``bind_host_to_gpu_node``."""
    try:
        if done.get("prev_cpus"):
            os.sched_setaffinity(0, set(done["prev_cpus"]))
        nr = {"x86_64": 238, "aarch64": 237}.get(platform.machine())
        if done.get("mempolicy") and nr is not None:
            ctypes.CDLL(None, use_errno=True).syscall(nr, 0, None, 0)      # MPOL_DEFAULT
    except Exception:
        pass
