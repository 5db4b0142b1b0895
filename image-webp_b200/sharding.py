"""Multi-GPU plumbing for the batched decode: images are independent, so the batch shards by
image index across ranks with NO data-path collective (reference: nothing is distributed; the
crate's decoders are independent objects, decoder.rs:312-331).  torch.distributed is used only
for the barrier and the max-over-ranks of the step time."""
import os


def env_rank_world():
    """(rank, local_rank, world_size) from the torchrun environment; (0, 0, 1) when not launched by it."""
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")))


def shard_range(n_total, rank, world):
    """Contiguous range of image indices for `rank`: sizes differ by at most one, earlier ranks get the extra."""
    base, extra = divmod(n_total, world)
    start = rank * base + min(rank, extra)
    return start, start + base + (1 if rank < extra else 0)


def weak_shard(per_gpu, rank):
    """Weak scaling: every rank decodes `per_gpu` images of its own (global indices)."""
    return rank * per_gpu, (rank + 1) * per_gpu


def init_distributed(backend=None):
    """Join the process group torchrun set up (nccl on GPUs, gloo on CPU).  No-op for one process."""
    rank, local_rank, world = env_rank_world()
    if world == 1:
        return None
    import torch
    import torch.distributed as dist
    if backend is None:
        backend = "nccl" if torch.cuda.is_available() else "gloo"
    if backend == "nccl":
        torch.cuda.set_device(local_rank)
    if not dist.is_initialized():
        dist.init_process_group(backend=backend, rank=rank, world_size=world)
    return dist


def barrier(dist):
    if dist is not None:
        dist.barrier()


def aggregate(dist, units_local, ms_local):
    """Whole-job figures: total units over all ranks and the MAX of the per-rank times."""
    if dist is None:
        return float(units_local), float(ms_local)
    import torch
    dev = "cuda" if dist.get_backend() == "nccl" else "cpu"
    t_units = torch.tensor([float(units_local)], dtype=torch.float64, device=dev)
    t_ms = torch.tensor([float(ms_local)], dtype=torch.float64, device=dev)
    dist.all_reduce(t_units, op=dist.ReduceOp.SUM)
    dist.all_reduce(t_ms, op=dist.ReduceOp.MAX)
    return float(t_units.item()), float(t_ms.item())


def all_min(dist, value):
    """Minimum of a per-rank number over all ranks (e.g. 1.0 / 0.0 for 'every rank's outputs were exact')."""
    if dist is None:
        return float(value)
    import torch
    dev = "cuda" if dist.get_backend() == "nccl" else "cpu"
    t = torch.tensor([float(value)], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    return float(t.item())


def bind_near_gpu(pci_bus_id):
    """Pin this process to the CPUs of the NUMA node the GPU hangs off, so that its threads and the
    pinned buffers it allocates afterwards are local to the GPU's PCIe root (with 8 ranks on two sockets
    an unbound rank sends half of its D2H traffic across the socket link).  Returns the CPU list used,
    or None when the topology is not exposed (single node, container without sysfs)."""
    import os
    try:
        with open("/sys/bus/pci/devices/%s/numa_node" % pci_bus_id) as f:
            node = int(f.read().strip())
        if node < 0:
            return None
        with open("/sys/devices/system/node/node%d/cpulist" % node) as f:
            spec = f.read().strip()
        cpus = set()
        for part in spec.split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return sorted(cpus)
    except (OSError, ValueError):
        return None
