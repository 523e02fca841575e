"""tools/exp_d2h.py: device -> pinned host copy rate of one 33 MB frame per rank while 1, 2, 4 ... WORLD ranks copy at the
same time, with the pinned buffer (a) wherever the allocating thread happened to run and (b) allocated and first touched
on the NUMA node of the rank's GPU.  Explains (or not) the collapse of the full-frame e2e figure at 8 ranks
(VERDICT r1 item 9).  Run under torchrun on the multi-GPU box; prints one table from rank 0."""
import os
import torch
import torch.distributed as dist

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))


def gpu_numa_node():
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local)
        bus = pynvml.nvmlDeviceGetPciInfo(h).busId
        bus = bus.decode() if isinstance(bus, bytes) else bus
        path = "/sys/bus/pci/devices/" + bus.lower()[-12:] + "/numa_node"
        return int(open(path).read()), bus
    except Exception as e:  # noqa: BLE001
        return -1, str(e)


def node_cpus(node):
    out = []
    for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
        a, _, b = part.partition("-")
        out += list(range(int(a), int(b or a) + 1))
    return out


NBYTES = 3840 * 2160 * 4
dev = torch.empty(NBYTES, dtype=torch.uint8, device="cuda")
node, bus = gpu_numa_node()
start_aff = sorted(os.sched_getaffinity(0))


def measure(host, active, reps=40):
    """GB/s of this rank while ranks < active copy concurrently"""
    st = torch.cuda.Stream()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    g = 0.0
    if rank < active:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(st):
            for _ in range(5):
                host.copy_(dev, non_blocking=True)
            e0.record(st)
            for _ in range(reps):
                host.copy_(dev, non_blocking=True)
            e1.record(st)
        torch.cuda.synchronize()
        g = NBYTES * reps / (e0.elapsed_time(e1) * 1e-3) / 1e9
    if world > 1:
        dist.barrier()
    return g


rows = []
for label in ("default placement", "GPU's NUMA node"):
    if label != "default placement":
        if node < 0:
            continue
        try:
            os.sched_setaffinity(0, node_cpus(node))
        except OSError as e:
            rows.append(f"rank {rank}: cannot bind to node {node}: {e}")
    host = torch.empty(NBYTES, dtype=torch.uint8).pin_memory()
    host.fill_(1)
    active = 1
    while active <= world:
        g = measure(host, active)
        t = torch.tensor([g], dtype=torch.float64, device="cuda")
        all_g = [torch.zeros_like(t) for _ in range(world)]
        if world > 1:
            dist.all_gather(all_g, t)
        else:
            all_g = [t]
        if rank == 0:
            vals = [x.item() for x in all_g[:active]]
            rows.append(f"{label:18s} ranks copying {active}: per-rank GB/s min {min(vals):6.1f} max {max(vals):6.1f} sum {sum(vals):7.1f}")
        active *= 2
    del host
info = torch.tensor([node], device="cuda")
nodes = [torch.zeros_like(info) for _ in range(world)]
if world > 1:
    dist.all_gather(nodes, info)
else:
    nodes = [info]
if rank == 0:
    online = open('/sys/devices/system/node/online').read().strip() if os.path.exists('/sys/devices/system/node/online') else '?'
    print(f"frame bytes {NBYTES}; GPU NUMA nodes by rank {[int(x.item()) for x in nodes]}; rank 0 start affinity {len(start_aff)} cpus; nodes online {online}")
    print("\n".join(rows), flush=True)
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
