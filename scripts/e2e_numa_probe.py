"""Why the end-to-end (host -> device -> host) leg does not scale with the number of ranks on one box.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29544 scripts/e2e_numa_probe.py

For k = 1, 2, 4, 8 concurrently active ranks: pinned-host -> device, device -> pinned-host and both at once, per rank,
for copy sizes 4 / 32 / 256 MiB, with the pinned buffers first-touched (a) wherever the rank runs and (b) after pinning
the rank to the CPUs `nvidia-smi topo` lists for its GPU (first-touch then places the pages on that NUMA node).  Also
records the topology (CPU affinity / NUMA node of every GPU, host NUMA nodes).  Rank 0 prints one JSON document."""
import json, os, re, subprocess, sys, time
import torch
import torch.distributed as dist

rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
local = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=dev)


def barrier():
    if world > 1:
        dist.barrier()


def gpu_affinity(idx):
    try:
        out = subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True, timeout=30).stdout
        for ln in out.splitlines():
            if ln.startswith(f"GPU{idx}\t") or ln.startswith(f"GPU{idx} "):
                cols = [c for c in re.split(r"\t+|\s{2,}", ln.strip()) if c]
                return cols[-3] if len(cols) >= 3 else None, out
        return None, out
    except Exception as e:                                   # noqa: BLE001
        return None, str(e)


def cpus_of(spec):
    cpus = set()
    for part in (spec or "").split(","):
        if "-" in part:
            a, b = part.split("-")
            cpus.update(range(int(a), int(b) + 1))
        elif part.strip().isdigit():
            cpus.add(int(part))
    return cpus


def run_case(active, nbytes, host_in, host_out, d_in, d_out, mode):
    reps = max(2, min(20, (1 << 30) // nbytes))
    n = nbytes // 4
    s1, s2 = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
    barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    if active:
        for _ in range(reps):
            if mode in ("h2d", "both"):
                with torch.cuda.stream(s1):
                    d_in[:n].copy_(host_in[:n], non_blocking=True)
            if mode in ("d2h", "both"):
                with torch.cuda.stream(s2):
                    host_out[:n].copy_(d_out[:n], non_blocking=True)
        torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    barrier()
    return (reps * nbytes / dt / 1e9) if active else 0.0


affinity, topo = gpu_affinity(local)
results = []
MAXB = 256 << 20
d_in = torch.empty(MAXB // 4, device=dev); d_out = torch.ones(MAXB // 4, device=dev)
for placement in ("default", "gpu_affinity"):
    if placement == "gpu_affinity":
        cp = cpus_of(affinity) & set(os.sched_getaffinity(0))
        if cp:
            os.sched_setaffinity(0, cp)
    host_in = torch.empty(MAXB // 4, pin_memory=True); host_in.fill_(1.0)       # first touch here
    host_out = torch.empty(MAXB // 4, pin_memory=True); host_out.fill_(0.0)
    for k in (1, 2, 4, 8):
        if k > world:
            continue
        for nbytes in (4 << 20, 32 << 20, 256 << 20):
            for mode in ("h2d", "d2h", "both"):
                gbs = run_case(rank < k, nbytes, host_in, host_out, d_in, d_out, mode)
                t = torch.tensor([gbs], device=dev, dtype=torch.float64)
                allv = [torch.zeros_like(t) for _ in range(world)] if world > 1 else [t]
                if world > 1:
                    dist.all_gather(allv, t)
                if rank == 0:
                    per = [float(v) for v in allv[:k]]
                    results.append({"placement": placement, "active_ranks": k, "copy_MiB": nbytes >> 20, "mode": mode,
                                    "per_rank_GBs_min": min(per), "per_rank_GBs_max": max(per), "aggregate_GBs": sum(per) * (2 if mode == "both" else 1)})
    del host_in, host_out
if rank == 0:
    numa = subprocess.run("lscpu | grep -i -E 'numa|^CPU\\(s\\)|model name|socket'", shell=True, capture_output=True, text=True).stdout
    print(json.dumps({"world": world, "gpu0_cpu_affinity": affinity, "lscpu": numa.strip().splitlines(), "topology": topo.splitlines()[:14],
                      "note": "aggregate for mode 'both' counts both directions; per-rank figures are per direction", "results": results}))
if world > 1:
    dist.destroy_process_group()
