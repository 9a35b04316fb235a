"""Probe: torch symmetric memory + copy-engine pushes as an all-gather that needs no SMs (run under torchrun)."""
import os, time, torch, torch.distributed as dist
import torch.distributed._symmetric_memory as symm
rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); lr = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr); dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
n_local, row = 72, 8 * 128 * 1723
buf = symm.empty((world * n_local, row), dtype=torch.float32, device=dev)
hdl = symm.rendezvous(buf, dist.group.WORLD.group_name)
print(rank, "rendezvous ok; multicast:", hdl.multicast_ptr != 0, flush=True)
mine = buf[rank * n_local:(rank + 1) * n_local]
mine.fill_(float(rank + 1))
torch.cuda.synchronize(); dist.barrier()
peers = [hdl.get_buffer(r, (world * n_local, row), torch.float32) for r in range(world)]
comm = torch.cuda.Stream()
def push():
    with torch.cuda.stream(comm):
        for k in range(1, world):
            r = (rank + k) % world
            peers[r][rank * n_local:(rank + 1) * n_local].copy_(mine, non_blocking=True)
        hdl.barrier(channel=0)
for _ in range(2): push()
torch.cuda.synchronize(); dist.barrier()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
with torch.cuda.stream(comm):
    e0.record()
for _ in range(5): push()
with torch.cuda.stream(comm):
    e1.record()
torch.cuda.synchronize(); dist.barrier()
ok = all(bool((buf[r * n_local:(r + 1) * n_local] == r + 1).all()) for r in range(world))
ms = e0.elapsed_time(e1) / 5
print(rank, "ok" if ok else "MISMATCH", f"{ms:.3f} ms per gather, {mine.numel()*4*(world-1)/ms/1e6:.1f} GB/s out", flush=True)
dist.destroy_process_group()
