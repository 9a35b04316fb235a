"""Probe: can a copy engine write to the NVSwitch multicast mapping of a symmetric buffer (one copy -> every rank)?"""
import ctypes, os, torch, torch.distributed as dist
import torch.distributed._symmetric_memory as symm
rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); lr = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr); dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
rt = ctypes.CDLL("libcudart.so.12")
rt.cudaMemcpyAsync.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_void_p]
n_local, row = 72, 8 * 128 * 1723
buf = symm.empty((world * n_local, row), dtype=torch.float32, device=dev)
hdl = symm.rendezvous(buf, dist.group.WORLD.group_name)
src = torch.full((n_local, row), float(rank + 1), device=dev)     # staging outside the symmetric buffer
buf.zero_()
torch.cuda.synchronize(); dist.barrier()
mc = hdl.multicast_ptr
print(rank, "multicast_ptr", hex(mc), flush=True)
comm = torch.cuda.Stream()
nbytes = src.numel() * 4
def push():
    with torch.cuda.stream(comm):
        rc = rt.cudaMemcpyAsync(ctypes.c_void_p(mc + rank * nbytes), ctypes.c_void_p(src.data_ptr()), nbytes, 3,
                                ctypes.c_void_p(comm.cuda_stream))
        assert rc == 0, rc
        hdl.barrier(channel=0)
push(); push()
torch.cuda.synchronize(); dist.barrier()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
with torch.cuda.stream(comm): e0.record()
for _ in range(5): push()
with torch.cuda.stream(comm): e1.record()
torch.cuda.synchronize(); dist.barrier()
ok = all(bool((buf[r * n_local:(r + 1) * n_local] == r + 1).all()) for r in range(world))
ms = e0.elapsed_time(e1) / 5
print(rank, "ok" if ok else "MISMATCH", f"{ms:.3f} ms per multicast gather ({nbytes/ms/1e6:.1f} GB/s out, {nbytes*(world-1)/ms/1e6:.1f} GB/s in)", flush=True)
dist.destroy_process_group()
