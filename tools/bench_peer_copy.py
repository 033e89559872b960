"""NVLink transfer rates between two ranks of one box, as the y-slab solver uses them (run under torchrun, 2+ ranks):
copy-engine copies into the IPC-mapped arena of the next rank (torch copy_ on a wrapped peer pointer = cudaMemcpyAsync) and
SM stores (a torch elementwise kernel writing the peer tensor), for the chunk sizes of the 512^3-per-GPU solve."""
import ctypes as C, os, sys
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, ".")
from microhh_b200 import host as H, lib as L

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl", device_id=dev)
z = (np.arange(8) + 0.5)
g = H.Grid(64, 64 * world, 8, 1., 1., 8., z, npy=world, rank_y=rank, device=dev)
nbytes = 1 << 30
g.arena = H.PeerArena(g, nbytes + (1 << 20))
buf = g.arena.alloc((nbytes // 8,))

class Comm: pass
comm = Comm(); comm.world, comm.rank, comm.group = world, rank, None
g.arena.connect(comm)
peers = g.arena.peer_ptrs(buf)
nxt = (rank + 1) % world
dst = torch.as_tensor(H._RawCuda(peers[nxt], nbytes), device=dev).view(torch.float64)
print(rank, "peer tensor device", dst.device, flush=True)
src = torch.randn(nbytes // 8, dtype=torch.float64, device=dev)
local = torch.empty_like(src)

def timeit(fn, n=5):
    for _ in range(2): fn()
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n

for mb in (64, 256, 512, 1024):
    n = mb * (1 << 20) // 8
    t_ce = timeit(lambda: dst[:n].copy_(src[:n], non_blocking=True))
    t_sm = timeit(lambda: torch.add(src[:n], 1.0, out=dst[:n]))
    t_loc = timeit(lambda: local[:n].copy_(src[:n], non_blocking=True))
    if rank == 0:
        print(f"{mb:5d} MiB: copy engine -> peer {mb / 1024 * 1.0737 / t_ce * 1e3:7.1f} GB/s ({t_ce:.3f} ms) | SM stores -> peer {mb / 1024 * 1.0737 / t_sm * 1e3:7.1f} GB/s ({t_sm:.3f} ms)"
              f" | local D2D copy {mb / 1024 * 1.0737 / t_loc * 1e3:7.1f} GB/s ({t_loc:.3f} ms)", flush=True)
dist.barrier()
dist.destroy_process_group()
