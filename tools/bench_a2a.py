"""all_to_all_single bandwidth of the pressure solver's exchange (torchrun, one rank per GPU): 1.07 GB per rank."""
import os, torch, torch.distributed as dist
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
n = 512 ** 3
a = torch.rand(n, dtype=torch.float64, device="cuda"); b = torch.empty_like(a)
for parts in (1, 4):
    seg = n // parts
    for _ in range(2):
        for c in range(parts): dist.all_to_all_single(b[c * seg:(c + 1) * seg], a[c * seg:(c + 1) * seg])
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        for c in range(parts): dist.all_to_all_single(b[c * seg:(c + 1) * seg], a[c * seg:(c + 1) * seg])
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    sent = n * 8 * (world - 1) / world
    if rank == 0:
        print(f"world={world} parts={parts} NCHAN={os.environ.get('NCCL_MIN_P2P_NCHANNELS')} {ms:.3f} ms per exchange, {sent / ms / 1e6:.0f} GB/s sent per GPU", flush=True)
dist.destroy_process_group()
