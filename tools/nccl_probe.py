"""torchrun --nproc-per-node N tools/nccl_probe.py : all-reduce latency / bus bandwidth at the message sizes of one
data-parallel train step (BN statistics, metric sums, the two gradient buckets)."""
import os, torch, torch.distributed as dist
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
sizes = [("bn stats f64 x128", 128, torch.float64), ("metrics f64 x16", 16, torch.float64),
         ("attn colsum f32 x5175", 5175, torch.float32), ("head grads 1.4M f32", 1417000, torch.float32),
         ("decoder bucket 6.2M f32", 6200000, torch.float32), ("encoder bucket 9.3M f32", 9336000, torch.float32),
         ("encoder bucket as bf16", 9336000, torch.bfloat16), ("all grads 15.5M f32", 15500000, torch.float32)]
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for name, n, dt in sizes:
    t = torch.ones(n, dtype=dt, device="cuda")
    for _ in range(5):
        dist.all_reduce(t)
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    reps = 20
    e0.record()
    for _ in range(reps):
        dist.all_reduce(t)
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / reps
    nbytes = n * t.element_size()
    busbw = nbytes * 2 * (world - 1) / world / (us * 1e-6) / 1e9
    if rank == 0:
        print(f"N={world} {name:28s} {nbytes/1e6:9.3f} MB  {us:9.1f} us  busbw {busbw:7.1f} GB/s", flush=True)
dist.destroy_process_group()
