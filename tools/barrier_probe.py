"""Tuning aid: latency of tiny NCCL collectives used as stream-ordered barriers (run under torch.distributed.run)."""
import os
import torch
import torch.distributed as dist

lr = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
s = torch.cuda.Stream()
torch.cuda.set_stream(s)
tok = torch.zeros(1, dtype=torch.int32, device="cuda")
tok8 = torch.zeros(dist.get_world_size(), dtype=torch.int32, device="cuda")
big = torch.zeros(64 << 20, dtype=torch.uint8, device="cuda")
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for name, fn in [("all_reduce 4B", lambda: dist.all_reduce(tok)), ("all_gather 4B", lambda: dist.all_gather_into_tensor(tok8, tok)),
                 ("all_reduce 4B after 20us kernel", lambda: (big[: 8 << 20].zero_(), dist.all_reduce(tok))),
                 ("all_reduce 4B after 2.7ms of kernels", lambda: ([big.zero_() for _ in range(90)], dist.all_reduce(tok)))]:
    for _ in range(5):
        fn()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    e0.record()
    for _ in range(50):
        fn()
    e1.record(); torch.cuda.synchronize()
    if dist.get_rank() == 0:
        print(f"{name:40s} {e0.elapsed_time(e1) / 50 * 1000:9.1f} us per iteration")
dist.destroy_process_group()
