import os, torch, torch.distributed as dist
local=int(os.environ["LOCAL_RANK"]); torch.cuda.set_device(local); dev=torch.device("cuda",local)
dist.init_process_group("nccl", device_id=dev)
for mb in (25, 100, 343):
    t=torch.ones(int(mb*1e6/4), device=dev)
    for _ in range(3): dist.all_reduce(t)
    torch.cuda.synchronize(); dist.barrier()
    e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): dist.all_reduce(t)
    e1.record(); torch.cuda.synchronize()
    if dist.get_rank()==0: print(f"all_reduce {mb} MB fp32: {e0.elapsed_time(e1)/10*1e3:.0f} us  ({mb/ (e0.elapsed_time(e1)/10):.0f} GB/s algbw)")
dist.destroy_process_group()
