import os, time, torch, torch.distributed as dist
rank=int(os.environ["RANK"]); local=int(os.environ["LOCAL_RANK"]); world=int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
for mb in (2.5, 20):
    n=int(mb*1e6/4)
    inp=torch.ones(n,dtype=torch.int32,device="cuda"); out=torch.empty(n*world,dtype=torch.int32,device="cuda")
    for _ in range(5): dist.all_gather_into_tensor(out,inp)
    torch.cuda.synchronize(); t0=time.perf_counter()
    for _ in range(20): dist.all_gather_into_tensor(out,inp)
    torch.cuda.synchronize(); dt=(time.perf_counter()-t0)/20
    if rank==0: print("allgather %.1f MB per rank x %d: %.3f ms" % (mb, world, dt*1e3), flush=True)
dist.destroy_process_group()
