# Probe (torchrun, N >= 2): how long does the 12.4 MB gradient all-reduce of config 3 take through NCCL and through
# PyTorch's symmetric-memory kernels (one-shot / two-shot over peer pointers, multimem = NVLS through the switch)?
import os, sys, torch, torch.distributed as dist
import torch.distributed._symmetric_memory as symm_mem
rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); local = int(os.environ.get("LOCAL_RANK", rank))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n = int(sys.argv[1]) if len(sys.argv) > 1 else 3_100_000
g = torch.randn(n, device="cuda")
def timeit(fn, iters=30):
    for _ in range(5): fn()
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3
res = {"nccl_us": timeit(lambda: dist.all_reduce(g))}
try:
    t = symm_mem.empty(n, dtype=torch.float32, device="cuda")
    hdl = symm_mem.rendezvous(t, dist.group.WORLD.group_name)
    res["multicast"] = bool(hdl.has_multicast_support) if hasattr(hdl, "has_multicast_support") else None
    res["mc_ptr"] = hex(hdl.multicast_ptr)
    t.copy_(g)
    name = dist.group.WORLD.group_name
    for op in ("one_shot_all_reduce", "two_shot_all_reduce_", "multimem_one_shot_all_reduce", "multimem_all_reduce_"):
        try:
            f = getattr(torch.ops.symm_mem, op)
            res[op + "_us"] = timeit(lambda: f(t, "sum", name))
        except Exception as e:
            res[op] = "ERR " + str(e)[:120]
    # correctness of the multimem path against NCCL on fresh data
    a = torch.arange(n, device="cuda", dtype=torch.float32) * (rank + 1) * 1e-3
    t.copy_(a); b = a.clone(); dist.all_reduce(b)
    try:
        torch.ops.symm_mem.multimem_all_reduce_(t, "sum", name); torch.cuda.synchronize()
        res["multimem_max_abs_diff"] = float((t - b).abs().max())
    except Exception as e:
        res["multimem_check"] = "ERR " + str(e)[:120]
except Exception as e:
    res["symm_mem"] = "ERR " + repr(e)[:300]
if rank == 0: print("PROBE", world, n, res, flush=True)
dist.barrier(); dist.destroy_process_group()
