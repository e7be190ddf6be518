"""2-GPU probe: NCCL all_reduce of the 62*P-float gradient bucket alone, and whether torch symmetric memory
(peer pointers, multicast pointer) is available on this box.  torchrun --nproc-per-node 2 tools/allreduce_probe.py"""
import os
import sys
import torch
import torch.distributed as dist

rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl")
n = 62 * 1000000
x = torch.randn(n, device="cuda")


def timeit(fn, iters=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


ms = timeit(lambda: dist.all_reduce(x))
if rank == 0:
    print("NCCL all_reduce %d MB: %.3f ms (%.0f GB/s algorithmic)" % (n * 4 / 1e6, ms, n * 4 / ms / 1e6), flush=True)
try:
    import torch.distributed._symmetric_memory as symm
    t = symm.empty(n, dtype=torch.float32, device="cuda")
    hdl = symm.rendezvous(t, dist.group.WORLD.group_name)
    if rank == 0:
        print("symm mem ok: buffer_ptrs", [hex(p) for p in hdl.buffer_ptrs], "multicast_ptr", hex(hdl.multicast_ptr),
              "signal pads", len(hdl.signal_pad_ptrs), [a for a in dir(hdl) if not a.startswith("_")], flush=True)
    t.copy_(x)
    for name in ("multimem_all_reduce_", "two_shot_all_reduce_", "one_shot_all_reduce"):
        try:
            op = getattr(torch.ops.symm_mem, name)
            ms = timeit(lambda: op(t, "sum", dist.group.WORLD.group_name))
            if rank == 0:
                print("symm_mem.%s: %.3f ms" % (name, ms), flush=True)
        except Exception as ex:  # noqa: BLE001
            if rank == 0:
                print("symm_mem.%s failed: %s" % (name, str(ex)[:200]), flush=True)
except Exception as ex:  # noqa: BLE001
    if rank == 0:
        print("symmetric memory unavailable:", str(ex)[:300], flush=True)
dist.destroy_process_group()
