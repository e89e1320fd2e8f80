#!/usr/bin/env python
"""Pinned host -> device copy rate with every GPU of the box uploading at once (the cap of bench.py's `e2e` figure).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P scripts/h2d_bench.py

Every rank copies the size of one bench step (65,536 channels x 9,600 int16 = 1.26 GB) from pinned host memory to its GPU,
`reps` times back to back, all ranks between the same two barriers; CUDA events time each rank's copies.  Two kinds of
pinned memory: cudaHostAllocDefault (what m17b_host_alloc gives) and cudaHostAllocWriteCombined.  Rank 0 prints one JSON line.
"""
import ctypes, json, os, sys, time
import torch
import torch.distributed as dist

def main():
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    nbytes = 65536 * 9600 * 2
    reps = 8
    rt = ctypes.CDLL("/usr/local/cuda/lib64/libcudart.so.12")
    dev = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    out = {}
    for name, flag in (("default", 0), ("write_combined", 4)):
        p = ctypes.c_void_p()
        assert rt.cudaHostAlloc(ctypes.byref(p), ctypes.c_size_t(nbytes), ctypes.c_uint(flag)) == 0
        ctypes.memset(p, 1, nbytes)                          # first touch on this rank's CPUs
        for alone in (True, False):
            torch.cuda.synchronize()
            if world > 1: dist.barrier()
            torch.cuda.synchronize()
            gbs = None
            if not alone or rank == 0:
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(reps):
                    assert rt.cudaMemcpyAsync(ctypes.c_void_p(dev.data_ptr()), p, ctypes.c_size_t(nbytes), 1, ctypes.c_void_p(0)) == 0
                e1.record(); torch.cuda.synchronize()
                gbs = reps * nbytes / (e0.elapsed_time(e1) * 1e-3) / 1e9
            if world > 1: dist.barrier()
            t = torch.tensor([gbs if gbs is not None else 0.0], dtype=torch.float64, device="cuda")
            if world > 1:
                lst = [torch.zeros_like(t) for _ in range(world)]
                dist.all_gather(lst, t)
                vals = [float(v.item()) for v in lst]
            else:
                vals = [float(t.item())]
            out[f"{name}_{'rank0_alone' if alone else 'all_ranks_at_once'}_gbs_per_gpu"] = [round(v, 2) for v in vals]
        rt.cudaFreeHost(p)
    if rank == 0:
        print(json.dumps({"what": "pinned host->device copy rate, GB/s per GPU", "n_gpus": world, "bytes_per_copy": nbytes, "reps": reps,
                          "host_cpus": os.cpu_count(), **out}), flush=True)
    if world > 1:
        dist.barrier(); dist.destroy_process_group()

main()
