"""Raw host↔device copy bandwidth with N processes at once, one per GPU (no library involved): what bounds bench.py's
e2e line at N GPUs.  Launch under torchrun:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/pcie_matrix.py

Every rank copies a 1 GiB pinned buffer to its GPU (and back) REPS times between two barriers; rank 0 prints the
per-rank and aggregate rates.  Variants: plain pinned memory (cudaHostAlloc default, what rpt_host_alloc gives),
write-combined pinned memory (H2D only), and the rank pinned to the CPUs NVML reports as local to its GPU."""
import ctypes
import json
import os
import time

import torch
import torch.distributed as dist

REPS = 8
BYTES = 1 << 30


def cudart():
    return ctypes.CDLL("libcudart.so.12") if os.path.exists("/usr/local/cuda/lib64/libcudart.so.12") else ctypes.CDLL(
        "libcudart.so")


def host_alloc(rt, nbytes, flags):
    p = ctypes.c_void_p()
    rc = rt.cudaHostAlloc(ctypes.byref(p), ctypes.c_size_t(nbytes), ctypes.c_uint(flags))
    assert rc == 0, rc
    return p


def main():
    world, rank, lr = int(os.environ["WORLD_SIZE"]), int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(lr)
    dev = torch.device("cuda", lr)
    dist.init_process_group("nccl", device_id=dev)
    rt = cudart()
    rt.cudaMemcpyAsync.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_void_p]
    d = torch.empty(BYTES, dtype=torch.uint8, device=dev)
    stream = torch.cuda.Stream(device=dev)
    results = {}
    affinity_note = None
    for variant in ("pinned", "write_combined", "pinned_cpu_affinity"):
        if variant == "pinned_cpu_affinity":
            try:
                import pynvml
                pynvml.nvmlInit()
                h = pynvml.nvmlDeviceGetHandleByIndex(lr)
                words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
                cpus = [64 * i + b for i, w in enumerate(words) for b in range(64) if (w >> b) & 1]
                os.sched_setaffinity(0, cpus)
                affinity_note = f"{len(cpus)} cpus ({cpus[0]}..{cpus[-1]})"
            except Exception as exc:  # noqa: BLE001
                affinity_note = f"unavailable: {exc}"
        flags = 4 if variant == "write_combined" else 0  # cudaHostAllocWriteCombined
        hp = host_alloc(rt, BYTES, flags)
        ctypes.memset(hp, 1, BYTES)  # touch: first-touch NUMA placement happens under the current affinity
        for direction in ("h2d", "d2h"):
            if variant == "write_combined" and direction == "d2h":
                continue
            kind = 1 if direction == "h2d" else 2
            dst, src = (d.data_ptr(), hp.value) if direction == "h2d" else (hp.value, d.data_ptr())
            rt.cudaMemcpyAsync(dst, src, BYTES, kind, ctypes.c_void_p(stream.cuda_stream))
            torch.cuda.synchronize()
            dist.barrier()
            t0 = time.perf_counter()
            for _ in range(REPS):
                rt.cudaMemcpyAsync(dst, src, BYTES, kind, ctypes.c_void_p(stream.cuda_stream))
            stream.synchronize()
            dt = time.perf_counter() - t0
            mine = torch.tensor([REPS * BYTES / dt / 1e9], dtype=torch.float64, device=dev)
            allr = [torch.zeros_like(mine) for _ in range(world)]
            dist.all_gather(allr, mine)
            dist.barrier()
            t_all = torch.tensor([dt], dtype=torch.float64, device=dev)
            dist.all_reduce(t_all, op=dist.ReduceOp.MAX)
            results[f"{variant}_{direction}"] = {"per_rank_gbs": [round(float(x.item()), 1) for x in allr],
                                                 "aggregate_gbs": round(world * REPS * BYTES / float(t_all.item()) / 1e9, 1)}
        rt.cudaFreeHost(hp)
    if rank == 0:
        print(json.dumps({"n_processes": world, "bytes_per_copy": BYTES, "reps": REPS, "host_cpus": os.cpu_count(),
                          "cpu_affinity_of_rank0": affinity_note, "results": results}), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
