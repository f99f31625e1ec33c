"""Host link microbenchmark: the copy pattern of one e2e step of bench.py (64 frames of 752x480 up from pinned host memory
on one stream, keypoints + descriptors down on another) with no kernel in between, on 1 GPU or on every rank of a
`python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/ubench_link.py` launch (all ranks at
once, max over ranks = the N-rank ceiling of the host entry point). bench.py measures the same thing through the e2e
leg's own pinned buffers and prints it as e2e.h2d_ceiling_frames_per_s; `--from-bench FILE...` rewrites those lines as
profiles/r2_link_{N}gpu.json."""
import json
import os
import sys
import time

import torch


def from_bench(paths):
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for p in paths:
        d = json.loads([l for l in open(p) if l.startswith("{")][-1])
        e = d["e2e"]
        out = {"n_gpus": d["n_gpus"], "source": os.path.relpath(os.path.abspath(p), root) + " (e2e leg of bench.py)",
               "h2d_bytes_per_step": e["h2d_bytes_per_step"], "d2h_bytes_per_step": e["d2h_bytes_per_step"],
               "copy_only_frames_per_s": e["h2d_ceiling_frames_per_s"], "e2e_frames_per_s": e["value"],
               "frac_of_link_ceiling": e["frac_of_link_ceiling"],
               "aggregate_host_traffic_GBps": (e["h2d_bytes_per_step"] + e["d2h_bytes_per_step"]) * e["h2d_ceiling_frames_per_s"] / 64 / 1e9,
               "limiter": "cudaMemcpyAsync from / to pinned host memory (PCIe + the VM's host memory system); no collective on this path"}
        with open(os.path.join(root, "profiles", "r2_link_%dgpu.json" % d["n_gpus"]), "w") as f:
            json.dump(out, f, indent=1)
        print(out)


def main():
    if len(sys.argv) > 1 and sys.argv[1] == "--from-bench":
        return from_bench(sys.argv[2:])
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank, local = int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    B, h, w, cap = 64, 480, 752, 1104
    pin = [torch.empty((B, h, w), dtype=torch.uint8).pin_memory() for _ in range(2)]
    okp = [torch.empty((B, cap, 6), dtype=torch.float32).pin_memory() for _ in range(2)]
    ode = [torch.empty((B, cap, 32), dtype=torch.uint8).pin_memory() for _ in range(2)]
    d_in = torch.empty((B, h, w), dtype=torch.uint8, device=dev)
    d_kp = torch.zeros((B, cap, 6), dtype=torch.float32, device=dev)
    d_de = torch.zeros((B, cap, 32), dtype=torch.uint8, device=dev)
    s_in, s_out = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)

    def run(k):
        for i in range(k):
            with torch.cuda.stream(s_in):
                d_in.copy_(pin[i & 1], non_blocking=True)
            with torch.cuda.stream(s_out):
                okp[i & 1].copy_(d_kp, non_blocking=True); ode[i & 1].copy_(d_de, non_blocking=True)
        torch.cuda.synchronize()

    run(5)
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    run(100)
    ms = torch.tensor([1e3 * (time.perf_counter() - t0) / 100], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    if rank == 0:
        up, down = pin[0].numel(), okp[0].numel() * 4 + ode[0].numel()
        print(json.dumps({"n_gpus": world, "ms_per_step_max_over_ranks": float(ms), "copy_only_frames_per_s": B * world / (float(ms) * 1e-3),
                          "h2d_bytes_per_step": up, "d2h_bytes_per_step": down,
                          "aggregate_host_traffic_GBps": (up + down) * world / (float(ms) * 1e-3) / 1e9}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
