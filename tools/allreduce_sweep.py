"""Gradient all-reduce bandwidth sweep: our kernels vs NCCL, 1 KB - 1 GB.

Launch with torch.distributed.run (one rank per GPU).  For each size: in-place all-reduce on
a symmetric bucket (NVLS and P2P two-shot; one-shot for small), CUDA-event timed, max over
ranks; the same tensor size through ``torch.distributed.all_reduce`` (NCCL).  Prints a table
and writes JSON lines to gpurun_out/allreduce_sweep_<N>gpu.jsonl.
busbw = algbw * 2(n-1)/n; roofline fraction is busbw / 770 GB/s (measured peer copy, B200_PROFILING.md).
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as tdist

from torch_distributed_sandbox_b200 import distributed as dist


def timeit(fn, iters, sync):
    for _ in range(3):
        fn()
    sync()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters):
        fn()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / iters


def main():
    rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    tdist.init_process_group("nccl", device_id=dev)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    be = dist._default().backend
    sizes = [1 << k for k in range(10, 31, 2)] + [53_376, 720_000_000]
    maxn = max(sizes) // 4
    sym = be.alloc_symmetric(maxn, torch.float32)
    plain = torch.zeros(maxn, device=dev)
    rows = []

    def sync():
        torch.cuda.synchronize()
        dist.barrier()

    def mx(v):
        t = torch.tensor([v], device=dev)
        tdist.all_reduce(t, op=tdist.ReduceOp.MAX)
        return float(t)

    for nbytes in sorted(sizes):
        n = nbytes // 4 // 4 * 4
        assert n * 4 == nbytes or nbytes == 53_376, nbytes
        iters = 200 if nbytes <= (1 << 20) else (30 if nbytes <= (1 << 26) else 8)
        row = {"bytes": n * 4, "world": world}
        view = sym[:n]
        algos = ["oneshot"] if n * 4 <= be.oneshot_max_bytes else ((["nvls"] if be.multicast else []) + ["p2p"])
        for algo in algos:
            blocks_list = [64] if algo == "oneshot" else [32, 64, 128]
            best = None
            for blocks in blocks_list:
                if algo == "oneshot":
                    f = lambda: be.all_reduce(plain[:n], dist.ReduceOp.SUM)
                else:
                    f = lambda: be.all_reduce(view, dist.ReduceOp.SUM, algo=algo, blocks=blocks)
                ms = mx(timeit(f, iters, sync))
                if best is None or ms < best[0]:
                    best = (ms, blocks)
            row[f"ours_{algo}_ms"], row[f"ours_{algo}_blocks"] = best
        if n * 4 >= 700_000_000 and n % 10 == 0:
            # the DDP bucket itself: one-pass tile kernel (reduce + 1/world + SGD + weight push), counters pre-satisfied
            if "w" not in globals():
                globals()["w"] = be.alloc_symmetric(n, torch.float32)
                globals()["rdy"] = be.alloc_symmetric(256, torch.int32, multicast=False)
                globals()["rdy"].fill_(1 << 30)
                sync()
            from torch_distributed_sandbox_b200 import ops
            slab = ops.fc_bwd_tile_slab_floats(8)
            for blocks in (8, 16, 32):
                f = lambda: be.all_reduce_tiles(view.view(10, -1), w.view(10, -1), rdy, rows=10, slab_floats=slab, target=1,
                                                scale=1.0 / world, lr=1e-4, channel=0, blocks=blocks)
                row[f"ours_onepass_fused_sgd_{blocks}ctas_ms"] = mx(timeit(f, iters, sync))
        f = lambda: tdist.all_reduce(plain[:n])
        row["nccl_ms"] = mx(timeit(f, iters, sync))
        ours = min(v for k, v in row.items() if k.startswith("ours_") and k.endswith("_ms") and "onepass" not in k)
        fac = 2 * (world - 1) / world
        row["ours_busbw_gbs"] = n * 4 / ours / 1e6 * fac
        row["nccl_busbw_gbs"] = n * 4 / row["nccl_ms"] / 1e6 * fac
        row["ours_frac_of_770"] = row["ours_busbw_gbs"] / 770.0
        rows.append(row)
        if rank == 0:
            print(json.dumps(row), flush=True)
    if rank == 0:
        os.makedirs("gpurun_out", exist_ok=True)
        with open(f"gpurun_out/allreduce_sweep_{world}gpu.jsonl", "w") as f:
            for r in rows:
                f.write(json.dumps(r) + "\n")
    dist.destroy_process_group()
    tdist.destroy_process_group()


if __name__ == "__main__":
    main()
