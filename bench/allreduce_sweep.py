"""All-reduce bus-bandwidth sweep over the VGG-11 bucket / tensor sizes (BASELINE.json config 5,
SURVEY.md S4): our peer-memory kernels (P2P and NVLS multicast) against NCCL's all_reduce on the
same GPUs.  Launch with torchrun.  Times with CUDA events on every rank, takes the max over
ranks, reports algorithm time, bus bandwidth (2(W-1)/W * bytes / time) and the fraction of the
measured 770 GB/s per-direction NVLink peer-copy rate (B200_PROFILING.md).
"""
import json
import os
import sys
from pathlib import Path

import torch
import torch.distributed as dist

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from ddpb200.parallel.comm import SymmetricBuffer  # noqa: E402

SIZES = {  # bytes (fp32 gradients)
    "fc1.bias (40 B)": 40, "bn vector (2 KB)": 2048, "bucket 2 (0.30 MB)": 76032 * 4,
    "conv 512x512x3x3 (9.4 MB)": 2359296 * 4, "bucket 0 (9.46 MB)": 2365962 * 4,
    "bucket 1 (27.2 MB)": 6789120 * 4, "all gradients (36.9 MB)": 9231114 * 4,
}
NVLINK_GBS = 770.0


def main():
    rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    max_floats = (max(SIZES.values()) // 4 + 63) // 64 * 64
    symm = SymmetricBuffer(max_floats, dev)
    rows = []
    nccl_buf = torch.randn(max_floats, device=dev)

    def timed(fn, iters=30, warm=5):
        for _ in range(warm):
            fn()
        torch.cuda.synchronize()
        dist.barrier()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(iters):
            fn()
        e.record()
        torch.cuda.synchronize()
        t = torch.tensor([s.elapsed_time(e) / iters * 1e3], device=dev)   # us
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    for name, nbytes in SIZES.items():
        n = (nbytes // 4 + 3) // 4 * 4
        row = {"what": name, "bytes": nbytes}
        for blocks in (8, 32, 128):
            for mc in ([False, True] if symm.has_multicast else [False]):
                key = f"ours_{'nvls' if mc else 'p2p'}_{blocks}cta_us"
                row[key] = timed(lambda: symm.allreduce_avg(0, n, blocks=blocks, use_mc=mc))
        row["nccl_us"] = timed(lambda: dist.all_reduce(nccl_buf[:n]))
        best = min(v for k, v in row.items() if k.startswith("ours_"))
        row["ours_best_us"] = best
        bus = 2.0 * (world - 1) / world * nbytes
        row["ours_busbw_gbs"] = bus / best / 1e3
        row["nccl_busbw_gbs"] = bus / row["nccl_us"] / 1e3
        row["ours_frac_of_nvlink_770"] = row["ours_busbw_gbs"] / NVLINK_GBS
        # Rooflines (bytes that must cross one GPU's NVLink per direction / 770 GB/s):
        #   two-shot over peer memory: 2 (W-1)/W S;   NVLS (ld_reduce of the own slice + multicast store): S + S/W
        #   (SURVEY.md par. 5.8) -- the NVLS kernel is scored against ITS OWN, lower, roof
        roof_p2p_us = bus / NVLINK_GBS / 1e3
        roof_nvls_us = (nbytes + nbytes / world) / NVLINK_GBS / 1e3
        nvls = [v for k, v in row.items() if k.startswith("ours_nvls")]
        p2p = [v for k, v in row.items() if k.startswith("ours_p2p")]
        row["roof_p2p_two_shot_us"], row["roof_nvls_us"] = roof_p2p_us, roof_nvls_us
        if nvls:
            row["ours_nvls_frac_of_nvls_roof"] = roof_nvls_us / min(nvls)
        row["ours_p2p_frac_of_p2p_roof"] = roof_p2p_us / min(p2p)
        # the batched multi-range kernel on the reference's per-parameter shape is timed once below
        rows.append(row)
    # the reference's part2b shape: 34 per-parameter all-reduces (32 of them <= 2 KB) -- one launch per tensor
    # versus the batched table kernel (one launch, small tensors one-shot)
    import ddpb200
    from ddpb200.engine import ALIGN

    sizes = [p.numel() for p in ddpb200.VGG11().parameters()]
    ranges, off = [], 0
    for nel in sizes:
        ranges.append((off, (nel + 3) // 4 * 4))
        off += (nel + ALIGN - 1) // ALIGN * ALIGN
    mc = symm.has_multicast
    per_param = {
        "what": "34 parameter tensors of VGG-11 (part2b shape)",
        "one_launch_per_tensor_us": timed(lambda: [symm.allreduce_avg(o, n, blocks=32, use_mc=mc) for o, n in ranges], iters=10),
        "batched_table_kernel_us": timed(lambda: symm.allreduce_multi(ranges, blocks=128, use_mc=mc), iters=30),
        "one_fused_range_us": timed(lambda: symm.allreduce_avg(0, off, blocks=128 if not mc else 32, use_mc=mc), iters=30),
        "nccl_34_calls_us": timed(lambda: [dist.all_reduce(nccl_buf[o:o + n]) for o, n in ranges], iters=10),
    }
    if rank == 0:
        out = {"world": world, "backend": symm.backend, "multicast": symm.has_multicast, "rows": rows,
               "per_parameter": per_param}
        print(json.dumps(out, indent=1))
        p = Path(os.environ.get("DDPB_SWEEP_OUT", "gpurun_out/allreduce_sweep.json"))
        p.parent.mkdir(parents=True, exist_ok=True)
        p.write_text(json.dumps(out, indent=1))
    symm.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
