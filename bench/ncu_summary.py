"""profiles/r2_ncu_summary.txt from the raw-metric CSV of `ncu --set full` over bench/ncu_target.py (default
list, DDPB_NCU_REPS=1): per kernel the duration, tensor-pipe and DRAM utilisation, DRAM bytes against the
algorithmic bytes, and the roofline fraction against MEASURED_PEAKS.json (cuBLAS bf16 burst rate for a kernel
timed alone, copy bandwidth).  ncu serialises and cold-starts every launch: read shares, not absolutes;
the in-graph times are in profiles/r2_call43_layer_times_*.

    python bench/ncu_summary.py profiles/r2_call46_ncu_raw_metrics.csv > profiles/r2_ncu_summary.txt
"""
import csv
import json
import re
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
peaks = json.loads((ROOT / "MEASURED_PEAKS.json").read_text()) if (ROOT / "MEASURED_PEAKS.json").exists() else {}
TF = peaks.get("bf16_tflops", 1590.0) * 1e12
BW = peaks.get("hbm_gbs", 6650.0) * 1e9
N = 256
LAYERS = [(32, 3, 64, True), (16, 64, 128, True), (8, 128, 256, False), (8, 256, 256, True),
          (4, 256, 512, False), (4, 512, 512, True), (2, 512, 512, False), (2, 512, 512, True)]


def conv(li, kind):
    S, Ci, Co, pool = LAYERS[li]
    P = N * S * S
    k = 64 if li == 0 else 9 * Ci
    flops = 2.0 * P * Co * k
    cin = 64 if li == 0 else Ci
    if kind == "fwd":     # x + w in, conv_out out, conv_out back in + act out (BatchNorm tail)
        by = P * cin * 2 + Co * k * 2 + 2 * P * Co * 2 + (P // 4 if pool else P) * Co * 2
    elif kind == "dgrad":
        by = P * Co * 2 + Co * k * 2 + P * Ci * 2
    else:
        by = P * Co * 2 + P * cin * 2 + Co * k * 4
    return flops, by


def bn(li, passes_x, passes_dy, writes_dx):
    S, Ci, Co, pool = LAYERS[li]
    P = N * S * S
    Po = P // 4 if pool else P
    return 0.0, passes_x * P * Co * 2 + passes_dy * Po * Co * 2 + writes_dx * P * Co * 2


# (label, flops, algorithmic bytes) in launch order of ncu_target.py's default list
ROWS = [("input kernel (crop/flip/normalise -> layer-0 operand)", 0, N * 32 * 32 * 3 + N * 32 * 32 * 64 * 2),
        ("fwd L0 conv+BN (persistent, 64-wide)",) + conv(0, "fwd"), ("fwd L1 conv+BN (128-wide)",) + conv(1, "fwd"),
        ("fwd L3 conv+BN (256-wide)",) + conv(3, "fwd"), ("fwd L5 conv+BN (128-wide)",) + conv(5, "fwd"),
        ("fwd L7 conv+BN, split-K x4 summed in the kernel",) + conv(7, "fwd")[:1] +
        (N * 4 * 512 * 2 + 512 * 4608 * 2 + 2 * 3 * N * 4 * 512 * 4 + 2 * N * 4 * 512 * 2 * 1.25,),
        ("dgrad L1 (64-wide)",) + conv(1, "dgrad"), ("dgrad L3 (256-wide)",) + conv(3, "dgrad"),
        ("dgrad L5 (128-wide)",) + conv(5, "dgrad"),
        ("dgrad L7 split-K x4",) + conv(7, "dgrad")[:1] + (N * 4 * 512 * 2 + 512 * 4608 * 2 + 4 * N * 4 * 512 * 4,),
        ("dgrad L7 split-K finalize", 0, 5 * N * 4 * 512 * 4),
        ("wgrad L0 (256 pixel splits)",) + conv(0, "wgrad")[:1] + (N * 1024 * 64 * 4 + 256 * 128 * 64 * 4,),
        ("wgrad L0 split reduction", 0, 256 * 64 * 64 * 4),
        ("wgrad L3 (8 pixel splits)",) + conv(3, "wgrad")[:1] + (2 * N * 64 * 256 * 2 + 8 * 256 * 2304 * 4,),
        ("wgrad L3 split reduction", 0, 9 * 256 * 2304 * 4), ("wgrad L5 (direct)",) + conv(5, "wgrad"),
        ("wgrad L6 (direct)",) + conv(6, "wgrad"),
        ("(test set-up: BatchNorm forward L0)",) + bn(0, 1, 1, 0), ("BN bwd L0 reduce",) + bn(0, 1, 1, 0),
        ("BN bwd L0 apply",) + bn(0, 1, 1, 1),
        ("(test set-up: BatchNorm forward L1)",) + bn(1, 1, 1, 0), ("BN bwd L1 reduce",) + bn(1, 1, 1, 0),
        ("BN bwd L1 apply",) + bn(1, 1, 1, 1),
        ("(test set-up: BatchNorm forward L3)",) + bn(3, 1, 1, 0), ("BN bwd L3 reduce",) + bn(3, 1, 1, 0),
        ("BN bwd L3 apply",) + bn(3, 1, 1, 1),
        ("(test set-up: BatchNorm forward L6)",) + bn(6, 1, 1, 0), ("BN bwd L6 (channel-owner, one launch)",) + bn(6, 1, 1, 1),
        ("classifier + cross-entropy forward", 2.0 * N * 512 * 10, N * 512 * 2 + 5120 * 4),
        ("classifier backward (+ loss reduction)", 4.0 * N * 512 * 10, 2 * N * 512 * 2 + 2 * 5120 * 4),
        ("SGD momentum + weight decay + bf16 shadow (9.2 M parameters)", 0, 9_233_024 * (3 * 4 + 2 * 4 + 2))]


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    hdr = rows[0]
    ix = {k: hdr.index(k) for k in ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
                                    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
                                    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
                                    "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
                                    "launch__grid_size", "smsp__issue_active.avg.pct_of_peak_sustained_active"]}
    data = rows[2:]
    assert len(data) == len(ROWS), (len(data), len(ROWS))
    print(f"ncu --set full, one cold launch per kernel, batch {N}; peaks: {TF / 1e12:.0f} TFLOP/s bf16 (cuBLAS burst), "
          f"{BW / 1e9:.0f} GB/s copy ({'measured' if peaks else 'fallback'})")
    print(f"{'what':58s} {'kernel':34s} {'us':>6s} {'tens%':>5s} {'dram%':>5s} {'issue%':>6s} {'warps%':>6s} {'regs':>4s} {'grid':>5s} "
          f"{'DRAM MB':>8s} {'algo MB':>8s} {'roof us':>7s} {'frac':>5s}")
    for r, (label, flops, by) in zip(data, ROWS):
        f = lambda k: float(r[ix[k]].replace(",", ""))  # noqa: E731
        name = re.sub(r"\(.*", "", r[ix["Kernel Name"]]).replace("ddpb::", "").replace("void ", "")[:34]
        us = f("gpu__time_duration.sum")
        dram = f("dram__bytes_read.sum") + f("dram__bytes_write.sum")
        roof = max(flops / TF, by / BW) * 1e6
        print(f"{label:58s} {name:34s} {us:6.1f} {f('sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active'):5.1f} "
              f"{f('gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed'):5.1f} "
              f"{f('smsp__issue_active.avg.pct_of_peak_sustained_active'):6.1f} "
              f"{f('sm__warps_active.avg.pct_of_peak_sustained_active'):6.1f} {int(f('launch__registers_per_thread')):4d} "
              f"{int(f('launch__grid_size')):5d} {dram:8.2f} {by / 1e6:8.2f} {roof:7.1f} {roof / us:5.2f}")


if __name__ == "__main__":
    main()
