"""Peaks MEASURED_PEAKS.json does not carry: dense TF32 throughput (cuBLAS, what the ranking GEMM's
roofline is quoted against) and NVLink peer-to-peer copy bandwidth (what the multi-GPU exchanges are quoted against).
    python scripts/measure_peaks.py            # 1 GPU: tf32 (and fp32 SIMT for reference)
    python scripts/measure_peaks.py --p2p      # >= 2 GPUs visible: also device 0 -> device 1 copy bandwidth
Writes / updates profiles/r02_extra_peaks.json."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "profiles", "r02_extra_peaks.json")


def best_ms(fn, reps=10):
    fn()
    torch.cuda.synchronize()
    best = 1e30
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b))
    return best


def main():
    res = json.load(open(OUT)) if os.path.exists(OUT) else {}
    dev = torch.device("cuda", 0)
    n = 8192
    x = torch.randn((n, n), device=dev)
    y = torch.randn((n, n), device=dev)
    torch.backends.cuda.matmul.allow_tf32 = True
    ms = best_ms(lambda: torch.matmul(x, y))
    res["tf32_tflops"] = 2 * n ** 3 / (ms * 1e-3) / 1e12
    torch.backends.cuda.matmul.allow_tf32 = False
    ms = best_ms(lambda: torch.matmul(x, y), reps=3)
    res["fp32_simt_tflops"] = 2 * n ** 3 / (ms * 1e-3) / 1e12
    res["how_tf32"] = "torch.matmul fp32 8192^3, allow_tf32=True (cuBLAS), best of 10, CUDA events"
    res["gpu_name"] = torch.cuda.get_device_name(0)
    if "--p2p" in sys.argv and torch.cuda.device_count() >= 2:
        a = torch.empty(1 << 30, dtype=torch.uint8, device="cuda:0")
        b = torch.empty(1 << 30, dtype=torch.uint8, device="cuda:1")
        ms = best_ms(lambda: b.copy_(a, non_blocking=True))
        res["nvlink_store_gbs"] = (1 << 30) / (ms * 1e-3) / 1e9
        ms2 = best_ms(lambda: (b.copy_(a, non_blocking=True), a.copy_(b, non_blocking=True)))
        res["nvlink_bidir_gbs"] = 2 * (1 << 30) / (ms2 * 1e-3) / 1e9
        res["how_nvlink"] = "cudaMemcpyPeerAsync of 1 GiB device 0 -> device 1 (torch copy_), best of 10, CUDA events on device 0's stream"
    json.dump(res, open(OUT, "w"), indent=1)
    print(json.dumps(res))


if __name__ == "__main__":
    main()
