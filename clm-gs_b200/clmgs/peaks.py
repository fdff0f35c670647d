"""Measured roofline denominators of the box the process runs on.

    hbm        : MEASURED_PEAKS.json (driver-written) -- not measured here
    host link  : cudaMemcpyAsync pinned host <-> device, 1 GiB, best of 10
    FP32 SIMT  : FMA chain kernel, all SMs, best of 5
    host DRAM  : STREAM triad on the host threads the host Adam uses, best of 5
"""
import os
import time

import numpy as np
import torch

from ._lib import call


def measure_host_link(device, nbytes=1 << 30, reps=10):
    h = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
    d = torch.empty(nbytes, dtype=torch.uint8, device=device)
    h.fill_(1)
    st = torch.cuda.current_stream().cuda_stream
    out = {}
    for name, dst, src in (("h2d_gbs", d, h), ("d2h_gbs", h, d)):
        best = 1e30
        for _ in range(reps + 1):
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            call("clmgs_measure_memcpy", dst, src, nbytes, st)
            e.record()
            e.synchronize()
            best = min(best, s.elapsed_time(e))
        out[name] = round(nbytes / best / 1e6, 2)
    del h, d
    return out


def measure_fma(device, iters=4096, reps=5):
    sms = torch.cuda.get_device_properties(device).multi_processor_count
    grid = sms * 2
    sink = torch.zeros(1, device=device)
    st = torch.cuda.current_stream().cuda_stream
    best = 1e30
    for _ in range(reps + 1):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        call("clmgs_measure_fma", grid, iters, sink, st)
        e.record()
        e.synchronize()
        best = min(best, s.elapsed_time(e))
    flops = grid * 1024.0 * iters * 128 * 2
    return round(flops / best / 1e9, 2)  # TFLOP/s


def measure_host_triad(n_threads, n=1 << 27, reps=5):
    a = np.zeros(n, np.float32)
    b = np.ones(n, np.float32)
    c = np.ones(n, np.float32)
    best = 1e30
    for _ in range(reps + 1):
        t0 = time.perf_counter()
        call("clmgs_measure_host_triad", a, b, c, n, 3.0, int(n_threads))
        best = min(best, time.perf_counter() - t0)
    return round(12.0 * n / best / 1e9, 2)


def measure_all(device, host_threads=0):
    host_threads = host_threads or (os.cpu_count() or 1)
    out = measure_host_link(device)
    out["fma_tflops"] = measure_fma(device)
    out["host_triad_gbs"] = measure_host_triad(host_threads)
    out["host_threads"] = host_threads
    out["how"] = ("cudaMemcpyAsync 1 GiB pinned<->device best of 10; FMA chain kernel 2 CTAs x 1024 threads per SM best "
                  "of 5; STREAM triad 3 x 512 MiB on the host threads best of 5 (12 B/element counted)")
    return out
