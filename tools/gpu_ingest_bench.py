"""Time the ingestion kernels (k_pcm16, k_resample) on one GPU against their HBM / FMA bounds.

    python tools/gpu_ingest_bench.py [out.json]

Workload: 64 clips x 10 s (BASELINE configs[1] sized at the source rate).  Inputs resident in HBM, CUDA events on
the launching stream, 3 warm-ups, 20 timed launches, an L2 flush (256 MB write) between launches.
torchaudio's own GPU resample (library code: conv1d) is timed beside it for scale, the CPU call on a bounded
sample as the reference arm.
"""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from xtts_b200 import Resampler, decode_pcm16  # noqa: E402


def timed(fn, flush, n=20, warm=3):
    for _ in range(warm):
        fn()
    ms = []
    for _ in range(n):
        flush.fill_(1.0)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ms.append(a.elapsed_time(b))
    return float(np.median(ms)), float(np.min(ms))


def main():
    import torchaudio
    dev = torch.device("cuda:0")
    flush = torch.empty(64 * 1024 * 1024, dtype=torch.float32, device=dev)
    peaks = {}
    p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    if os.path.exists(p):
        peaks = json.load(open(p))
    res = {"peaks_file": peaks, "rows": []}
    B, secs = 64, 10
    for ch in (1, 2):
        F = 44100 * secs
        pcm = torch.randint(-32768, 32767, (B, F, ch), dtype=torch.int16, device=dev)
        med, best = timed(lambda: decode_pcm16(pcm), flush)
        byts = B * F * (2 * ch + 4)
        res["rows"].append({"kernel": "k_pcm16", "channels": ch, "ms": med, "ms_min": best, "bytes": byts,
                            "GB/s": byts / med / 1e6, "audio_s_per_s": B * secs / med * 1e3})
    for o, n in ((44100, 22050), (48000, 24000), (24000, 22050), (44100, 24000), (16000, 22050)):
        L = o * secs
        x = torch.rand((B, L), device=dev) * 2 - 1
        rs = Resampler(o, n)
        med, best = timed(lambda: rs(x), flush)
        Lo = rs.out_len(L)
        byts = B * (L + Lo) * 4
        flop = 2.0 * B * Lo * (2 * rs.width + rs.orig)
        tms, _ = timed(lambda: torchaudio.functional.resample(x, o, n), flush, n=5, warm=2)
        y = rs(x)
        ref = torchaudio.functional.resample(x, o, n)
        err = float((y - ref).abs().max())
        xc = x[:4].cpu()
        torch.set_num_threads(os.cpu_count() or 1)
        t0 = time.perf_counter()
        torchaudio.functional.resample(xc, o, n)
        cpu_s = time.perf_counter() - t0
        res["rows"].append({"kernel": "k_resample", "orig": rs.orig, "new": rs.new, "taps": 2 * rs.width + rs.orig,
                            "ms": med, "ms_min": best, "bytes": byts, "GB/s": byts / med / 1e6,
                            "TFLOP/s": flop / med / 1e9, "audio_s_per_s": B * secs / med * 1e3,
                            "torchaudio_gpu_ms": tms, "max_abs_diff_vs_torchaudio_gpu": err,
                            "torchaudio_cpu_audio_s_per_s": 4 * secs / cpu_s, "cpu_threads": torch.get_num_threads()})
    for o, n, ch in ((44100, 22050, 2), (48000, 24000, 1), (44100, 24000, 2)):
        F = o * secs
        pcm = torch.randint(-32768, 32767, (B, F, ch), dtype=torch.int16, device=dev)
        rs = Resampler(o, n)
        med, best = timed(lambda: rs.from_pcm16(pcm, clip=True), flush)
        byts = B * (F * 2 * ch + rs.out_len(F) * 4)
        med2, _ = timed(lambda: rs(decode_pcm16(pcm), clip=True), flush)
        res["rows"].append({"kernel": "fused pcm16 -> resample", "orig": rs.orig, "new": rs.new, "channels": ch,
                            "ms": med, "ms_min": best, "bytes": byts, "GB/s": byts / med / 1e6,
                            "audio_s_per_s": B * secs / med * 1e3, "two_step_ms": med2})
    for r in res["rows"]:
        print(json.dumps(r))
    if len(sys.argv) > 1:
        os.makedirs(os.path.dirname(os.path.abspath(sys.argv[1])), exist_ok=True)
        json.dump(res, open(sys.argv[1], "w"), indent=1)


if __name__ == "__main__":
    main()
