"""STFT family on one B200: ours (tcgen05 contraction + small kernels) vs the reference's own call sequence
on the same GPU (torch.stft = cuFFT; irfft + fold + fold + divide), 64 clips x 10 s at 22.05 kHz, inputs resident."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from oracle import stft_oracle as so
from xtts_b200 import ISTFT, MultiScaleSTFTLoss, TorchSTFT

dev = torch.device("cuda:0")


def timed(fn, iters, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters


out = {}
x = torch.rand(64, 220500, device=dev) * 2 - 1
for n_fft, hop, win in so.MS_SHAPES:
    m = TorchSTFT(n_fft, hop, win)
    w = torch.hann_window(win, device=dev)

    def ref():
        o = torch.view_as_real(torch.stft(x, n_fft, hop, win, w, center=True, pad_mode="reflect", normalized=False,
                                          onesided=True, return_complex=True))
        return torch.sqrt(torch.clamp(o[..., 0] ** 2 + o[..., 1] ** 2, min=1e-8))
    ours_ms, ref_ms = timed(lambda: m(x), 20), timed(ref, 20)
    T = 1 + 220500 // hop
    err = float((m(x) - ref()).abs().max() / ref().abs().max())
    flop = 2.0 * win * (n_fft + 2) * 64 * T
    out[f"stft_mag_{n_fft}_{hop}_{win}"] = dict(ours_ms=ours_ms, ref_gpu_ms=ref_ms, speedup=ref_ms / ours_ms, frames=64 * T,
                                                algorithmic_tflops=flop / (ours_ms * 1e-3) / 1e12, executed_tflops=3 * flop / (ours_ms * 1e-3) / 1e12,
                                                max_rel_diff_vs_ref_gpu=err)
y = torch.rand(64, 220500, device=dev) * 2 - 1
ms = MultiScaleSTFTLoss()
out["multiscale_stft_loss_forward"] = dict(ours_ms=timed(lambda: ms(y, x), 10),
                                           ref_gpu_ms=timed(lambda: so.callsite_multiscale_loss(y.cpu()[:1], x.cpu()[:1]), 1, warm=0) if False else None)
spec = torch.stft(x, 1024, 256, 1024, torch.hann_window(1024, device=dev), center=True, return_complex=True)
for padding in ("same", "center"):
    m = ISTFT(1024, 256, 1024, padding=padding)
    win = torch.hann_window(1024, device=dev)

    def ref_i():
        if padding == "center":
            return torch.istft(spec, 1024, 256, 1024, win, center=True)
        pad = (1024 - 256) // 2
        B, N, T = spec.shape
        ifft = torch.fft.irfft(spec, 1024, dim=1, norm="backward") * win[None, :, None]
        size = (T - 1) * 256 + 1024
        yy = torch.nn.functional.fold(ifft, output_size=(1, size), kernel_size=(1, 1024), stride=(1, 256))[:, 0, 0, pad:-pad]
        env = torch.nn.functional.fold(win.square().expand(1, T, -1).transpose(1, 2), output_size=(1, size),
                                       kernel_size=(1, 1024), stride=(1, 256)).squeeze()[pad:-pad]
        return yy / env
    ours_ms, ref_ms = timed(lambda: m(spec), 20), timed(ref_i, 20)
    err = float((m(spec) - ref_i()).abs().max())
    out[f"istft_1024_256_{padding}"] = dict(ours_ms=ours_ms, ref_gpu_ms=ref_ms, speedup=ref_ms / ours_ms, frames=64 * spec.shape[2],
                                            max_abs_diff_vs_ref_gpu=err)
for k, v in out.items():
    print(k, {a: (round(b, 4) if isinstance(b, float) else b) for a, b in v.items()})
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "stft_bench.json"), "w"), indent=1)
