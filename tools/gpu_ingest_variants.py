"""On the GPU box: time the 2:1 decimator of every build/variants/v*.so (float32 and fused stereo int16, 64 x 10 s at 44.1 kHz,
L2 flushed between launches) and compare with variant 0's output."""
import glob, json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def worker(so):
    import torch
    from xtts_b200 import _cabi
    _cabi.LIB_PATH = so
    from xtts_b200 import Resampler
    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev).manual_seed(3)
    x = torch.rand(64, 441000, device=dev, generator=g) * 2 - 1
    pcm = torch.randint(-32768, 32767, (64, 441000, 2), dtype=torch.int16, device=dev, generator=g)
    flush = torch.empty(64 * 1024 * 1024, dtype=torch.float32, device=dev)
    rs = Resampler(44100, 22050)
    res = {"so": os.path.basename(so)}
    for name, fn, nbytes in (("f32", lambda: rs(x), 64 * 441000 * 4 + 64 * 220500 * 4),
                             ("pcm16x2", lambda: rs.from_pcm16(pcm), 64 * 441000 * 4 + 64 * 220500 * 4)):
        y = fn()
        torch.cuda.synchronize()
        ref = f"/tmp/ingest_ref_{name}.pt"
        if os.path.exists(ref):
            res[name + "_maxdiff_vs_v0"] = float((y - torch.load(ref).to(dev)).abs().max())
        else:
            torch.save(y.cpu(), ref)
        ms = []
        for i in range(23):
            flush.fill_(0.0)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); fn(); b.record()
            torch.cuda.synchronize()
            if i >= 3:
                ms.append(a.elapsed_time(b))
        ms.sort()
        res[name + "_us"] = round(ms[len(ms) // 2] * 1e3, 1)
        res[name + "_TB/s"] = round(nbytes / (ms[len(ms) // 2] * 1e-3) / 1e12, 2)
    print("RESULT " + json.dumps(res), flush=True)


if len(sys.argv) > 2 and sys.argv[1] == "--one":
    worker(sys.argv[2])
else:
    for f in glob.glob("/tmp/ingest_ref_*.pt"):
        os.remove(f)
    names = {l.split(":")[0]: l.split(":", 1)[1].strip() for l in open(os.path.join(ROOT, "build", "variants", "list.txt"))}
    for so in sorted(glob.glob(os.path.join(ROOT, "build", "variants", "v*.so")), key=lambda p: int(os.path.basename(p)[1:-3])):
        r = subprocess.run([sys.executable, __file__, "--one", so], capture_output=True, text=True, timeout=180)
        line = [l for l in r.stdout.splitlines() if l.startswith("RESULT ")]
        d = json.loads(line[-1][7:]) if line else {"so": os.path.basename(so), "error": (r.stderr or r.stdout)[-400:]}
        d["flags"] = names.get(os.path.basename(so)[:-3], "")
        print(json.dumps(d), flush=True)
