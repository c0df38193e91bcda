"""On the GPU box: time every build/variants/v*.so (built by tools/variants.sh) on the V1 and V2 64 x 10 s batches and
compare each variant's output with variant 0's (max abs difference), one subprocess per library.

    python tools/gpu_variants.py            # all variants
    python tools/gpu_variants.py --one build/variants/v3.so   # worker
"""
import glob, json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def worker(so):
    import ctypes
    import torch
    from xtts_b200 import _cabi
    _cabi.LIB_PATH = so
    from xtts_b200 import MelPlan
    lib = _cabi.lib()
    dev = torch.device("cuda:0")
    res = {"so": os.path.basename(so)}
    g = torch.Generator(device=dev).manual_seed(7)
    cases = {
        "v1": (dict(sample_rate=22050, n_fft=1024, hop_length=256, win_length=1024, n_mels=80, f_min=0, f_max=8000,
                    power=2.0, norm="slaney"), 220500),
        "v2": (dict(sample_rate=24000, n_fft=1024, hop_length=256, win_length=1024, n_mels=100, f_min=0, f_max=None,
                    power=1.0, norm=None, clamp=1e-7), 240000),
    }
    only = os.environ.get("VARIANTS_CASES")          # e.g. "v2": variants built for one parametrisation only
    for name, (kw, L) in cases.items():
        if only and name not in only.split(","):
            continue
        plan = MelPlan(device=dev, **kw)
        xs = [torch.rand(64, L, device=dev, generator=g) * 2 - 1 for _ in range(4)]
        for i in range(5):
            plan.forward(xs[i & 3])
        plan.status()
        ref_path = f"/tmp/variant_ref_{name}.pt"
        y0 = plan.forward(xs[0])[0]
        torch.cuda.synchronize()
        if os.path.exists(ref_path):
            res[name + "_maxdiff_vs_v0"] = float((y0 - torch.load(ref_path).to(dev)).abs().max())
        else:
            torch.save(y0.cpu(), ref_path)
        n = 200
        plan.profile_enable(n)
        for i in range(n):
            plan.forward(xs[i & 3])
        ms = sorted(plan.profile_collect(n))
        k = len(ms)
        plan.profile_enable(0)
        res[name + "_us_median"] = round(ms[k // 2] * 1e3, 2)
        res[name + "_us_min"] = round(ms[0] * 1e3, 2)
        if os.environ.get("VARIANTS_SUSTAINED"):
            # the bench's regime: thousands of launches back to back, the board at its power cap (SM clock ~1.75 GHz)
            plan.profile_enable(0)
            ns = int(os.environ["VARIANTS_SUSTAINED"])
            for rep in range(2):
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                for i in range(ns):
                    plan.forward(xs[i & 3])
                b.record()
                torch.cuda.synchronize()
            res[name + "_sustained_step_us"] = round(a.elapsed_time(b) / ns * 1e3, 2)
        plan.status()
    print("RESULT " + json.dumps(res), flush=True)


def main():
    if len(sys.argv) > 2 and sys.argv[1] == "--one":
        return worker(sys.argv[2])
    names = {}
    lst = os.path.join(ROOT, "build", "variants", "list.txt")
    if os.path.exists(lst):
        for line in open(lst):
            k, _, v = line.partition(":")
            names[k.strip()] = v.strip()
    for f in glob.glob("/tmp/variant_ref_*.pt"):
        os.remove(f)
    out = []
    sos = sorted(glob.glob(os.path.join(ROOT, "build", "variants", "v*.so")), key=lambda p: int(os.path.basename(p)[1:-3]))
    for so in sos:
        try:
            r = subprocess.run([sys.executable, __file__, "--one", so], capture_output=True, text=True, timeout=180)
            line = [l for l in r.stdout.splitlines() if l.startswith("RESULT ")]
            if line:
                d = json.loads(line[-1][7:])
            else:
                d = {"so": os.path.basename(so), "error": (r.stderr or r.stdout)[-600:]}
        except subprocess.TimeoutExpired:
            d = {"so": os.path.basename(so), "error": "timeout"}
        d["flags"] = names.get(os.path.basename(so)[:-3], "")
        print(json.dumps(d), flush=True)
        out.append(d)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(out, open(os.path.join(ROOT, "gpurun_out", "variants.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
