"""Where does k_logmel's time go?  Remove pipeline legs one at a time (results invalid, timing only)."""
import os, sys, ctypes
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from xtts_b200 import MelPlan, _cabi
dev = torch.device("cuda:0")
if os.environ.get("TTTS_LAB_SO"):       # time another build of the lab library (e.g. the previous commit's)
    _cabi.LAB_PATH = os.environ["TTTS_LAB_SO"]
lib = _cabi.lab()          # leg removal exists in the lab build only (include/ttts_mel_lab.h)
V2 = "--v2" in sys.argv          # the MelSpectrogramFeatures parametrisation (24 kHz, 100 mel, 512 live bins, power 1)
if V2:
    sys.argv.remove("--v2")
    plan = MelPlan(sample_rate=24000, n_fft=1024, hop_length=256, win_length=1024, n_mels=100, f_min=0, f_max=None,
                   power=1.0, norm=None, clamp=1e-7, device=dev, lab=True)
    x = torch.rand(64, 240000, device=dev) * 2 - 1
else:
    plan = MelPlan(sample_rate=22050, n_fft=1024, hop_length=256, win_length=1024, n_mels=80, f_min=0, f_max=8000,
                   power=2.0, norm="slaney", device=dev, lab=True)
    x = torch.rand(64, 220500, device=dev) * 2 - 1
names = {0: "full", 1: "no A handoff", 2: "no basis loads", 4: "no epilogue", 3: "MMA+epilogue only", 5: "no A, no epi",
         6: "no basis, no epi", 7: "MMA issue only", 8: "no drain math", 16: "no log pass", 24: "TMEM drain only", 32: "no MMAs", 36: "no MMAs, no epilogue", 38: "converters only (no MMA/basis/epi)", 34: "no MMAs, no basis",
         35: "epilogue only", 43: "epilogue only, no drain math", 51: "epilogue only, no log pass", 59: "epilogue only: TMEM loads + barriers",
         71: "MMA issue only, no per-k-step commits", 39: "bare issue loop (no MMAs)", 103: "bare loop, no commits",
         128: "no raw-audio loads", 132: "no raw loads, no epilogue", 134: "no raw, no basis, no epilogue"}
FLAGS = [int(a) for a in sys.argv[1:]] or [0, 4, 8, 16, 24, 7, 32, 36, 38, 34, 0]
for flags in FLAGS:
    lib.ttts_mel_set_experiment(flags)
    for _ in range(3):
        plan.forward(x)
    torch.cuda.synchronize()
    plan.profile_enable(20)
    for _ in range(20):
        plan.forward(x)
    ms = sorted(plan.profile_collect(20))
    n = len(ms)
    print(f"flags={flags} ({names.get(flags, "?"):>20s}): k_logmel median {ms[n//2]*1e3:7.1f} us  min {ms[0]*1e3:7.1f} us", flush=True)
    plan.profile_enable(0)
lib.ttts_mel_set_experiment(0)
