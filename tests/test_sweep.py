"""Host logic of the dataset sweep: buckets, rank sharding, the statistics all-reduce (gloo, world_size 2)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import mel_oracle as orc
from xtts_b200.sweep import (Bucket, allreduce_stats, frame_count, make_buckets, norms_from_stats, shard_buckets)


def _lengths(n, seed=1234):
    rng = np.random.default_rng(seed)
    return [int(round(22050 * u)) for u in rng.uniform(1.0, 30.0, size=n)]


def test_frame_count_matches_oracle():
    for L in (513, 1024, 22050, 220500, 661500, 2646000):
        assert frame_count(L) == orc.frame_count(L)
    assert frame_count(24000, pad=384) == orc.frame_count(24000, orc.V2_SAME)


def test_buckets_partition_and_bound():
    lens = _lengths(256)
    buckets = make_buckets(lens, 2 * 22050, max_clips=64)
    seen = sorted(i for b in buckets for i in b.indices)
    assert seen == list(range(256))                                  # every clip exactly once
    for b in buckets:
        ls = [lens[i] for i in b.indices]
        assert max(ls) == b.max_len and len(ls) <= 64
        assert max(ls) // (2 * 22050) == min(ls) // (2 * 22050)      # one 2 s band per bucket
        assert b.frames == sum(orc.frame_count(L) for L in ls)
        assert b.max_len * len(ls) <= 64 * 220500
    assert len({b.max_len // (2 * 22050) for b in buckets}) <= 15    # <= 15 bands for 1..30 s
    waste = sum(b.max_len * len(b.indices) for b in buckets) / sum(lens)
    assert waste < 1.08                                              # padding overhead of the buckets


def test_empty_and_single():
    assert make_buckets([], 1000) == []
    b = make_buckets([700], 1000)
    assert b == [Bucket((0,), 700, orc.frame_count(700))]


@pytest.mark.parametrize("world", [1, 2, 4, 8])
def test_sharding_balances_frames(world):
    lens = _lengths(2000, seed=4321)
    buckets = make_buckets(lens, 2 * 22050, max_clips=64)
    owned = shard_buckets(buckets, world)
    flat = sorted(j for o in owned for j in o)
    assert flat == list(range(len(buckets)))                         # disjoint cover
    loads = [sum(buckets[j].frames for j in o) for o in owned]
    assert max(loads) - min(loads) <= max(b.frames for b in buckets) # within one batch
    assert max(loads) / (sum(loads) / world) < 1.05


def test_norms_from_stats():
    lm = np.random.default_rng(0).standard_normal((3, 80, 50))
    acc = torch.from_numpy(orc.mel_stats(lm, [50, 20, 7]))
    mean, std = norms_from_stats(acc)
    assert np.allclose(mean.numpy(), orc.norms_from_stats(acc.numpy()), rtol=1e-6)
    allv = np.concatenate([lm[0], lm[1][:, :20], lm[2][:, :7]], axis=1)
    assert np.allclose(std.numpy(), allv.std(axis=1), rtol=1e-5)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, lens, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        buckets = make_buckets(lens, 2 * 22050, max_clips=8)
        owned = shard_buckets(buckets, world)
        acc = torch.zeros(161, dtype=torch.float64)
        for j in owned[rank]:                      # the oracle stands in for the GPU kernels on this CPU box
            for i in buckets[j].indices:
                x = orc.synth("uniform", 1, lens[i], 9000 + i)
                lm = orc.callsite_logmel(x, orc.V1, None).numpy()
                acc += torch.from_numpy(orc.mel_stats(lm, [lm.shape[2]]))
        allreduce_stats(acc)
        ret[rank] = acc.numpy()
    finally:
        dist.destroy_process_group()


def test_two_rank_allreduce_equals_single_rank():
    lens = [int(v) for v in np.random.default_rng(5).integers(600, 9000, size=24)]
    world = 2
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), lens, ret), nprocs=world, join=True)
    single = np.zeros(161)
    for i, L in enumerate(lens):
        lm = orc.callsite_logmel(orc.synth("uniform", 1, L, 9000 + i), orc.V1, None).numpy()
        single += orc.mel_stats(lm, [lm.shape[2]])
    assert np.array_equal(ret[0], ret[1])                               # every rank derives identical norms
    assert ret[0][160] == single[160] == sum(orc.frame_count(L) for L in lens)
    assert np.allclose(ret[0], single, rtol=1e-12, atol=0)              # fp64 all-reduce == single-rank result


def test_mel_pth_sink_writes_the_reference_format(tmp_path):
    """mel_extract.py:26-27: torch.save(mel.cpu().detach(), f'{wav_file}.mel.pth') with mel [1, n_mels, T] fp32."""
    from xtts_b200.sweep import mel_pth_done, mel_pth_sink
    paths = [str(tmp_path / f"clip{i}.wav") for i in range(4)]
    mel = torch.arange(2 * 5 * 7, dtype=torch.float32).reshape(2, 5, 7)
    sink, done = mel_pth_sink(paths), mel_pth_done(paths)
    assert not any(done(i) for i in range(4))
    sink([3, 1], mel, torch.tensor([7, 4], dtype=torch.int32))
    assert [done(i) for i in range(4)] == [False, True, False, True]
    a, b = torch.load(paths[3] + ".mel.pth"), torch.load(paths[1] + ".mel.pth")
    assert a.shape == (1, 5, 7) and a.dtype == torch.float32 and torch.equal(a[0], mel[0])
    assert b.shape == (1, 5, 4) and torch.equal(b[0], mel[1, :, :4])
    assert a.is_contiguous() and a.untyped_storage().nbytes() == a.numel() * 4     # no padded batch behind the file
    import os
    assert not [f for f in os.listdir(tmp_path) if f.endswith(".tmp")]


def _toy_encoder(dim):
    """Stands in for the DVAE encoder (a model, outside the path): 4 frames -> 1 row, a fixed projection to dim."""
    w = torch.linspace(-1.0, 1.0, 5 * 4 * dim).reshape(5 * 4, dim)

    def encode(mel_bct, n_frames):
        B, C, T = mel_bct.shape
        x = mel_bct[:, :, :T // 4 * 4].reshape(B, C, T // 4, 4).permute(0, 2, 1, 3).reshape(B, T // 4, C * 4)
        return (x @ w.to(x.device)).contiguous(), n_frames // 4
    return encode


def test_melvq_pth_sink_writes_the_reference_format(tmp_path):
    """extract_vq.py:26-29: torch.save(code.tolist(), path + '.melvq.pth') with code [1, T'] int64."""
    from oracle import vq_oracle as vo
    from xtts_b200.sweep import melvq_pth_sink
    paths = [str(tmp_path / f"clip{i}.wav") for i in range(3)]
    embed = vo.synth_codebook(16, 64, 3)
    mel = torch.randn(2, 5, 12, generator=torch.Generator().manual_seed(1))
    search = lambda logits: vo.callsite_quantize(logits.numpy(), embed)[2]       # CPU stand-in for the kernel
    sink = melvq_pth_sink(paths, _toy_encoder(16), None, search=search)
    sink([2, 0], mel, torch.tensor([12, 9]))
    a, b = torch.load(paths[2] + ".melvq.pth"), torch.load(paths[0] + ".melvq.pth")
    assert isinstance(a, list) and len(a) == 1 and len(a[0]) == 3 and all(isinstance(c, int) for c in a[0])
    assert len(b) == 1 and len(b[0]) == 2
    logits, _ = _toy_encoder(16)(mel, torch.tensor([12, 9]))
    assert a == vo.callsite_quantize(logits[0:1].numpy(), embed)[2].tolist()
    import os
    assert not [f for f in os.listdir(tmp_path) if f.endswith(".tmp")]


@pytest.mark.gpu
def test_melvq_pth_sink_on_the_gpu_matches_the_cpu_call_sequence(tmp_path):
    from oracle import vq_oracle as vo
    from xtts_b200.sweep import melvq_pth_sink
    from xtts_b200.vq import Quantize
    paths = [str(tmp_path / f"clip{i}.wav") for i in range(2)]
    embed = vo.synth_codebook(64, 512, 3)
    q = Quantize(64, 512)
    q.embed.copy_(torch.from_numpy(embed))
    q = q.eval().cuda()
    mel = torch.randn(2, 5, 400, generator=torch.Generator().manual_seed(1))
    enc = _toy_encoder(64)

    def enc_same_bits(m, n):                   # the projection on the CPU, so that both sides search identical rows
        logits, t = enc(m.cpu(), n.cpu())
        return logits.cuda(), t
    melvq_pth_sink(paths, enc_same_bits, q)([0, 1], mel.cuda(), torch.tensor([400, 123]).cuda())
    logits, _ = enc(mel, torch.tensor([400, 123]))
    want = vo.callsite_quantize(logits.numpy(), embed)[2]
    assert torch.load(paths[0] + ".melvq.pth") == want[0:1].tolist()
    assert torch.load(paths[1] + ".melvq.pth") == want[1:2, :30].tolist()


@pytest.mark.gpu
def test_pcm16_corpus_sweep_matches_clip_by_clip_reference(tmp_path):
    """WAV files -> pcm16_fetch -> MelSweep -> .mel.pth files, against load_audio + the V2 call per clip."""
    import numpy as np
    from oracle import ingest_oracle as ing
    from oracle import mel_oracle as orc
    from xtts_b200 import MelSpectrogramFeatures
    from xtts_b200.sweep import MelSweep, mel_pth_done, mel_pth_sink, pcm16_fetch
    frames = [30000, 52000, 44100, 61000, 30500, 8000]
    paths = []
    for i, f in enumerate(frames):
        p = str(tmp_path / f"u{i}.wav")
        ing.write_wav(p, ing.synth_pcm16(f, 1 + i % 2, 900 + i, "uniform"), 44100)
        paths.append(p)
    mod = MelSpectrogramFeatures()
    plan = mod._plan_for(torch.device("cuda:0"))
    lengths = [ing.out_len(f, 44100, 24000) for f in frames]
    sweep = MelSweep(plan, lengths, sample_rate=24000, max_clips=4)
    fetch = pcm16_fetch(lambda i: ing.read_wav(paths[i])[0], 44100, 24000, "cuda:0")
    acc = sweep.run_stats(fetch, done=mel_pth_done(paths), sink=mel_pth_sink(paths))
    assert float(acc[-1]) == sum(frame_count(l) for l in lengths)
    for i, p in enumerate(paths):
        got = torch.load(p + ".mel.pth")
        wav = ing.callsite_load_audio(ing.read_wav(p)[0], 44100, 24000)
        ref = orc.callsite_logmel(wav, orc.V2, None)
        assert got.shape == ref.shape and got.dtype == torch.float32
        assert float((got - ref).abs().max()) <= 2e-3     # log of |.| amplifies the 3e-7 resample tolerance near the clamp
    # resume: everything exists now.  The statistics still cover the whole corpus (a resumed sweep must derive
    # the same mel_norms), nothing is written again, and the file pass does not even fetch
    stamps = [os.stat(p + ".mel.pth").st_mtime_ns for p in paths]
    acc2 = sweep.run_stats(fetch, done=mel_pth_done(paths), sink=mel_pth_sink(paths))
    assert torch.equal(acc2, acc)                                   # fixed-point statistics: bit-identical run to run
    assert stamps == [os.stat(p + ".mel.pth").st_mtime_ns for p in paths]
    calls = []
    assert sweep.run_files(lambda idx, m: calls.append(idx), mel_pth_sink(paths), done=mel_pth_done(paths)) == 0
    assert calls == []
    # the module that owns the plan still applies ITS norms (none here) after the sweep touched the plan
    # (V2's bar, DESIGN section 2: 1e-4 on the log-mel, or 2e-5 absolute on the mel magnitude where a one-bin filter is small)
    x = torch.rand(1, 24000, device="cuda:0", generator=torch.Generator(device="cuda:0").manual_seed(7)) * 2 - 1
    ref = orc.callsite_logmel(x.cpu(), orc.V2, None).to("cuda:0")
    assert bool(((mod(x) - ref).abs() <= torch.maximum(torch.full_like(ref, 1e-4), 2e-5 / ref.exp())).all())


def test_numa_binding_reads_sysfs_and_degrades_quietly(tmp_path, monkeypatch):
    from xtts_b200 import hostmem
    assert hostmem._parse_cpulist("0-3,8,10-11\n") == {0, 1, 2, 3, 8, 10, 11}
    assert hostmem._parse_cpulist("") == set()

    class Props:
        pci_domain_id, pci_bus_id, pci_device_id = 0, 0x1b, 0
    monkeypatch.setattr(torch.cuda, "get_device_properties", lambda i: Props())
    # no sysfs entry (VM / single socket): nothing changes
    before = os.sched_getaffinity(0)
    assert hostmem.bind_to_gpu_numa(0, sysfs=str(tmp_path)) == {"node": None, "cpus": len(before)}
    d = tmp_path / "bus/pci/devices/0000:1b:00.0"
    d.mkdir(parents=True)
    (d / "numa_node").write_text("-1\n")
    assert hostmem.gpu_numa_node(0, sysfs=str(tmp_path)) is None
    (d / "numa_node").write_text("1\n")
    assert hostmem.gpu_numa_node(0, sysfs=str(tmp_path)) == 1
    n = tmp_path / "devices/system/node/node1"
    n.mkdir(parents=True)
    keep = sorted(before)[:1]
    (n / "cpulist").write_text(f"{keep[0]},100000\n")           # one CPU we may use, one that does not exist
    try:
        assert hostmem.bind_to_gpu_numa(0, sysfs=str(tmp_path)) == {"node": 1, "cpus": 1}
        assert os.sched_getaffinity(0) == set(keep)
    finally:
        os.sched_setaffinity(0, before)
