"""bench.py pieces that can be checked without a GPU: the clocks sampler (NVML child process with the recipe's
nvidia-smi query as the fall-back) and the "no device" behaviour of both arms."""
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def test_sampler_falls_back_when_nvml_is_missing():
    import bench
    s = bench.make_sampler(0)
    s.start()
    out = s.stop()
    assert set(out) >= {"sm_mhz", "sm_max_mhz", "reasons"}
    if isinstance(s, bench.ClockSampler):               # this box: neither NVML nor nvidia-smi answer
        assert out["sm_mhz"] is None or out["sm_mhz"] > 0


def test_nvml_sampler_keeps_only_samples_of_the_timed_region_and_decodes_reasons(monkeypatch):
    """A stand-in child that speaks the sampler's line protocol: samples stamped before start() or after stop()
    must not count, and the reasons bitmask must come back as the recipe's names."""
    import bench
    fake = r'''
import sys, time
print("ready", 1965, flush=True)
i = 0
while True:
    i += 1
    # sw_power_cap (0x4) on every sample, hw_slowdown (0x8) only on the very first one (before start())
    print(time.time(), 1900 + (i % 3), 0x4 | (0x8 if i == 1 else 0), 250.0, flush=True)
    time.sleep(0.002)
'''
    monkeypatch.setattr(bench.NvmlSampler, "CHILD", fake)
    monkeypatch.setattr(bench.torch.cuda, "get_device_properties", lambda i: (_ for _ in ()).throw(RuntimeError("no gpu")))
    s = bench.NvmlSampler(0)
    time.sleep(0.05)                   # samples before the timed region
    s.start()
    time.sleep(0.08)
    out = s.stop()
    assert out["source"] == "nvml" and out["sm_max_mhz"] == 1965.0
    assert 2 <= out["samples"] <= 80, out
    assert 1900 <= out["sm_mhz"] <= 1902
    assert out["reasons"] == ["sw_power_cap"], out          # the hw_slowdown sample predates start()
    assert out["power_w"] == 250.0


def test_both_arms_say_unavailable_without_a_gpu():
    import torch
    if torch.cuda.is_available():
        return
    for impl in ("ours", "reference"):
        r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", impl, "--steps", "2", "--warmup", "1"],
                           capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stderr[-400:]
        line = json.loads(r.stdout.strip().splitlines()[-1])
        assert line["impl"] == impl and "unavailable" in line
