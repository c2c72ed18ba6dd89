"""Which way of sampling clocks during a timed region disturbs the kernels?  Runs the config-4 kernel in 14 ms steps
for ~3 s under one sampler kind and prints the slow steps with the time since the nearest sampler query.

    python tools/hiccup_probe.py none|smi100|smi250|nvml100|nvml250
"""
import os
import subprocess
import sys
import threading
import time

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "stochastic-vector-oracle-consensus_b200"))
import torch  # noqa: E402
import svoc  # noqa: E402

kind = sys.argv[1]
T, N, D, F = int(os.environ.get("PROBE_T", "200000")), 256, 128, 51
x = svoc.generate(T, N, D, F, a=10.0, b=15.0, failing="colluding", seed=1, dtype=torch.float32).pred
for _ in range(5):
    svoc.consensus_float(x, F, want_qr=False, want_essence=False)
torch.cuda.synchronize()
queries = []
stop = False


def smi(period):
    p = subprocess.Popen(["nvidia-smi", "-i", "0", "--query-gpu=clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
                          "clocks_event_reasons.sw_power_cap", "--format=csv,noheader,nounits", "-lms", str(period)],
                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
    for line in p.stdout:
        queries.append(time.time())
        if stop:
            break
    p.terminate()


def nvml(period):
    import pynvml
    pynvml.nvmlInit()
    h = pynvml.nvmlDeviceGetHandleByIndex(0)
    while not stop:
        t = time.time()
        pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)
        pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(h)
        queries.append(t)
        time.sleep(period / 1000.0)


if kind.startswith("smi"):
    threading.Thread(target=smi, args=(int(kind[3:]),), daemon=True).start()
elif kind.startswith("nvml"):
    threading.Thread(target=nvml, args=(int(kind[4:]),), daemon=True).start()
time.sleep(0.6)
steps = int(os.environ.get("PROBE_STEPS", "200"))
ev = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
t_host = []
ev[0].record()
for i in range(steps):
    svoc.consensus_float(x, F, want_qr=False, want_essence=False)
    ev[i + 1].record()
    if i % 8 == 7:
        torch.cuda.synchronize()          # keep the host close to the device so host timestamps mean something
    t_host.append(time.time())
torch.cuda.synchronize()
stop = True
ms = [ev[i].elapsed_time(ev[i + 1]) for i in range(steps)]
med = sorted(ms)[steps // 2]
slow = [(i, round(ms[i], 1)) for i in range(steps) if ms[i] > 1.3 * med]
print(kind, "median", round(med, 2), "queries", len([q for q in queries if q >= t_host[0] - 0.1]), "slow steps", slow[:10],
      "extra ms", round(sum(ms[i] - med for i, _ in slow), 1))
