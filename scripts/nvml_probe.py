"""How much does an NVML clock query cost a host-synchronised GPU loop?  (bench.py samples clocks during the timed
region; this measures which calls are cheap enough to do that without distorting the number.)"""
import threading
import time

import pynvml
import torch

pynvml.nvmlInit()
h = pynvml.nvmlDeviceGetHandleByIndex(0)
x = torch.zeros(1 << 22, device="cuda")
stop = threading.Event()
laps = []


def load():
    torch.cuda.set_device(0)
    while not stop.is_set():
        t0 = time.perf_counter()
        for _ in range(40):
            x.add_(1.0)
        torch.cuda.synchronize()
        laps.append((time.perf_counter(), time.perf_counter() - t0))


t = threading.Thread(target=load)
t.start()
time.sleep(1.0)
calls = {
    "GetClockInfo(SM)": lambda: pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM),
    "GetCurrentClocksEventReasons": lambda: pynvml.nvmlDeviceGetCurrentClocksEventReasons(h),
    "GetMaxClockInfo(SM)": lambda: pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM),
    "GetClock(SM, CURRENT)": lambda: pynvml.nvmlDeviceGetClock(h, pynvml.NVML_CLOCK_SM, pynvml.NVML_CLOCK_ID_CURRENT),
    "GetPowerUsage": lambda: pynvml.nvmlDeviceGetPowerUsage(h),
    "GetTemperature": lambda: pynvml.nvmlDeviceGetTemperature(h, pynvml.NVML_TEMPERATURE_GPU),
}
for name, fn in calls.items():
    time.sleep(0.5)
    a = time.perf_counter()
    base = [d for (ts, d) in laps if ts > a - 0.4]
    dts = []
    for _ in range(10):
        t0 = time.perf_counter()
        try:
            v = fn()
        except Exception as e:
            v = repr(e)
        dts.append(time.perf_counter() - t0)
        time.sleep(0.02)
    b = time.perf_counter()
    during = [d for (ts, d) in laps if a < ts <= b]
    print("%-32s value %s  call ms: median %.3f max %.3f | loop lap ms: before median %.3f max %.3f, during median %.3f max %.3f"
          % (name, v, 1e3 * sorted(dts)[5], 1e3 * max(dts), 1e3 * sorted(base)[len(base) // 2], 1e3 * max(base),
             1e3 * sorted(during)[len(during) // 2], 1e3 * max(during)), flush=True)
stop.set()
t.join()
