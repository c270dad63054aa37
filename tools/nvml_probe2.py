"""Which NVML query disturbs a busy GPU?  Times a fixed batch of matmuls with CUDA events while one NVML query is issued from the host
in the middle of the batch; prints how many batches ran long.  (bench.py saw 66-140 ms "steps" of a 46 ms sort in 1 run of 3 with
one clock + throttle-reason sample per step, and none without the sampler.)"""
import time, sys, os
import torch, pynvml as nv, numpy as np
nv.nvmlInit(); h = nv.nvmlDeviceGetHandleByIndex(0)
x = torch.randn(8192, 8192, device="cuda", dtype=torch.bfloat16)
calls = {
 "none": lambda: None,
 "clock_sm": lambda: nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM),
 "reasons": lambda: nv.nvmlDeviceGetCurrentClocksEventReasons(h),
}
for name, f in calls.items():
    ts = []
    for i in range(150):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(30): (x @ x)
        f()
        for _ in range(30): (x @ x)
        e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts = np.array(ts)
    med = np.median(ts)
    print(name, "median %.2f ms  max %.2f  >1.2x median: %d of %d" % (med, ts.max(), (ts > 1.2 * med).sum(), len(ts)), [round(t, 1) for t in ts if t > 1.2 * med][:8])
