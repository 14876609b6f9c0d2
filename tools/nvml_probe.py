import time, torch, pynvml as nv
nv.nvmlInit(); h = nv.nvmlDeviceGetHandleByIndex(0)
x = torch.randn(8192, 8192, device="cuda")
def t(f, n=10):
    ts=[]
    for _ in range(n):
        a=time.perf_counter(); f(); ts.append((time.perf_counter()-a)*1e3)
    return [round(v,2) for v in ts]
for load in (False, True):
    if load:
        for _ in range(200): y = x @ x
    print("load" if load else "idle", "clock", t(lambda: nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)))
    if load:
        for _ in range(200): y = x @ x
    print("load" if load else "idle", "reasons", t(lambda: nv.nvmlDeviceGetCurrentClocksEventReasons(h)))
    if load:
        for _ in range(200): y = x @ x
    print("load" if load else "idle", "power", t(lambda: nv.nvmlDeviceGetPowerUsage(h)))
    torch.cuda.synchronize()
