"""Diagnostic: which NVLink byte counters this box exposes (NVML field values, nvidia-smi nvlink -gt d), read around a
256 MB peer copy cuda:0 -> cuda:1."""
import subprocess
import sys

import pynvml
import torch

pynvml.nvmlInit()
pr = torch.cuda.get_device_properties(0)
bus = f"{pr.pci_domain_id:08x}:{pr.pci_bus_id:02x}:{pr.pci_device_id:02x}.0"
print("bus id", bus, flush=True)
try:
    h = pynvml.nvmlDeviceGetHandleByPciBusId(bus.encode())
except Exception as e:
    print("by bus id failed:", repr(e))
    h = pynvml.nvmlDeviceGetHandleByIndex(0)


def read():
    out = {}
    for name, req in (("agg", [(138, 0xFFFFFFFF), (139, 0xFFFFFFFF)]), ("link0", [(138, 0), (139, 0)]), ("plain", [138, 139]),
                      ("raw", [(140, 0xFFFFFFFF), (141, 0xFFFFFFFF)])):
        try:
            vals = pynvml.nvmlDeviceGetFieldValues(h, req)
            out[name] = [(v.nvmlReturn, v.valueType, int(v.value.ullVal)) for v in vals]
        except Exception as e:
            out[name] = repr(e)
    try:
        r = subprocess.run(["nvidia-smi", "nvlink", "-gt", "d", "-i", "0"], capture_output=True, text=True, timeout=30)
        out["smi"] = r.stdout[-900:] + r.stderr[-300:]
    except Exception as e:
        out["smi"] = repr(e)
    return out


a = read()
if torch.cuda.device_count() > 1:
    x = torch.empty(64 * 1024 * 1024, dtype=torch.float32, device="cuda:0")
    y = torch.empty_like(x, device="cuda:1")
    for _ in range(4):
        y.copy_(x)
    torch.cuda.synchronize(0)
    torch.cuda.synchronize(1)
b = read()
for k in a:
    print("==", k)
    print(" before:", a[k])
    print(" after: ", b[k])
