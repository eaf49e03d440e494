#!/usr/bin/env python
"""NVLink data counters of every visible GPU (NVML field values, KiB summed over the links).

usage: nvlink_bytes.py            -> one JSON line {"gpu0": {"tx_kib": .., "rx_kib": ..}, ...}
       nvlink_bytes.py A.json B.json -> per-GPU difference B - A in MB
Used around a run of known length to compare the bytes that actually crossed NVLink with the halo
volume the decomposition predicts (profiles/scaling.md)."""
import json
import sys


def read():
    import pynvml as nv
    nv.nvmlInit()
    out = {}
    tx = getattr(nv, "NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_TX", 138)
    rx = getattr(nv, "NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_RX", 139)
    for i in range(nv.nvmlDeviceGetCount()):
        h = nv.nvmlDeviceGetHandleByIndex(i)
        rec = {}
        try:
            vals = nv.nvmlDeviceGetFieldValues(h, [(tx, 0xFFFFFFFF), (rx, 0xFFFFFFFF)])
            for name, v in zip(("tx_kib", "rx_kib"), vals):
                rec[name] = int(v.value.ullVal) if v.nvmlReturn == 0 else None
                if v.nvmlReturn != 0:
                    rec[name + "_err"] = int(v.nvmlReturn)
        except Exception as e:          # old pynvml: plain field ids, scope 0
            rec["error"] = repr(e)[:200]
        out[f"gpu{i}"] = rec
    return out


if len(sys.argv) == 3:
    a, b = (json.load(open(f)) for f in sys.argv[1:3])
    for g in sorted(a):
        d = {k: (b[g][k] - a[g][k]) * 1024 / 1e6 for k in ("tx_kib", "rx_kib")
             if a[g].get(k) is not None and b[g].get(k) is not None}
        print(g, {k.replace("_kib", "_MB"): round(v, 1) for k, v in d.items()} or "counters not available")
else:
    print(json.dumps(read()))
