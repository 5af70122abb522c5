#!/usr/bin/env python3
"""NVML common-ancestor level of every GPU pair (10 same board, 20 one switch, 30 host bridge, 40 NUMA node, 50 system)
and the rank placement atray_b200.multi.spread_devices derives from it."""
import os
import sys

import pynvml

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pynvml.nvmlInit()
n = pynvml.nvmlDeviceGetCount()
hs = [pynvml.nvmlDeviceGetHandleByIndex(i) for i in range(n)]
for i in range(n):
    print(i, pynvml.nvmlDeviceGetPciInfo(hs[i]).busId, [pynvml.nvmlDeviceGetTopologyCommonAncestor(hs[i], hs[j]) if i != j else 0 for j in range(n)])
try:
    from atray_b200 import multi
    for w in (1, 2, 4, 8):
        if w <= n:
            print("world", w, "->", multi.spread_devices(w))
except Exception as e:  # noqa: BLE001
    print("spread_devices:", e)
