#!/usr/bin/env python
"""NVLink traffic of the fused halo push, from the driver's own per-GPU counters (NVML field values
NVLINK_THROUGHPUT_DATA_TX / _RX, KiB, all links summed) — read before and after `reps` solves, outside any
profiler — against the algorithmic halo bytes (every owned face cell with a neighbour rank is stored once,
8 B, into that neighbour's ghost layer; flags and the forwarded maxima are a few hundred bytes per layer).

usage: python scripts/nvlink_bytes.py N steps ngpus [dims] [reps]"""
import ctypes as C
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pynvml  # noqa: E402
import wave3d_b200 as w  # noqa: E402
from wave3d_b200.parallel.decomp import Decomposition  # noqa: E402

N, K, ng = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
dims = w.parallel.decomp.parse_dims(sys.argv[4]) if len(sys.argv) > 4 and sys.argv[4] != "auto" else None
reps = int(sys.argv[5]) if len(sys.argv) > 5 else 5
pynvml.nvmlInit()
hs = [pynvml.nvmlDeviceGetHandleByIndex(i) for i in range(ng)]
FIDS = [pynvml.NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_TX, pynvml.NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_RX]


def counters():
    out = []
    for h in hs:
        vals = pynvml.nvmlDeviceGetFieldValues(h, [(f, 0xFFFFFFFF) for f in FIDS])  # scope: all links
        row = []
        for v in vals:
            if v.nvmlReturn != 0:
                row.append(-1)
            elif v.valueType == pynvml.NVML_VALUE_TYPE_UNSIGNED_LONG_LONG:
                row.append(int(v.value.ullVal))
            elif v.valueType == pynvml.NVML_VALUE_TYPE_UNSIGNED_LONG:
                row.append(int(v.value.ulVal))
            elif v.valueType == pynvml.NVML_VALUE_TYPE_UNSIGNED_INT:
                row.append(int(v.value.uiVal))
            else:
                row.append(int(v.value.sllVal))
        out.append(row)
    return out


p = w.Problem(N=N, T=K * 0.00125, timesteps=K)
with w.WaveSolver(p, nranks=ng, dims=dims, backend="cuda", devices=list(range(ng))) as s:
    s.run()
    c0 = counters()
    for _ in range(reps):
        r = s.run()
    c1 = counters()
    dec = Decomposition(N, tuple(r.dims))
    alg = []
    for rk in range(ng):
        b, fe = dec.box(rk), dec.face_elems(rk)
        alg.append(sum(8 * fe[f] for f in range(6) if b.nbr[f] >= 0 and b.nbr[f] != rk))
    rows = []
    for g in range(ng):
        tx = (c1[g][0] - c0[g][0]) * 1024 / (reps * (K + 1))  # K updates + the layer-0 push
        rx = (c1[g][1] - c0[g][1]) * 1024 / (reps * (K + 1))
        rows.append({"gpu": g, "nvlink_tx_bytes_per_layer": tx, "nvlink_rx_bytes_per_layer": rx,
                     "algorithmic_halo_bytes_per_layer": alg[g], "tx_over_algorithmic": tx / alg[g] if alg[g] else None})
    print("raw counters before/after:", c0, c1, file=sys.stderr)
    print(json.dumps({"N": N, "steps": K, "gpus": ng, "dims": list(r.dims), "reps": reps, "gcells_per_s": r.gcells_per_s,
                      "multi_layer_launch": s.multistep_active, "per_gpu": rows,
                      "note": "NVML NVLINK_THROUGHPUT_DATA counters (KiB, all links); payload only, protocol overhead "
                              "excluded; 8-byte face stores travel as partial 32-byte sectors"}, indent=1))
