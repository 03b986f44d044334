#!/usr/bin/env python3
"""One-line digest of a bench.py JSON output file."""
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    if "unavailable" in d:
        print("unavailable:", d["unavailable"][:200])
    else:
        c = d.get("config", {})
        print(f"{d.get('impl')} value={d['value']:.2f} ms/step={d['ms_per_step']:.4f} e2e={(d.get('e2e') or {}).get('value')} "
              f"par={c.get('parallelism')} tr={str(c.get('transport'))[:6]} clk={d.get('clocks')}")
except Exception as e:  # noqa: BLE001
    print("unparsed:", e)
