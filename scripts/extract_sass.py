#!/usr/bin/env python3
"""Write the SASS of selected kernels (cuobjdump) + an opcode histogram that proves the Blackwell
paths (UTMALDG = cp.async.bulk.tensor, SYNCS = mbarrier, DFMA/DADD/DMUL = fp64 pipe).
usage: extract_sass.py <object or .so> <substring of the mangled/demangled kernel name> <out.sass>"""
import collections, re, subprocess, sys

obj, pat, out = sys.argv[1], sys.argv[2], sys.argv[3]
txt = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout
blocks = re.split(r"(?=\n\s*Function : )", txt)
sel = [b for b in blocks if re.search(r"Function : (\S+)", b) and pat in
       subprocess.run(["c++filt", re.search(r"Function : (\S+)", b).group(1)], capture_output=True, text=True).stdout]
with open(out, "w") as f:
    for b in sel:
        name = re.search(r"Function : (\S+)", b).group(1)
        dem = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip()
        ops = collections.Counter()
        lines = []
        for ln in b.splitlines():
            m = re.match(r"\s+/\*([0-9a-f]{4,6})\*/\s+(.*?);\s*/\*", ln)
            if m:
                ins = m.group(2)
                lines.append(f"/*{m.group(1)}*/  {ins}")
                mm = re.match(r"(@!?U?P\d+\s+)?([A-Z0-9_]+)", ins.strip())
                if mm:
                    ops[mm.group(2)] += 1
        f.write(f"// {dem}\n// {len(lines)} instructions; opcode histogram: " +
                ", ".join(f"{k}:{v}" for k, v in ops.most_common()) + "\n")
        f.write("\n".join(lines) + "\n\n")
print(f"{len(sel)} kernel(s) -> {out}")
