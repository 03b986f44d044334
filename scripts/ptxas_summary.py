#!/usr/bin/env python3
"""Summarise an `nvcc -Xptxas -v` log: kernel, registers, stack, spills, smem."""
import re, subprocess, sys

def demangle(n):
    try:
        return subprocess.run(["c++filt", n], capture_output=True, text=True).stdout.strip()
    except Exception:
        return n

def main(path):
    lines = open(path).read().splitlines()
    i = 0
    while i < len(lines):
        m = re.search(r"Compiling entry function '(\S+)'", lines[i])
        if m:
            name = demangle(m.group(1))
            name = re.sub(r"\(.*", "", name).replace("void w3d::", "")
            st = re.search(r"(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads", lines[i + 2])
            us = re.search(r"Used (\d+) registers.*?(?:(\d+) bytes smem)?$", lines[i + 3])
            print(f"{name:34s} regs={us.group(1):>3s} stack={st.group(1):>3s} spill_st={st.group(2):>3s} spill_ld={st.group(3):>3s} smem={us.group(2)}")
            i += 3
        i += 1

if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else "build/wave3d_kernels.ptxas.log")
