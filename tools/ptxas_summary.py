#!/usr/bin/env python
"""Summarise `nvcc -Xptxas -v` logs (easydl_b200/csrc/*.ptxas.log): registers, spills, smem per kernel."""
import glob, re, subprocess, sys, os

def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.split("\n")
    return out

def main():
    here = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "easydl_b200", "csrc")
    rows = []
    for log in sorted(glob.glob(os.path.join(here, "*.ptxas.log"))):
        cur = None; spill = ""
        for line in open(log):
            m = re.search(r"Compiling entry function '([^']+)'", line)
            if m: cur = m.group(1); spill = ""; continue
            m = re.search(r"(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads", line)
            if m: spill = "stack=%s spill=%s/%s" % m.groups(); continue
            m = re.search(r"Used (\d+) registers(?:, used (\d+) barriers)?(?:, (\d+) bytes smem)?", line)
            if m and cur:
                rows.append((os.path.basename(log).split(".")[0], cur, int(m.group(1)), m.group(3) or "0", spill)); cur = None
    names = demangle([r[1] for r in rows])
    for (f, _, regs, smem, spill), n in zip(rows, names):
        n = re.sub(r"\(anonymous namespace\)::", "", n)
        n = re.sub(r"\(.*$", "", n)
        print(f"{f:8s} regs={regs:3d} smem={smem:>6s} {spill:28s} {n[:110]}")

if __name__ == "__main__":
    main()
