#!/usr/bin/env python
"""Summarise `nvcc -Xptxas -v` output: registers, stack and spills per kernel.

usage: python tools/ptxas_report.py build.log
"""
import re
import subprocess
import sys


def main(path):
    rows, name = [], None
    for line in open(path):
        m = re.search(r"Compiling entry function '(\S+)'", line)
        if m:
            name = m.group(1)
            stack = spill = None
            continue
        m = re.search(r"(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads", line)
        if m and name:
            stack, spill = m.group(1), f"{m.group(2)}/{m.group(3)}"
            continue
        m = re.search(r"Used (\d+) registers", line)
        if m and name:
            rows.append((name, m.group(1), stack, spill))
            name = None
    if not rows:
        print("no kernels found")
        return
    out = subprocess.run(["c++filt"] + [r[0] for r in rows], capture_output=True, text=True,
                         stdin=subprocess.DEVNULL, timeout=60).stdout.splitlines()
    for dem, (_, regs, stack, spill) in zip(out, rows):
        dem = re.sub(r"\(anonymous namespace\)::", "", dem)
        dem = re.sub(r"^void ", "", dem).split("(")[0]
        print(f"{dem[:84]:84s} regs={regs:>3s} stack={stack} spill={spill}")


if __name__ == "__main__":
    main(sys.argv[1])
