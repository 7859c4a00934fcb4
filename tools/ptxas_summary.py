#!/usr/bin/env python
"""Registers / stack / spills per kernel from the -Xptxas -v logs csrc/Makefile keeps next to the objects."""
import glob
import os
import re
import subprocess
import sys

d = sys.argv[1] if len(sys.argv) > 1 else os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "vermilion_b200", "csrc")
for f in sorted(glob.glob(os.path.join(d, "*.ptxas.log"))):
    txt = open(f).read()
    for m in re.finditer(r"Compiling entry function '([^']+)'[^\n]*\n[^\n]*\n\s*(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads\n[^\n]*Used (\d+) registers", txt):
        name = subprocess.run(["c++filt", m.group(1)], stdout=subprocess.PIPE, text=True).stdout.strip()
        name = re.sub(r"\(.*", "", name)
        print(f"{os.path.basename(f)[:-12]:12s} {name[:60]:60s} regs {m.group(5):>3s} stack {m.group(2):>5s} spill st/ld {m.group(3)}/{m.group(4)}")
