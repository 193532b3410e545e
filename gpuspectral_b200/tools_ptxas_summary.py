"""Summarises the -Xptxas -v logs written by the Makefile (registers / stack / spills per kernel)."""
import glob
import os
import re
import subprocess
import sys

here = os.path.dirname(os.path.abspath(__file__))
for path in sorted(glob.glob(os.path.join(here, "build", "*.ptxas.log"))):
    text = open(path).read()
    blocks = text.split("Compiling entry function ")[1:]
    for b in blocks:
        name = b.split("'")[1]
        try:
            name = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip()
        except OSError:
            pass
        name = re.sub(r"\(.*", "", name).replace("gsb::(anonymous namespace)::", "").replace("void ", "")
        regs = re.search(r"Used (\d+) registers", b)
        stack = re.search(r"(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads", b)
        print(f"{os.path.basename(path)[:-10]:10s} {name[:44]:44s} regs={regs.group(1):>3s} stack={stack.group(1):>4s} spill_st={stack.group(2)} spill_ld={stack.group(3)}")
