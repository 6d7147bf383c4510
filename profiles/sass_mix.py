#!/usr/bin/env python
"""Instruction mix of one kernel from `cuobjdump -sass` (usage: sass_mix.py <so> <substring of mangled name>)."""
import collections
import re
import subprocess
import sys

so, pat = sys.argv[1], sys.argv[2]
txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
for block in txt.split("Function : ")[1:]:
    name = block.split("\n", 1)[0]
    if pat not in name:
        continue
    ops = re.findall(r"^\s+/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", block, re.M)
    c = collections.Counter(ops)
    print(name, "total", len(ops))
    print("  " + ", ".join(f"{k}:{v}" for k, v in c.most_common(22)))
