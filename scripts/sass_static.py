#!/usr/bin/env python
"""Static opcode histogram of one kernel in the built .so (cuobjdump -sass), e.g.
   python scripts/sass_static.py 'demod_blocked_kernelILi7ELb0ELb0ELb1E'   -- per 16 samples/thread in the main loop."""
import subprocess, sys, collections, re
so = "lora-sdr-lightweight-standalone-library-_b200/liblora_phy_b200.so"
txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
pat = sys.argv[1]
blocks = txt.split("Function : ")
for b in blocks[1:]:
    name = b.split("\n", 1)[0]
    if pat not in name: continue
    h = collections.Counter()
    n = 0
    for line in b.splitlines():
        m = re.match(r"\s+/\*[0-9a-f]{4,5}\*/\s+(@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m:
            op = m.group(2).split(".")[0]
            h[op] += 1; n += 1
    print(name[:100], "total", n)
    print("  " + "  ".join(f"{k}:{v}" for k, v in h.most_common(45)))
