"""Registers / spills / static shared memory per kernel from rampack_b200/build.log (ptxas -v)."""
import re, subprocess, sys, os
log = open(os.path.join(os.path.dirname(__file__), "..", "rampack_b200", "build.log")).read()
pat = sys.argv[1] if len(sys.argv) > 1 else "k_sweep"
ents = re.findall(r"Compiling entry function '([^']+)' for 'sm_100a'\n(?:.*\n)*?.*?(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads\n.*?Used (\d+) registers(?:, used \d+ barriers)?(?:, (\d+) bytes smem)?", log)
for name, stack, ss, sl, regs, smem in ents:
    dn = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip()
    if pat in dn:
        print(f"{dn[:60]:60s} regs={regs} stack={stack} spill={ss}/{sl} smem={smem}")
