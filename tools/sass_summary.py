#!/usr/bin/env python
"""Per-kernel SASS evidence: counts of the mnemonics that show what a kernel is built from (TMA bulk copies, packed FP32,
mbarrier, 128-bit loads/stores, match.any ...) from `cuobjdump -sass` of build/*.o, with registers / spills / shared memory
from the ptxas logs (build/*.ptxas.log).  Usage: python tools/sass_summary.py > profiles/r02_sass_summary.txt"""
import glob
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
MNEMONICS = ["UBLKCP", "UTMALDG", "UTMASTG", "SYNCS", "FFMA2", "FFMA", "FMUL", "FADD", "LDG.E.128", "STG.E.128", "LDGSTS", "LDS.128",
             "STS.128", "MATCH.ANY", "REDUX", "ATOMS", "ATOMG", "RED.E", "BAR.SYNC", "HMMA", "UTC"]


def demangle(name):
    try:
        full = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip()
        m = re.search(r"(k_\w+(?:<[^>]*>)?)", full)
        return m.group(1) if m else full[:70]
    except Exception:
        return name[-70:]


def main():
    print("kernel | object | regs | spill st/ld B | smem B | " + " | ".join(MNEMONICS))
    for obj in sorted(glob.glob(os.path.join(ROOT, "build", "*.o"))):
        log = obj[:-2] + ".ptxas.log"
        info = {}
        if os.path.exists(log):
            cur = None
            for line in open(log):
                m = re.search(r"Compiling entry function '(\S+)'", line)
                if m:
                    cur = m.group(1)
                    info[cur] = {}
                m = re.search(r"(\d+) bytes spill stores, (\d+) bytes spill loads", line)
                if m and cur:
                    info[cur]["spill"] = f"{m.group(1)}/{m.group(2)}"
                m = re.search(r"Used (\d+) registers", line)
                if m and cur:
                    info[cur]["regs"] = m.group(1)
                    s = re.search(r"(\d+) bytes smem", line)
                    info[cur]["smem"] = s.group(1) if s else "0 (+dynamic)"
        sass = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout
        for block in sass.split("Function : ")[1:]:
            name = block.split("\n", 1)[0].strip()
            body = block
            counts = []
            for mn in MNEMONICS:
                counts.append(len(re.findall(r"\b" + re.escape(mn) + (r"\b" if mn[-1].isalnum() else ""), body)))
            i = info.get(name, {})
            print(f"{demangle(name)} | {os.path.basename(obj)} | {i.get('regs', '?')} | {i.get('spill', '?')} | {i.get('smem', '?')} | "
                  + " | ".join(str(c) for c in counts))


if __name__ == "__main__":
    main()
