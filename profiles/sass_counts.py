"""Opcode counts of the hot kernels from the built library (cuobjdump -sass), the evidence that the code the tests and the
bench load is what DESIGN.md describes: bulk copies (UBLKCP), tcgen05 (UTCHMMA / LDTM / UTCBAR), REDUX, the integer dot
products of the uint8 rows (IDP.4A), cluster / DSMEM traffic (UCGABAR, ST.E with .shared::cluster shows as STAS / SYNCS),
spills (STL / LDL).   python profiles/sass_counts.py > profiles/sass_r2.txt   (no GPU needed)"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "parallelizing-support-vector-machine-training-with-gpu-and-mpi_b200", "libsvmb200.so")
HOT = ["smo_persistent_kernelIhLi5ELi1ELi256", "smo_persistent_kernelIfLi7ELi4ELi256", "smo_persistent_kernelIdLi13ELi1ELi256",
       "decision_tc_pipelined_kernel", "decision_kernelIdLi13", "decision_kernelIfLi7ELi4ELb1", "peer_rendezvous_kernel", "gather_raw_rows_kernel",
       "u8_pack_kernel", "u8_norms_kernel"]
WATCH = ["UBLKCP", "UTCHMMA", "LDTM", "UTCBAR", "UTCATOMSWS", "REDUX", "IDP", "DFMA", "DADD", "PRMT", "LDS", "SHFL", "SYNCS", "STAS", "UCGABAR", "MAPA",
         "MEMBAR", "BAR", "LDG", "STG", "LD", "ST", "STL", "LDL", "MUFU", "F2F", "I2F"]

out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
funcs, name = collections.defaultdict(list), None
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        name = m.group(1)
        continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and name:
        funcs[name].append(m.group(1))
head = subprocess.run(["git", "-C", ROOT, "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()
print(f"# cuobjdump -sass {os.path.relpath(LIB, ROOT)} (sm_100a), HEAD {head}; opcode counts per kernel (static: instructions in the binary)")
for key in HOT:
    for fn, ops in funcs.items():
        if key in fn:
            c = collections.Counter(o.split(".")[0] for o in ops)
            full = collections.Counter(ops)
            print(f"\n{fn}\n  instructions {len(ops)}")
            print("  " + "  ".join(f"{w} {c[w]}" for w in WATCH if c[w]))
            extra = [f"{k} {v}" for k, v in full.items() if k.startswith(("IDP.4A", "UBLKCP", "UTCHMMA", "LDTM", "REDUX", "SYNCS", "UTCBAR", "UCGABAR", "CGABAR"))]
            if extra:
                print("  " + "  ".join(sorted(extra)))
            break
