#!/usr/bin/env python3
"""oracle/gen_stubs.py -- TEST INFRASTRUCTURE (part of the oracle/_ref build recipe).

Emits a C file with aborting stubs for every MR:: symbol that the pre-linked reference objects
leave undefined.  Those symbols are referenced only by Mesh methods that the hot path never calls
(hole filling, point-cloud trees, IO helpers ...); stubbing them lets libmrref.so load with RTLD_NOW.
Usage: gen_stubs.py prelink.so > stubs.c
"""
import subprocess
import sys

out = subprocess.run(["nm", "-D", "-u", sys.argv[1]], capture_output=True, text=True, check=True).stdout
syms = []
for line in out.splitlines():
    parts = line.split()
    if not parts:
        continue
    s = parts[-1].split("@")[0]
    if "2MR" in s[:12] and s.startswith("_Z"):
        syms.append(s)
print("#include <stdio.h>\n#include <stdlib.h>")
for i, s in enumerate(sorted(set(syms))):
    if s.startswith(("_ZTV", "_ZTI", "_ZTS")):
        print(f'char mrref_stub_data_{i}[512] __asm__("{s}");')
        print(f"char mrref_stub_data_{i}[512] = {{0}};")
    else:
        print(f'void mrref_stub_{i}(void) __asm__("{s}");')
        print(f'void mrref_stub_{i}(void) {{ fprintf(stderr, "oracle/_ref: unlinked reference symbol called: {s}\\n"); abort(); }}')
