"""Digest of the SASS of every kernel in libb2mm.so: the mnemonics that prove which hardware paths are used.

    python scripts/sass_digest.py > profiles/r1_sass_digest.txt
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "pytorch_distributed_matmul_benchmark_b200", "libb2mm.so")
KEY = re.compile(r"^(UTCHMMA|UTCQMMA|UTCMMA|UTCIMMA|UTCBAR|UTMALDG|UTMASTG|UTMAPF|UTMACMDFLUSH|UTCATOMSWS|LDTM|STTM|UGETNEXTWORKID|"
                 r"SYNCS|LDGMC|STG\.E\.\w*MMLS|REDG|RED|ATOMG|ACQBULK|PREEXIT|UCGABAR|CGAERRBAR|LDL|STL|FFMA|FADD|HMMA|LDGSTS|LDSM)")

sass = subprocess.check_output(["cuobjdump", "-sass", LIB], text=True, stderr=subprocess.DEVNULL)
demangled = {}
try:
    names = re.findall(r"Function : (\S+)", sass)
    out = subprocess.check_output(["cu++filt"] + sorted(set(names)), text=True).splitlines()
    demangled = dict(zip(sorted(set(names)), out))
except (OSError, subprocess.CalledProcessError):
    pass

print(f"source: cuobjdump -sass {os.path.relpath(LIB, ROOT)} (sm_100a); per kernel: instruction count and the mnemonics that matter")
print("UTCHMMA = tcgen05.mma (kind::f16 / tf32), .2CTA = cta_group::2; UTMALDG/UTMASTG = TMA tensor load/store; LDTM = tcgen05.ld;")
print("UTCBAR = tcgen05.commit; UGETNEXTWORKID = clusterlaunchcontrol.try_cancel; LDGMC = multimem.ld_reduce; ACQBULK/PREEXIT =")
print("griddepcontrol.wait / launch_dependents; LDL/STL = local-memory traffic: in these kernels only the argument buffers of the")
print("watchdog printf calls and of the out-of-line ragged-N tail (ptxas -v: 0 bytes of spill stores / loads for every kernel).\n")
for block in sass.split("Function : ")[1:]:
    name = block.split("\n", 1)[0].strip()
    ops = collections.Counter()
    n = 0
    for line in block.splitlines():
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Za-z0-9_.]*)", line)
        if not m:
            continue
        n += 1
        k = KEY.match(m.group(1))
        if k:
            ops[m.group(1) if m.group(1).startswith(("UTC", "UTMA", "LDTM", "LDGMC", "UGETNEXT")) else k.group(1)] += 1
    pretty = demangled.get(name, name)
    pretty = re.sub(r"\(CUtensorMap_st.*", "(...)", pretty)
    print(f"{pretty[:170]}\n    {n} instructions; " + ", ".join(f"{k} x{v}" for k, v in sorted(ops.items())) + "\n")
