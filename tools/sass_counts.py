#!/usr/bin/env python
"""SASS mnemonic counts of the tcgen05 GEMM kernels: `python tools/sass_counts.py build/gemm_tc.o > profiles/r0x_sass_gemm.txt`
(cuobjdump -sass; shows that the kernels issue tcgen05.mma / tcgen05.ld / TMA bulk + tensor copies and how many)."""
import re
import subprocess
import sys

obj = sys.argv[1] if len(sys.argv) > 1 else "build/gemm_tc.o"
txt = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True, check=True).stdout
cols = [("UTCHMMA.2CTA", r"UTCHMMA\.2CTA"), ("UTCHMMA(1cta)", r"UTCHMMA(?!\.2CTA)"), ("LDTM", r"\bLDTM"), ("UBLKCP", r"\bUBLKCP"),
        ("UTMASTG", r"\bUTMASTG"), ("UTCBAR", r"\bUTCBAR"), ("SYNCS", r"\bSYNCS"), ("STG.128", r"STG\.E\.128"), ("LDG.128", r"LDG\.E\.128"),
        ("STS.128", r"STS\.128"), ("LDS.128", r"LDS\.128")]
print("# SASS mnemonic counts of the tcgen05 GEMM kernels (cuobjdump -sass %s, sm_100a)" % obj)
print("# UTCHMMA* = tcgen05.mma (TF32 kind), .2CTA = cta_group::2; LDTM = tcgen05.ld (TMEM -> registers); UBLKCP = cp.async.bulk (TMA bulk copy);")
print("# UTMASTG = cp.async.bulk.tensor store (TMA, the NT epilogue's 32x32 boxes); UTCBAR = tcgen05.commit -> mbarrier; SYNCS = mbarrier ops")
print("kernel," + ",".join(c for c, _ in cols))
for m in re.finditer(r"Function : (\S+)\n(.*?)(?=\n\s*Function : |\Z)", txt, re.S):
    name, body = m.group(1), m.group(2)
    short = re.sub(r"^_ZN2pn\d+", "", name)[:34]
    print(short + "," + ",".join(str(len(re.findall(rx, body))) for _, rx in cols))
