"""Count the Blackwell-specific SASS mnemonics per kernel of the built library (no GPU needed):
UTCHMMA = tcgen05.mma, LDTM = tcgen05.ld, UTCBAR = tcgen05.commit, UTMALDG / UBLKCP = TMA tensor / bulk copies,
SYNCS = mbarrier ops, FHFMA = fma.rn.f32.bf16.   python tools/sass_evidence.py > profiles/<round>_sass_mnemonics.txt"""
import collections
import os
import re
import subprocess
import sys

root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(root, "colxlip_b200", "libcolxlip_b200.so")
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
pat = re.compile(r"\b(UTCHMMA|UTCQMMA|UTCBAR|UTCATOMSWS|UTMALDG|UTMASTG|UBLKCP|UBLKPF|UTCCP|LDTM|STTM|SYNCS|FHFMA|HMMA|"
                 r"REDUX|MATCH|ELECT)\b")
fn, cnt = None, collections.OrderedDict()
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        fn = m.group(1)
        cnt[fn] = collections.Counter()
    elif fn:
        cnt[fn].update(pat.findall(line))
print("# cuobjdump -sass", os.path.relpath(lib, root), "(sm_100a): occurrences of Blackwell-specific mnemonics per kernel")
for f, c in cnt.items():
    name = subprocess.run(["c++filt", f], capture_output=True, text=True).stdout.strip()
    print(name[:110])
    print("    " + ("  ".join(f"{k}={v}" for k, v in sorted(c.items())) if c else "(none: plain SIMT kernel)"))
