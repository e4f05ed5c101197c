"""Per-kernel counts of the SASS mnemonics that prove the Blackwell paths (cuobjdump -sass of the
built library): UTC*MMA = tcgen05.mma, LDTM = tcgen05.ld, UTCBAR = tcgen05.commit, DMMA = FP64 tensor
cores (mma.sync.m8n8k4.f64), LDGSTS = cp.async.   python tools/sass_evidence.py > profiles/r02_sass_evidence.txt"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(ROOT, "keypointmoseq_b200", "libkpms_b200.so")
out = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
demangle = lambda n: subprocess.run(["c++filt", n], capture_output=True, text=True).stdout.strip()
pat = re.compile(r"\b(UTC[A-Z]*MMA|LDTM|STTM|UTCBAR|UTMALDG|UTMASTG|DMMA|HMMA|LDGSTS|DFMA|FFMA|SHFL|MUFU)\b")
cur, counts, first = None, collections.OrderedDict(), {}
for line in out.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        cur = m.group(1)
        counts[cur] = collections.Counter()
        continue
    if cur is None:
        continue
    m = pat.search(line)
    if m:
        counts[cur][m.group(1)] += 1
        first.setdefault((cur, m.group(1)), line.strip()[:110])
print("# cuobjdump -sass keypointmoseq_b200/libkpms_b200.so : mnemonic counts per kernel (kernels with tensor-core or cp.async instructions)")
for fn, c in counts.items():
    if not any(k in c for k in ("UTCHMMA", "UTCQMMA", "UTCIMMA", "LDTM", "DMMA", "LDGSTS", "UTCBAR")) and not any(k.startswith("UTC") for k in c):
        continue
    name = demangle(fn)
    name = re.sub(r"\(.*", "", name)
    print("%-70s %s" % (name[:70], "  ".join("%s=%d" % kv for kv in sorted(c.items()))))
print("\n# one example line per (kernel, mnemonic) for the tensor-core instructions")
for (fn, mn), line in first.items():
    if mn.startswith("UTC") or mn in ("LDTM", "DMMA"):
        print("%-45s %s" % (re.sub(r"\(.*", "", demangle(fn))[:45], line))
