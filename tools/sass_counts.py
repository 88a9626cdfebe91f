"""Static SASS mnemonic counts per kernel of libdq4ml.so -> profiles/<round>_sass_counts.json

    python tools/sass_counts.py r2

Counts the instructions that show which hardware path a kernel uses: UTCHMMA / UTCQMMA (tcgen05.mma), LDTM / STTM
(TMEM), UTMALDG (TMA tensor-map load), UBLKCP (cp.async.bulk), LDGSTS (cp.async), SYNCS (mbarrier), DMMA (fp64 tensor
core), plus LDG / STG / ATOM / RED / BAR / SHFL for orientation.
"""
import collections
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WANT = ("UTCHMMA", "UTCQMMA", "UTCMMA", "LDTM", "STTM", "UTMALDG", "UBLKCP", "LDGSTS", "SYNCS", "DMMA", "LDG", "STG", "ATOM", "RED",
        "BAR", "SHFL", "REDUX")


def main():
    rnd = sys.argv[1] if len(sys.argv) > 1 else "r2"
    so = os.path.join(ROOT, "net.jgp.labs.sparkdq4ml_b200", "libdq4ml.so")
    out = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True, check=True).stdout
    kernels = collections.OrderedDict()
    cur = None
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            kernels[cur] = collections.Counter()
            continue
        if cur is None:
            continue
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
        if m:
            op = m.group(1)
            if op in WANT:
                kernels[cur][op] += 1
    demangled = {}
    try:
        names = "\n".join(kernels.keys())
        dem = subprocess.run(["c++filt"], input=names, capture_output=True, text=True, check=True).stdout.splitlines()
        demangled = dict(zip(kernels.keys(), dem))
    except Exception:
        pass
    doc = {"how": "cuobjdump -sass net.jgp.labs.sparkdq4ml_b200/libdq4ml.so, instruction mnemonics counted per kernel (static counts); tools/sass_counts.py",
           "kernels": {demangled.get(k, k)[:240]: dict(v) for k, v in kernels.items() if v}}
    path = os.path.join(ROOT, "profiles", f"{rnd}_sass_counts.json")
    with open(path, "w") as f:
        json.dump(doc, f, indent=1)
    print(path, len(doc["kernels"]), "kernels")
    for k, v in doc["kernels"].items():
        if any(x in v for x in ("UTCHMMA", "UTMALDG", "UBLKCP", "DMMA")):
            print(" ", k[:110], dict(v))


if __name__ == "__main__":
    main()
