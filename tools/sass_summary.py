"""Per-kernel SASS mnemonic counts of the built library (no GPU needed):
    python tools/sass_summary.py > profiles/r02_sass_summary.txt
Proves which kernels carry tcgen05 (UTCHMMA / UTCQMMA, LDTM / STTM), TMA (UTMALDG / UTMASTG / UBLKCP), cp.async (LDGSTS), packed fp32
FMAs (FFMA2) and MUFU, straight from `cuobjdump -sass` of weclip-vit-comer_b200/libweclip_sm100.so."""
import collections
import os
import re
import subprocess
import sys

LIB = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "weclip-vit-comer_b200", "libweclip_sm100.so")
KEYS = ["UTCHMMA", "UTCQMMA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "LDGSTS", "SYNCS", "FFMA2", "FFMA", "MUFU", "HMMA",
        "LDS", "STS", "LDG", "STG", "RED", "ATOM", "STL", "LDL"]


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    per = collections.OrderedDict()
    cur = None
    for line in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            name = re.sub(r"\(.*", "", name).replace("void ", "").replace("wc::", "") or m.group(1)
            cur = per.setdefault(name, collections.Counter())
            cur["_n"] += 0
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m and cur is not None:
            op = m.group(1)
            cur["_n"] += 1
            for k in KEYS:
                if op == k or op.startswith(k + "."):
                    cur[k] += 1
                    break
    arch = subprocess.run(["cuobjdump", "-lelf", LIB], capture_output=True, text=True).stdout
    print("# cuobjdump -sass of libweclip_sm100.so; ELF images:", ", ".join(sorted(set(re.findall(r"sm_\d+a?", arch)))))
    print("# instruction counts per kernel (static SASS, all template instances listed)")
    cols = [k for k in KEYS if any(c[k] for c in per.values())]
    print(f"{'kernel':64s} {'instrs':>7s} " + " ".join(f"{k:>7s}" for k in cols))
    for name, c in sorted(per.items(), key=lambda kv: kv[0]):
        print(f"{name[:64]:64s} {c['_n']:7d} " + " ".join(f"{c[k]:7d}" for k in cols))


if __name__ == "__main__":
    sys.exit(main())
