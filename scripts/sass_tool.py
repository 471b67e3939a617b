#!/usr/bin/env python3
"""SASS inspection helper (developer tool): opcode histogram / spill locations of one kernel inside an object or .so.

  python scripts/sass_tool.py <file.o|.so> <substring of the mangled kernel name> [--hist] [--spills] [--dump]
"""
import re
import subprocess
import sys
from collections import Counter


def functions(path):
    out = subprocess.run(["cuobjdump", "-sass", path], capture_output=True, text=True, check=True).stdout
    cur, body = None, {}
    for line in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1); body[cur] = []
            continue
        m = re.match(r"\s*/\*([0-9a-f]{4,6})\*/\s+(.*?);\s*/\* 0x", line)
        if m and cur:
            body[cur].append((int(m.group(1), 16), m.group(2).strip()))
    return body


def opcode(txt):
    t = txt.split()
    if t[0].startswith("@"):
        t = t[1:]
    return t[0].split(".")[0]


def main():
    path, pat = sys.argv[1], sys.argv[2]
    flags = set(sys.argv[3:]) or {"--hist"}
    for name, ins in functions(path).items():
        if not all(p in name for p in pat.split("&")):
            continue
        print("==", name, len(ins), "instructions")
        if "--hist" in flags:
            c = Counter(opcode(t) for _, t in ins)
            print("  " + "  ".join(f"{k}:{v}" for k, v in c.most_common(40)))
        if "--spills" in flags:
            for a, t in ins:
                if opcode(t) in ("LDL", "STL"):
                    print(f"  {a:06x} {t}")
        if "--dump" in flags:
            for a, t in ins:
                print(f"  {a:06x} {t}")


if __name__ == "__main__":
    main()
