#!/usr/bin/env python
"""Dumps the SASS of one kernel of a built library and prints instruction statistics.

    python tools/sass_kernel.py <lib.so> [substring of the mangled kernel name] [--dump]
"""
import collections
import re
import subprocess
import sys


def kernel_sass(so, needle):
    txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
    out, on = [], False
    for ln in txt.splitlines():
        m = re.match(r"\s*Function : (\S+)", ln)
        if m:
            on = needle in m.group(1)
            if on:
                out.append(("#", m.group(1)))
            continue
        if on:
            m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
            if m:
                out.append((m.group(1), m.group(2)))
    return out


def main():
    so = sys.argv[1]
    needle = sys.argv[2] if len(sys.argv) > 2 and not sys.argv[2].startswith("--") else "render_kernelILi5ELi0ELb0ELb0ELb0"
    ins = kernel_sass(so, needle)
    body = [i for i in ins if i[0] != "#"]
    ops = collections.Counter(re.sub(r"^@!?U?P\d+\s+", "", t).split()[0].split(".")[0] for _, t in body)
    print(f"{[t for a, t in ins if a == '#']}: {len(body)} SASS instructions")
    print(", ".join(f"{k} {v}" for k, v in ops.most_common(28)))
    if "--dump" in sys.argv:
        for a, t in ins:
            print(a, t)


if __name__ == "__main__":
    main()
