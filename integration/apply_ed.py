#!/usr/bin/env python
"""Applies integration/alnmain_b200.ed -- an `ed` script in the format `diff -e` writes (commands bottom-up: `Na`, `Ni`, `N,Mc`,
`N,Md`, text closed by a lone `.`) -- to the reference's alnmain.c.  The script holds only the lines that are ADDED, so no line
of the reference lives in this repository; the line numbers are tied to the pristine file by its sha256.  (`patch -e` would
do the same but needs the `ed` program, which this image lacks.)

    python integration/apply_ed.py /path/to/reference/alnmain.c  out.c
"""
import hashlib
import os
import re
import sys

PRISTINE_SHA256 = "93dd58aec0c4a996d90cac397950e1e8732e0a772336f8c159bd077dd2be1409"     # lisanhu/LongReadMapper alnmain.c
HERE = os.path.dirname(os.path.abspath(__file__))


def apply_ed(lines, script):
    cmds = script.split("\n")
    i = 0
    while i < len(cmds):
        m = re.fullmatch(r"(\d+)(?:,(\d+))?([acid])", cmds[i])
        if not m:
            if cmds[i] == "":
                i += 1
                continue
            raise ValueError(f"bad ed command {cmds[i]!r}")
        a, b, op = int(m.group(1)), int(m.group(2) or m.group(1)), m.group(3)
        i += 1
        text = []
        if op in "aci":
            while cmds[i] != ".":
                text.append(cmds[i] + "\n")
                i += 1
            i += 1
        if op == "a":
            lines[a:a] = text
        elif op == "i":
            lines[a - 1:a - 1] = text
        elif op == "c":
            lines[a - 1:b] = text
        else:
            del lines[a - 1:b]
    return lines


def main(src, dst):
    data = open(src, "rb").read()
    got = hashlib.sha256(data).hexdigest()
    if got != PRISTINE_SHA256:
        raise SystemExit(f"{src}: sha256 {got} is not the alnmain.c this script was written against ({PRISTINE_SHA256})")
    out = apply_ed(data.decode().splitlines(keepends=True), open(os.path.join(HERE, "alnmain_b200.ed")).read())
    open(dst, "w").write("".join(out))


if __name__ == "__main__":
    main(*sys.argv[1:3])
