"""Prints the (shape, pattern) instantiations of shapes.h as S_I_N_pattern tokens (used by the Makefile)."""
import os
import re
import sys


def shapes():
    src = open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "shapes.h")).read()
    quick = len(sys.argv) > 1 and sys.argv[1] == "quick"
    body = src[src.index("#define BMPC_SHAPES"):] if quick else src[src.index("#else"):]
    if quick:
        body = body[:body.index("#else")]
    return [(int(a), int(b), int(c), d) for a, b, c, d in re.findall(r"X\((\d+),\s*(\d+),\s*(\d+),\s*(\w+)\)", body)]


if __name__ == "__main__":
    print(" ".join("%d_%d_%d_%s" % s for s in shapes()))
