"""Prints the template shapes of shapes.h as S_I_N tokens (used by the Makefile and build())."""
import os
import re

def shapes():
    src = open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "shapes.h")).read()
    line = [l for l in src.splitlines() if l.startswith("#define BMPC_SHAPES")][0]
    return [tuple(int(x) for x in m) for m in re.findall(r"X\((\d+),\s*(\d+),\s*(\d+)\)", line)]

if __name__ == "__main__":
    print(" ".join("%d_%d_%d" % s for s in shapes()))
