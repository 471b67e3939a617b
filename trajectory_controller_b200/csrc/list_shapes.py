"""Prints the (shape, pattern) instantiations of shapes.h as S_I_N_pattern tokens (used by the Makefile)."""
import os
import re


def shapes():
    src = open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "shapes.h")).read()
    body = src[src.index("#define BMPC_SHAPES"):]
    return [(int(a), int(b), int(c), d) for a, b, c, d in re.findall(r"X\((\d+),\s*(\d+),\s*(\d+),\s*(\w+)\)", body)]


if __name__ == "__main__":
    print(" ".join("%d_%d_%d_%s" % s for s in shapes()))
