"""Compile the oracle's C restatement (oracle/raycast_bruteforce.c) into oracle/_build/ with gcc.
The reference itself is pure Python: there is nothing to compile into oracle/_ref/ (DESIGN.md "Oracle")."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_build")
LIB = os.path.join(OUT, "liboracle_raycast.so")


def build(force=False):
    src = os.path.join(HERE, "raycast_bruteforce.c")
    os.makedirs(OUT, exist_ok=True)
    if force or not os.path.exists(LIB) or os.path.getmtime(LIB) < os.path.getmtime(src):
        subprocess.check_call(["gcc", "-O3", "-fopenmp", "-shared", "-fPIC", "-o", LIB, src, "-lm"])
    return LIB


if __name__ == "__main__":
    print(build())
