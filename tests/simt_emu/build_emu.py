"""Compile the product's .cu sources against the SIMT emulator (tests only, CPU)."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
CSRC = os.path.join(ROOT, "finitedifference_b200", "csrc")
OUT = os.path.join(HERE, "_burgers_emu.so")


def build(sources=("runtime.cu", "fom.cu", "ops.cu", "rom.cu", "rom_grid.cu", "gram.cu")):
    srcs = [os.path.join(CSRC, s) for s in sources]
    deps = srcs + [os.path.join(CSRC, "common.cuh"), os.path.join(HERE, "cuda_emu.h"),
                   os.path.join(ROOT, "include", "burgers_b200.h")]
    deps += [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cuh")]
    if os.path.exists(OUT) and all(os.path.getmtime(d) <= os.path.getmtime(OUT) for d in deps):
        return OUT
    objs = []
    for s in srcs:
        o = os.path.join(HERE, "_" + os.path.basename(s) + ".o")
        subprocess.run(["g++", "-std=c++20", "-O2", "-g", "-DBFD_EMULATE", "-I" + HERE, "-fPIC",
                        "-x", "c++", "-c", s, "-o", o], check=True)
        objs.append(o)
    subprocess.run(["g++", "-shared", "-o", OUT, *objs, "-pthread"], check=True)
    return OUT
