"""Build an experiment variant of the library: python scripts/build_variant.py NAME -DFOO=1 ... -> arcquant_b200/libarcq_NAME.so"""
import os, subprocess, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from arcquant_b200 import build as b
name, defs = sys.argv[1], sys.argv[2:]
objdir = os.path.join(os.path.dirname(b.LIB), "build", "variant_" + name)
os.makedirs(objdir, exist_ok=True)
procs, objs = [], []
for entry in b.SOURCES:
    src, extra, oname = entry if isinstance(entry, tuple) else (entry, [], entry.replace(".cu", ".o"))
    obj = os.path.join(objdir, oname)
    procs.append(subprocess.Popen([b._nvcc()] + b.NVCC_FLAGS + extra + defs + ["-c", os.path.join(b.CSRC, src), "-o", obj]))
    objs.append(obj)
assert all(p.wait() == 0 for p in procs)
out = os.path.join(os.path.dirname(b.LIB), "libarcq_%s.so" % name)
subprocess.check_call([b._nvcc(), "-shared", "-Wno-deprecated-gpu-targets", "-o", out] + objs)
print(out)
