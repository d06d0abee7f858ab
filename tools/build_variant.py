"""Development: build a variant of libcnx_b200.so with extra -D flags on gemm2.cu / dwconv7.cu into
gpurun_in/libcnx_<name>.so (git-ignored; travels to the GPU box) for same-box A/B runs through
CNX_LIB_PATH.   python tools/build_variant.py NAME [-DFLAG ...] [--src gemm2.cu]"""
import os, subprocess, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from image_classification_b200 import build as B

name = sys.argv[1]
srcs, flags = [], []
it = iter(sys.argv[2:])
for a in it:
    if a == "--src":
        srcs.append(next(it))
    else:
        flags.append(a)
srcs = srcs or ["gemm2.cu"]
B.build_library()
out_dir = os.path.join(os.path.dirname(B.HERE), "gpurun_in")
os.makedirs(out_dir, exist_ok=True)
objs = []
for s in B.SOURCES:
    o = os.path.join(B.OBJ_DIR, s.replace(".cu", ".o"))
    if s in srcs:
        o = os.path.join(out_dir, f"{name}_{s.replace('.cu', '.o')}")
        cmd = [B._nvcc(), *[f for f in B.NVCC_FLAGS if f not in ("-Xptxas", "-v")], *flags, "-c",
               os.path.join(B.CSRC, s), "-o", o]
        subprocess.run(cmd, check=True)
    objs.append(o)
lib = os.path.join(out_dir, f"libcnx_{name}.so")
subprocess.run([B._nvcc(), "-shared", "-o", lib, *objs, "-lcudart_static", "-ldl", "-lpthread", "-lrt"], check=True)
print(lib)
