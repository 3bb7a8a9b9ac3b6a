"""Build an A/B variant of librf_b200.so: tools/build_variant.py NAME N[,N...] -DFOO=1 [-DBAR=2 ...]

Recompiles rf_pipe_inst.cu (and rf_fused_inst.cu) for the listed sizes with the extra defines and links
randomfield_b200/librf_NAME.so from those objects plus the regular ones (run the normal
build first).  Select it at run time with RF_LIB=randomfield_b200/librf_NAME.so.
"""
import os, subprocess, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from randomfield_b200 import build as B

name, sizes, defs = sys.argv[1], [int(x) for x in sys.argv[2].split(",")], sys.argv[3:]
B.build()
objs = []
for f in sorted(os.listdir(B.BUILD)):
    if f.endswith(".o") and not f.startswith("var_"):
        objs.append(os.path.join(B.BUILD, f))
for n in sizes:
    obj = os.path.join(B.BUILD, "var_%s_pipe_%d.o" % (name, n))
    _, log = B._compile((os.path.join(B.CSRC, "rf_pipe_inst.cu"), obj, ["-DRF_PIPE_N=%d" % n] + defs, True))
    for line in log.splitlines():
        if "spill" in line and " 0 bytes spill stores" not in line or "error" in line:
            print(line)
    objs = [o for o in objs if os.path.basename(o) != "pipe_%d.o" % n] + [obj]
    if n <= 1024:   # the fused pass 2 + 3 kernels of this ny live in their own unit
        fobj = os.path.join(B.BUILD, "var_%s_fused_%d.o" % (name, n))
        B._compile((os.path.join(B.CSRC, "rf_fused_inst.cu"), fobj, ["-DRF_PIPE_N=%d" % n] + defs, False))
        objs = [o for o in objs if os.path.basename(o) != "fused_%d.o" % n] + [fobj]
out = os.path.join(B.HERE, "librf_%s.so" % name)
subprocess.run([B._nvcc(), "-shared", "-o", out, *objs, "-gencode", "arch=compute_100a,code=sm_100a",
                "-Xcompiler", "-fPIC", "-lcudart"], check=True)
print(out)
