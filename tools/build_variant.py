"""Tuning helper (not product): build the library with extra -D defines into tools/_variants/<name>.so.
    python tools/build_variant.py lb32 -DTDFO_SORT_LB_BATCH=32
Time a variant with TDFO_TUNE_LIB=tools/_variants/lb32.so python tools/tune_update.py"""
import subprocess, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from tdfo_b200 import build as b

name, defines = sys.argv[1], sys.argv[2:]
out_dir = Path(__file__).resolve().parent / "_variants"
obj_dir = out_dir / ("obj_" + name)
obj_dir.mkdir(parents=True, exist_ok=True)
objs = []
procs = []
for src in b._sources():
    obj = obj_dir / (src.stem + ".o")
    objs.append(obj)
    procs.append(subprocess.Popen([b._nvcc(), *b.NVCC_FLAGS, *defines, "-I", str(b.INCLUDE), "-c", str(src), "-o", str(obj)]))
assert all(p.wait() == 0 for p in procs)
lib = out_dir / (name + ".so")
subprocess.check_call([b._nvcc(), "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", str(lib), *map(str, objs)])
print(lib)
