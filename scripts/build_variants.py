"""Developer tool: builds libdli.so variants with -D experiment knobs into build/ (shipped to the GPU box by
gpurun, ignored by git).  Usage: python scripts/build_variants.py name=-DFOO=1,-DBAR ..."""
import os, subprocess, sys
from concurrent.futures import ThreadPoolExecutor
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from delaunay_linterp_b200 import build as b
os.makedirs(os.path.join(ROOT, "build"), exist_ok=True)
def one(spec):
    name, _, flags = spec.partition("=")
    out = os.path.join(ROOT, "build", "libdli_%s.so" % name)
    cmd = [b.nvcc_path()] + b.NVCC_FLAGS + [f for f in flags.split(",") if f] + ["-I", os.path.join(ROOT, "include"), "-o", out] + \
          [os.path.join(b.CSRC, s) for s in b.SOURCES]
    r = subprocess.run(cmd, capture_output=True, text=True)
    regs = [l for l in (r.stdout + r.stderr).splitlines() if "k_interp2_fatILb0" in l or "Used" in l]
    info = ""
    lines = (r.stdout + r.stderr).splitlines()
    for i, l in enumerate(lines):
        if "Compiling entry function" in l and "k_interp2_fatILb0" in l:
            info = " | ".join(x.strip() for x in lines[i + 1:i + 4])
    return name, r.returncode, info
with ThreadPoolExecutor(4) as ex:
    for name, rc, info in ex.map(one, sys.argv[1:]):
        print(name, "rc=%d" % rc, info, flush=True)
