"""Register / spill / shared-memory report of every kernel in libmpcb.so (nvcc -Xptxas -v), and optionally the SASS instruction
histogram of the step kernels (cuobjdump -sass).  usage: python scripts/ptxas_info.py [--sass] [extra nvcc flags]"""
import os, re, subprocess, sys, collections
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
from mpc_for_av_at_intersection_b200 import _lib
args = [a for a in sys.argv[1:] if a != "--sass"]
srcs = [os.path.join(_lib.CSRC, f) for f in sorted(os.listdir(_lib.CSRC)) if f.endswith(".cu")]
out = "/tmp/mpcb_ptxas_info.so"
r = subprocess.run(["nvcc"] + _lib.NVCC_FLAGS + args + ["-Xptxas", "-v", "-o", out] + srcs, capture_output=True, text=True)
if r.returncode:
    sys.exit(r.stderr[-3000:])
name = None
for ln in r.stderr.splitlines():
    m = re.search(r"Compiling entry function '(\S+)' for 'sm_100a'", ln)
    if m:
        name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        name = re.sub(r"\(.*", "", name)
        continue
    if name and "bytes stack frame" in ln:
        stack = ln.strip()
    m = re.search(r"Used (\d+) registers.*", ln)
    if m and name:
        print(f"{name:70s} regs {m.group(1):>3s} | {stack}")
        name = None
if "--sass" in sys.argv:
    sass = subprocess.run(["cuobjdump", "-sass", out], capture_output=True, text=True).stdout
    fn, hist = None, collections.defaultdict(collections.Counter)
    for ln in sass.splitlines():
        m = re.search(r"Function : (\S+)", ln)
        if m:
            fn = re.sub(r"\(.*", "", subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip())
            continue
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", ln)
        if m and fn:
            hist[fn][m.group(1).split(".")[0]] += 1
    for fn, h in hist.items():
        if "mpc_step" in fn or "coupling" in fn:
            tot = sum(h.values())
            print(f"\n{fn}: {tot} SASS instructions; " + " ".join(f"{k}:{v}" for k, v in h.most_common(14)))
