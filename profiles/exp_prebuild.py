"""Prebuilds (NVRTC, no GPU needed) the specialised kernel of a bench workload under the current OMO_* environment into the in-tree
cubin cache, so that profiles/exp_kernel.py does not compile on the GPU box.   python profiles/exp_prebuild.py tree64"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
if os.environ.get("OMO_EXP_PKG_ROOT"):
    sys.path.insert(0, os.path.abspath(os.environ["OMO_EXP_PKG_ROOT"]))
import bench  # noqa: E402
from optimo_b200.core import Core  # noqa: E402

name = sys.argv[1]
fk = bench.W_TRAJ.get(name, 0.0) != 0.0
t0 = time.time()
ok, log = Core(bench.make_scene(name)).probe_specialised_kernel(fk=fk, count=False, dump_prefix=(sys.argv[2] if len(sys.argv) > 2 else None))
print(name, {k: v for k, v in os.environ.items() if k.startswith("OMO_")}, "ok" if ok else "FAILED", f"{time.time() - t0:.1f} s", log[:400])
