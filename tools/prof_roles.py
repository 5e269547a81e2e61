"""Per-role phase timers of the fused kernel (diagnostics build: python cnmf_b200/build.py --prof).
usage: CNMF_LIB=cnmf_b200/libcnmf_b200_prof.so python tools/prof_roles.py f64|f32 [CNMF_TUNE string] [m]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ.setdefault("CNMF_LIB", os.path.join(ROOT, "cnmf_b200", "libcnmf_b200_prof.so"))
if len(sys.argv) > 2 and sys.argv[2]:
    os.environ["CNMF_TUNE"] = sys.argv[2] + ",GRAPH=0"
else:
    os.environ["CNMF_TUNE"] = "GRAPH=0"
import numpy as np
import torch
from cnmf_b200 import Plan, _lib, synth

dt = np.float32 if sys.argv[1] == "f32" else np.float64
m = int(sys.argv[3]) if len(sys.argv) > 3 else 301
n, k = 1_000_000, 6
code = _lib.F32 if dt == np.float32 else _lib.F64
tdt = torch.float32 if dt == np.float32 else torch.float64
ld = _lib.padded_ld(m, code)
dev = torch.device("cuda", 0)
X = torch.empty((n, ld), dtype=tdt, device=dev)
V = torch.empty_like(X)
hs = torch.from_numpy(synth.basis(k, m, 0)).to(dev)
_lib.check(_lib.load().cnmf_synth(code, X.data_ptr(), V.data_ptr(), n, 0, m, ld, k, hs.data_ptr(), 0, 3, 0.05, None))
rng = np.random.default_rng(1)
W = torch.from_numpy(rng.random((n, k)).astype(dt)).to(dev)
H = torch.from_numpy(rng.random((k, m)).astype(dt)).to(dev)
hist = torch.zeros(8, dtype=torch.float64, device=dev)
plan = Plan(n, m, k, dt, ld=ld)
print(sys.argv[1:], plan.info, flush=True)
print("--- launch 1: pass over (W0, H0), role W only copies; launches 2-3: full iterations", flush=True)
plan.solve(X, V, W, H, hist, 2)
torch.cuda.synchronize()
