"""GPU check of the register-resident role W (CNMF_TUNE=WREG=1): parity against the oracle on
shapes that have an instance (and odd row counts: partial tiles / chunks), then prints the plan."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import bringup_parity as fl  # noqa: E402
from cnmf_b200 import Plan  # noqa: E402

cases = [(5000, 301, 6, np.float64, 30), (5003, 301, 6, np.float32, 30), (4097, 301, 4, np.float64, 20),
         (3001, 301, 2, np.float64, 20), (2999, 300, 5, np.float64, 10), (40001, 184, 8, np.float32, 10),
         (7001, 301, 8, np.float32, 10), (1237, 290, 1, np.float64, 10), (9, 301, 6, np.float64, 5)]
for tune in ("WREG=1", "WREG=1,CR=4", "WREG=1,CR=1", "WREG=1,NOPROD=2"):
    os.environ["CNMF_TUNE"] = tune
    for (n, m, k, dt, it) in cases:
        p = Plan(n, m, k, dt)
        v = p.info["variant"]
        p.close()
        print("tune=%s variant=%d" % (tune, v), end="  ", flush=True)
        try:
            fl.parity(n, m, k, dt, it)
        except Exception as exc:  # noqa: BLE001
            print("FAILED %r" % (exc,), flush=True)
    if tune != "WREG=1":
        cases = cases[:3]
