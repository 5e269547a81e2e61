"""Hang-safe tuning sweep: runs `first_light.py --sweep` in a child process and restarts it
after the config that made it stall (a deadlocked kernel must not eat the GPU budget).
usage: python tools/sweep.py [--n ROWS] [--m M] [--k K] "f64:RP1=2;f32:MAXW=16,C2=1" """
import os
import subprocess
import sys
import threading
import time

HERE = os.path.dirname(os.path.abspath(__file__))
args = sys.argv[1:]
extra = []
while args and args[0].startswith("--"):
    extra += args[:2]
    args = args[2:]
cfgs = [c for c in args[0].split(";") if c]
STALL = 45.0
i = 0
while i < len(cfgs):
    p = subprocess.Popen([sys.executable, os.path.join(HERE, "first_light.py")] + extra + ["--sweep", ";".join(cfgs[i:])],
                         stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    last = [time.time()]
    done = [0]

    def reader():
        for line in p.stdout:
            last[0] = time.time()
            if line.startswith("timing") or line.startswith("FAILED"):
                done[0] += 1
            print(line, end="", flush=True)

    t = threading.Thread(target=reader, daemon=True)
    t.start()
    first = True
    while p.poll() is None:
        time.sleep(0.5)
        limit = STALL + (60.0 if first and done[0] == 0 else 0.0)  # torch import on a fresh box
        if time.time() - last[0] > limit:
            print("STALLED on %s -- killing" % cfgs[i + done[0]], flush=True)
            p.kill()
            p.wait()
            done[0] += 1
            break
    t.join(timeout=2)
    if p.returncode not in (None, 0) and done[0] == 0:
        done[0] = 1
    i += max(done[0], 1) if p.returncode != 0 else len(cfgs)
