"""Run a command in its own process group and SIGKILL the whole group after a time limit (torchrun workers
outlive a killed launcher otherwise).  usage: python tools/run_pg.py SECONDS cmd args..."""
import os
import signal
import subprocess
import sys

limit = float(sys.argv[1])
p = subprocess.Popen(sys.argv[2:], start_new_session=True)
try:
    rc = p.wait(timeout=limit)
except subprocess.TimeoutExpired:
    os.killpg(p.pid, signal.SIGKILL)
    p.wait()
    print(f"run_pg: killed process group after {limit:.0f} s", file=sys.stderr)
    rc = 124
sys.exit(rc)
