#!/usr/bin/env python
"""dev tool: host-side split of a complete config-1 search (example alignment, -a 1) with LVB_GPU_PROFILE=1"""
import os, subprocess, sys, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import cli_harness as H
td = tempfile.mkdtemp()
open(os.path.join(td, "infile"), "wb").write(H.example_input())
env = dict(os.environ, LVB_GPU_PROFILE=os.environ.get("LVB_GPU_PROFILE", "1"))
p = subprocess.run([H.CLI, "-s", "12345", "-a", "1"] + sys.argv[1:], cwd=td, capture_output=True, text=True, env=env)
print("\n".join(l for l in p.stdout.splitlines() if any(k in l for k in ("Wall", "Host time", "trees scored", "Rearrangements", "length"))))
print(p.stderr[-900:])
