#!/usr/bin/env python
"""The C4-like pipeline on one GPU at a reduced photon count, for launch lists. Usage: python tools/prof_c4.py [casted_total] [iters]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as graft  # noqa: E402

casted = int(sys.argv[1]) if len(sys.argv) > 1 else 16_000_000
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 1
pkg = graft.load_package()
pipelines = graft.load_package_module("pipelines")
ctx = pkg.Context(0)
r = pipelines.run_c4(pkg, ctx, None, 0, 1, casted_total=casted, iters=iters)
print({k: round(v, 2) for k, v in r.items() if k.endswith("_ms")})
