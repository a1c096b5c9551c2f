#!/usr/bin/env python
"""Set-up trace of the CLI (SSN_TRACE_SETUP) on a small simulated pair: python tools/cli_trace.py"""
import os, sys, tempfile
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from somatic_sniper_b200 import cli_tools as ct
w = tempfile.mkdtemp(prefix="ssn_trace_")
pair = ct.make_pair(w, [("chr1", 100000)], 60, 30, seed=1234, n_runs=0)
for i in range(2):
    r = ct.run_cli(ct.CLI, pair["fasta"], pair["tumor"], pair["normal"], os.path.join(w, "o"), [], env={"SSN_TRACE_SETUP": "1"}, stats=os.path.join(w, "s.json"))
    print("run %d wall %.3f s rc %d" % (i, r["wall_s"], r["rc"]))
    print(r["stderr"][-1500:] if isinstance(r["stderr"], str) else r["stderr"].decode()[-1500:])
    print(r["stats"])
