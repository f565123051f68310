#!/usr/bin/env python
"""ingest of a huge PHYLIP file.  Writes a synthetic n x m alignment (default 2000 x 1M =
2 GB of text), then times the reference's own reader + set-up (CReadFiles reads the file three times, then
matchange: oracle/_ref, ref_msa_load) against this repository's host reader (lvb_b200/host/data.cpp: one pass,
then matchange) on the same file; both must agree on n, m after matchange and the packed word count.
Host-only work; run on the GPU box for comparable cores.  Prints one JSON line."""
import json, os, sys, tempfile, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from lvb_b200 import host, synthetic

n = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
m = int(sys.argv[2]) if len(sys.argv) > 2 else 1_000_000
with tempfile.TemporaryDirectory() as td:
    path = os.path.join(td, "infile")
    t0 = time.time()
    synthetic.write_phylip(path, synthetic.survey_alignment(n, m, 1, 0.02))
    out = {"taxa": n, "sites": m, "file_bytes": os.path.getsize(path), "write_s": round(time.time() - t0, 2)}
    t0 = time.time()
    ours = host.HostMSA.load(path, 0)
    out["lvb_b200_host_reader_s"] = round(time.time() - t0, 3)
    out["dims_ours"] = [int(ours.n), int(ours.m), int(ours.nwords)]
    ours.close()
    import oracles
    if oracles.have_ref():
        t0 = time.time()
        ref = oracles.RefMSA.load(path, 0, threads=1)
        out["reference_reader_s"] = round(time.time() - t0, 3)
        out["dims_reference"] = [int(ref.n), int(ref.m), int(ref.nwords)]
        out["agree"] = out["dims_ours"] == out["dims_reference"]
print(json.dumps(out))
