"""gpurun_out/label_parity.jsonl (written by tests/label_gate.py during `pytest -m gpu`) -> profiles/r02_label_parity.md"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
src = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "label_parity.jsonl")
recs = [json.loads(l) for l in open(src) if l.strip()]
print("# Label parity of the GPU path against the oracle (round 2)\n")
print("Gate (tests/label_gate.py, BASELINE.json north_star): GPU labels equal the oracle's float64 direct solve except where the")
print("ORACLE's own two largest potentials are closer than 1e-6; every record below is one gated comparison of `pytest -m gpu`")
print("(solver tolerance 1e-6, default solver path: multilevel-preconditioned PCG).\n")
print("| comparison | vertices | PCG iterations | max rel. residual | mismatching vertices | largest oracle top-2 gap |")
print("|---|---|---|---|---|---|")
tot = 0
for r in recs:
    tot += r["mismatches"]
    print("| %s | %d | %s | %s | %d | %s |" % (r["tag"], r["V"], r.get("iterations", "-"),
          ("%.2e" % r["max_rel_residual"]) if "max_rel_residual" in r else "-", r["mismatches"],
          ("%.3e" % r["max_gap"]) if r["mismatches"] else "-"))
print("\n%d comparisons, %d mismatching vertices in total." % (len(recs), tot))
