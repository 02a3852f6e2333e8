"""Extracts the diag lines (STEP, KINETIC_ENRG, BAROTR_KE, MAX_ADV_CFL, MAX_VERT_CFL) of the reference's Rivers_ana regression
log into tests/golden/rivers_ana_diag.json.  Run in the build container, where /root/reference exists:
    python tests/golden/make_rivers_ana_diag.py [Rivers_ana|Pipes_ana]
"""
import json
import os

import sys
CASE = sys.argv[1] if len(sys.argv) > 1 else "Rivers_ana"
SRC = f"/root/reference/tests/{CASE}/benchmark.result_github_gnu"
rows = []
for n, l in enumerate(open(SRC), 1):
    f = l.split()
    if len(f) == 5 and f[0].isdigit() and "E" in f[1]:
        rows.append({"step": int(f[0]), "line": n, "KE": float(f[1]), "BAROTR_KE": float(f[2]), "MAX_ADV_CFL": float(f[3]), "MAX_VERT_CFL": float(f[4])})
json.dump({"source": f"tests/{CASE}/benchmark.result_github_gnu", "made_by": "tests/golden/make_rivers_ana_diag.py", "diag": rows},
          open(os.path.join(os.path.dirname(os.path.abspath(__file__)), CASE.lower() + "_diag.json"), "w"), indent=1)
print(len(rows), "diag lines")
