"""Extracts the 51-row S-coordinate table (N = 50, theta_s = theta_b = 6, hc = 25 m) and the mode-splitting line of the reference's
Pipes_real regression log into tests/golden/pipes_real_startup.json -- a start-up pin at the level count of the reference's
realistic cases (the time stepping of that case needs netCDF input that is not in the tree).  Run in the build container:
    python tests/golden/make_pipes_real_startup.py
"""
import json
import os
import re

SRC = "/root/reference/tests/Pipes_real/benchmark.result_github_gnu"
L = open(SRC).read().splitlines()
g = {"source": "tests/Pipes_real/benchmark.result_github_gnu", "made_by": "tests/golden/make_pipes_real_startup.py"}
for n, l in enumerate(L, 1):
    m = re.search(r"ndtfast =\s*(\d+)\s+nfast =\s*(\d+)", l)
    if m:
        g["ndtfast"], g["nfast"], g["nfast_line"] = int(m.group(1)), int(m.group(2)), n
    for key in ("theta_s", "theta_b", "hc"):
        m = re.match(r"\s*%s =\s*(\S+)" % key, l)
        if m:
            g[key] = float(m.group(1))
    if l.strip().startswith("level") and "Cs-curve" in l:
        tab = []
        k = 0
        while True:
            f = L[n + k].split()
            if len(f) < 3 or not f[0].isdigit():
                break
            tab.append([int(f[0]), float(f[1]), float(f[2])])
            k += 1
        g["Cs_w"] = tab
        g["Cs_w_lines"] = [n + 1, n + k]
json.dump(g, open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "pipes_real_startup.json"), "w"), indent=1)
