"""Extracts the start-up block of the reference's Rivers_ana regression log (mode-splitting line, Cs_w table, setup_grid2
report) into tests/golden/rivers_ana_startup.json.  Run in the build container, where /root/reference exists:
    python tests/golden/make_rivers_ana_startup.py
"""
import json
import os
import re

SRC = "/root/reference/tests/Rivers_ana/benchmark.result_github_gnu"
L = open(SRC).read().splitlines()
g = {"source": "tests/Rivers_ana/benchmark.result_github_gnu", "made_by": "tests/golden/make_rivers_ana_startup.py"}
for n, l in enumerate(L, 1):
    m = re.search(r"ndtfast =\s*(\d+)\s+nfast =\s*(\d+)", l)
    if m:
        g["ndtfast"], g["nfast"], g["nfast_line"] = int(m.group(1)), int(m.group(2)), n
    if l.strip().startswith("level") and "Cs-curve" in l:
        tab = []
        for k in range(11):
            f = L[n + k].split()
            tab.append([int(f[0]), float(f[1]), float(f[2])])
        g["Cs_w"] = tab
        g["Cs_w_lines"] = [n + 1, n + 11]
    m = re.search(r"hmin =\s*(\S+)\s+grdmin =\s*(\S+)\s+area =\s*(\S+)", l)
    if m:
        g["hmin"], g["grdmin"], g["area"] = map(float, m.groups())
        g["stats_line"] = n
    m = re.search(r"hmax =\s*(\S+)\s+grdmax =\s*(\S+)\s+volume =\s*(\S+)", l)
    if m:
        g["hmax"], g["grdmax"], g["volume"] = map(float, m.groups())
    m = re.search(r"Cg_max =\s*(\S+)\s+Cg_min =\s*(\S+)\s+Cu_Cor =\s*(\S+)", l)
    if m:
        g["Cg_max"], g["Cg_min"], g["Cu_Cor"] = map(float, m.groups())
json.dump(g, open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "rivers_ana_startup.json"), "w"), indent=1)
