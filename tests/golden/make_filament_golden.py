"""Regenerates tests/golden/filament_diag.json from the reference's own regression logs
(run in the build container, where /root/reference exists)."""
import json, re
out = {}
for tag in ("github_gnu", "github_ifx"):
    rows = []
    for l in open(f"/root/reference/tests/Filament/benchmark.result_{tag}"):
        m = re.match(r"^\s*(\d+)\s+([-0-9.E+]+)\s+([-0-9.E+]+)\s+([-0-9.E+]+)\s+([-0-9.E+]+)\s*$", l)
        if m:
            rows.append([m.group(i) for i in range(2, 6)])
    assert len(rows) == 21
    out[tag] = rows
json.dump({"source": "tests/Filament/benchmark.result_github_gnu:146-167 and benchmark.result_github_ifx:142-163 "
                     "(KINETIC_ENRG, BAROTR_KE, MAX_ADV_CFL, MAX_VERT_CFL; steps 0..20), copied as decimal strings",
           "nfast": 82, **out}, open("tests/golden/filament_diag.json", "w"), indent=1)
