"""Generate optiland_pr_b200/data/glasses.json from the refractiveindex.info YAML files that the
reference vendors (optiland/database/data-nk, CC0 public domain). Data only -- no code is copied.

    python oracle/make_glass_table.py
"""
import json
import os
import sys

import yaml

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DB = "/root/reference/optiland/database/data-nk/glass"
# sample-lens names -> catalogue file (the reference resolves "SK16" to N-SK16, see materials/material.py:216-245)
# which file each name resolves to was read off the reference at run time (Material(name).filename)
GLASSES = {"SK16": "hikari/SK16", "N-SK16": "schott/N-SK16", "F2": "schott/F2",
           "N-SSK2": "schott/N-SSK2", "N-SK2": "schott/N-SK2", "F5": "schott/F5",
           "N-SF11": "schott/N-SF11", "N-BK7": "schott/N-BK7", "SF1": "schott/SF1"}


def main():
    out = {}
    for name, fname in GLASSES.items():
        path = os.path.join(DB, fname + ".yml")
        if not os.path.exists(path):
            print("skip", name, file=sys.stderr)
            continue
        with open(path) as f:
            doc = yaml.safe_load(f)
        entry = {"source": f"refractiveindex.info glass/{fname}.yml (CC0)"}
        for d in doc["DATA"]:
            t = d["type"].strip()
            if t.startswith("formula"):
                entry["formula"] = int(t.split()[1])
                entry["coefficients"] = [float(v) for v in d["coefficients"].split()]
            elif t == "tabulated k":
                rows = [r.split() for r in d["data"].strip().splitlines()]
                entry["k_wavelength"] = [float(r[0]) for r in rows]
                entry["k"] = [float(r[1]) for r in rows]
        out[name] = entry
    dst = os.path.join(ROOT, "optiland_pr_b200", "data", "glasses.json")
    with open(dst, "w") as f:
        json.dump(out, f, indent=1)
    print("wrote", dst, len(out))


if __name__ == "__main__":
    main()
