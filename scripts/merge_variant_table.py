#!/usr/bin/env python
"""Install a freshly measured per-world variant table (bench/sweep.py --write-table) into
torch_distlearn_b200/parallel/variant_table.json under its world size, keeping the other worlds' entries."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
src, world = sys.argv[1], str(int(sys.argv[2]))
dst = os.path.join(ROOT, "torch_distlearn_b200", "parallel", "variant_table.json")
new = json.load(open(src))
table = json.load(open(dst)) if os.path.exists(dst) else {}
table[world] = {k: v for k, v in new.items() if k != "world"}
json.dump(table, open(dst, "w"), indent=1, sort_keys=True)
print("installed world", world, "->", dst)
