import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import golden_util as gu, oracle_lib as ol, parity
name = sys.argv[1] if len(sys.argv) > 1 else "clustered4k_periodic"
g = gu.load(name)
cells, tasks, nt = gu.tree_structs(g)
n = len(g["h"])
out = parity.run_gpu(g["ic"], g["params"], record=True, pair_capacity=n * 400, tree=(cells, tasks, nt))
prm = ol.make_params(record=True, **g["params"])
o = ol.run_step("oracle", prm, g["ic"], cells, tasks, nt, record=True)
A = set(map(tuple, out["pairs"][2].tolist())); B = set(map(tuple, o["pairs"][2].tolist()))
print("gpu", len(out["pairs"][2]), "unique", len(A), "oracle", len(B))
ids = g["id"]; pos = {int(v): k for k, v in enumerate(ids)}
leaf_of = np.zeros(n, dtype=int)
for c in range(len(cells)):
    if not cells[c].split:
        leaf_of[cells[c].first:cells[c].first + cells[c].count] = c
from collections import Counter
cnt = Counter(map(tuple, out["pairs"][2].tolist()))
dups = [(p, c) for p, c in cnt.items() if c > 1]
print("duplicates:", dups[:20])
for (i, j) in sorted(A - B)[:20]:
    a, b = pos[i], pos[j]
    dx = g["x"][a] - g["x"][b]; dx -= np.round(dx)
    r2 = float((dx ** 2).sum())
    hi, hj = out["h"][a], out["h"][b]
    print("extra", i, j, "leaf", leaf_of[a], leaf_of[b], "depth", cells[leaf_of[a]].depth, cells[leaf_of[b]].depth,
          "r2", r2, "hig2", float(hi) ** 2 * 3.3333, "hjg2", float(hj) ** 2 * 3.3333, "rev in ref", (j, i) in B)
print("missing", sorted(B - A)[:10])
