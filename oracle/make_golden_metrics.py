"""Generate tests/golden/ndcg.json by running the UNMODIFIED reference post_utils.py (:45-128).

TEST INFRASTRUCTURE ONLY.  Needs /root/reference.  post_utils imports matplotlib, gseapy and plotly at
module level (all absent here and unused by the numeric functions): they are stubbed with empty modules.

    python oracle/make_golden_metrics.py [--ref /root/reference]
"""
import argparse
import json
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ref", default="/root/reference")
    a = ap.parse_args()
    for name in ("matplotlib", "gseapy", "plotly", "plotly.express", "plotly.io", "plotly.graph_objects"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.path.insert(0, a.ref)
    from node2vec2rank import post_utils as pu
    import joblib
    # loky workers would re-import post_utils without the stubs: run the reference's Parallel(...) on threads
    ctx = joblib.parallel_backend("threading")
    ctx.__enter__()
    rng = np.random.default_rng(7)
    cases = []
    for n, m in [(10, 3), (257, 40), (1000, 106), (1000, 0), (3000, 1), (64, 64)]:
        items = [f"g{i}" for i in range(n)]
        ranked = [items[i] for i in rng.permutation(n)]
        relevant = sorted(rng.choice(items, size=m, replace=False).tolist()) + (["not_there"] if m else [])
        perms = [[items[i] for i in rng.permutation(n)] for _ in range(12)]
        ndcg = pu.relevance_and_ndcg(ranked, relevant)
        rel_vec = [1 if x in relevant else 0 for x in ranked]
        cases.append({
            # item names are "g<i>": rankings are stored as the integers i
            "n": n, "ranked": [int(x[1:]) for x in ranked], "relevant": relevant,
            "permutations": [[int(x[1:]) for x in p] for p in perms],
            "ndcg": float(ndcg), "dcg": float(pu.discounted_cumulative_gain(rel_vec)),
            "ndcg_from_vector": float(pu.normalized_discounted_cumulative_gain(rel_vec, binary_relevance=True)),
            "perm_ndcg": [float(pu.relevance_and_ndcg(p, relevant)) for p in perms],
            "p_value": float(pu.parallel_empirical_p_ndcg(ndcg, perms, relevant)),
            "p_value_vs_median": float(pu.parallel_empirical_p_ndcg(
                float(np.median([pu.relevance_and_ndcg(p, relevant) for p in perms])), perms, relevant)),
        })
    out = os.path.join(ROOT, "tests", "golden", "ndcg.json")
    json.dump({"cases": cases}, open(out, "w"))
    print("wrote", out, os.path.getsize(out), "bytes")


if __name__ == "__main__":
    main()
