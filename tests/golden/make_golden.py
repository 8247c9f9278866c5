"""Generates tests/golden/*.json from the reference tree (run in the build container only:
/root/reference does not exist on the GPU box).

  python tests/golden/make_golden.py

dec_tree_rules_kat.json  -- known answers extracted from the one result artefact the reference ships
                            for the tree path, resource/dec_tree_rules.json:
                            per decision path its predicate strings, population, infoContent (gini),
                            classValPr in file order, and the integer class counts they imply.
schemas.json             -- the four tree-path schemas shipped under resource/ (field metadata only),
                            used to pin split enumeration counts (App. C items 5-6).
"""
import json
import os

REF = "/root/reference/resource"
OUT = os.path.dirname(os.path.abspath(__file__))


def main():
    model = json.load(open(os.path.join(REF, "dec_tree_rules.json")))
    paths = []
    for p in model["decisionPaths"]:
        pop = p["population"]
        counts = {k: round(v * pop) for k, v in p["classValPr"].items()}
        assert sum(counts.values()) == pop
        paths.append({
            "predicateStrs": [q["predicateStr"] for q in p["predicates"]],
            "population": pop,
            "infoContent": p["infoContent"],
            "stopped": p["stopped"],
            "classValPr": list(p["classValPr"].items()),      # file order = HashMap order
            "classCounts": counts,
            "pathFields": sorted(p.keys()),
            "predicateFields": sorted(p["predicates"][-1].keys()),
            "lastPredicate": p["predicates"][-1],
        })
    json.dump({"source": "resource/dec_tree_rules.json", "paths": paths}, open(os.path.join(OUT, "dec_tree_rules_kat.json"), "w"), indent=1)
    schemas = {}
    for name in ("call_hangup.json", "teleComChurn.json", "emailCampaign.json", "patient.json"):
        schemas[name] = json.load(open(os.path.join(REF, name)))
    json.dump(schemas, open(os.path.join(OUT, "schemas.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
