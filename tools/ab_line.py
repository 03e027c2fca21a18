#!/usr/bin/env python
"""one line per A/B variant from its bench JSON"""
import json, sys
d = json.load(open(sys.argv[2]))
print(sys.argv[1], "gcups %.2f" % d["gcups"], "ms %.1f" % d["ms_per_step"], "e2e gcups %.2f" % d["e2e"]["gcups"],
      "pairs", d["config"].get("aligned_pairs_rank0"), d["config"].get("pairs_column_sums_rank0"))
