#!/usr/bin/env python
"""TEST INFRASTRUCTURE.  Dump the BS.2051 layout / facet DATA out of the compiled reference
(oracle/_ref) into oracle/data/bs2051.json for the numpy restatement (oracle/ear_oracle.py),
and the HOA t-design points into oracle/data/tdesign_5200.npy.  Build container only."""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
from oracle import ref  # noqa: E402

out = {}
for name in ref.layout_names():
    info = ref.layout_info(name)
    out[name] = {
        "names": info["names"], "is_lfe": [bool(x) for x in info["is_lfe"]],
        "pos": info["pos"].tolist(), "nominal": info["nominal"].tolist(),
        "facets": ref.facets(name) if name != "0+2+0" else None,
    }
here = os.path.dirname(__file__)
json.dump(out, open(os.path.join(here, "data", "bs2051.json"), "w"), indent=0)
np.save(os.path.join(here, "data", "tdesign_5200.npy"), ref.hoa_points())
print("wrote oracle/data/bs2051.json, tdesign_5200.npy")
