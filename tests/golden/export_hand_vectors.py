#!/usr/bin/env python3
"""Writes tests/golden/hand_vectors.json: the hand-typed vectors of hand_vectors.py in the hex/JSON form that
oracle/go_crosscheck/main.go replays (same field names as adapter_vectors.json.gz).  Run after editing hand_vectors.py."""
import json
import os

import hand_vectors as hv

HERE = os.path.dirname(os.path.abspath(__file__))
out = []
for name, payload, start, status, frame in hv.ADAPTER:
    rec = {"name": "hand_" + name, "in": payload.hex(), "start": start, "status": status, "frame": frame.hex(), "nobridge": True}
    if name == "G5_escapes_go121":
        rec["frame122"] = hv.G5_GO122_FRAME.hex()
    out.append(rec)
for k, (payload, status, subject, js) in enumerate(hv.BRIDGE):
    out.append({"name": "hand_bridge_%d" % k, "in": payload.hex(), "start": 0, "bstatus": status, "subject": subject.hex(),
                "json": js.hex(), "noadapter": True})
with open(os.path.join(HERE, "hand_vectors.json"), "w") as f:
    json.dump(out, f, indent=0)
print("wrote %d vectors" % len(out))
