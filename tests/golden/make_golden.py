"""Writes tests/golden/oracle_streams.json: size + SHA-256 of the oracle's stream for every
(corpus file, quality, lgwin, mode).  Run in the build container after the oracle has been
validated (round-trips through both Google C decoders): python tests/golden/make_golden.py"""
import hashlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import _oracle as O  # noqa: E402
from conftest import CORPUS, SYNTH, load_input  # noqa: E402

gold = {}
for name in CORPUS + SYNTH:
    data = load_input(name)
    for q in (1, 2, 3, 4, 5, 9, 10, 11):
        for lgwin, mode in ((22, 0), (22, 2), (24, 1)):
            s = O.encode(data, q, lgwin, mode)
            assert O.decode(s, len(data)) == data
            gold[f"{name}|{q}|{lgwin}|{mode}"] = [len(s), hashlib.sha256(s).hexdigest()]
json.dump(gold, open(os.path.join(ROOT, "tests", "golden", "oracle_streams.json"), "w"), indent=0, sort_keys=True)
print(len(gold), "entries")
