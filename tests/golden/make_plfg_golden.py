#!/usr/bin/env python
"""Generates tests/golden/plfg_golden.json by running the reference's OWN Python golden model of the PLFG,
PLFG() in /root/reference/generators/plfg/src/main/python/plfg_script.py:21-57, on a set of configurations.

The reference file imports matplotlib (absent here) and plots at import time, so only the source of the PLFG function is
executed: the text between `def PLFG` and the first top-level statement after it is exec'd unmodified.  Nothing from the
reference is copied into this repository; only the function's OUTPUT vectors are stored.

Run in the build container (where /root/reference exists):  python tests/golden/make_plfg_golden.py
"""
import json
import os
import re

REF = "/root/reference/generators/plfg/src/main/python/plfg_script.py"
HERE = os.path.dirname(os.path.abspath(__file__))


def load_reference_plfg():
    src = open(REF).read()
    m = re.search(r"^def PLFG .*?(?=^\S)", src, flags=re.S | re.M)
    ns = {}
    exec(compile(m.group(0), REF, "exec"), ns)
    return ns["PLFG"]


CASES = [
    # the module's own demo (plfg_script.py:61-68)
    dict(name="script_demo", startingPoint=0, numOfFrames=2, numOfChirps=2, numOfSegments=[2, 3],
         numOfSamples=[[30, 5], [40, 10, 20]], slopes=[[5, 0], [10, 0, -8]],
         segmentResets=[[False, True], [False, False, False]], interFrameNumOfSamples=50),
    # PLFGSpec.scala:42-62 (ramp 24 / reset 6 / triangle 32 up, 20 down)
    dict(name="plfgspec_kat", startingPoint=0, numOfFrames=1, numOfChirps=1, numOfSegments=[4],
         numOfSamples=[[24, 6, 32, 20]], slopes=[[2, 0, 10, -8]], segmentResets=[[False, True, False, False]],
         interFrameNumOfSamples=60),
    # RspChainVanillaTester.scala:86-94: one segment, slope 0, startingPoint 16 -> constant frequency word
    dict(name="constant_tone", startingPoint=16, numOfFrames=1, numOfChirps=1, numOfSegments=[1], numOfSamples=[[255]],
         slopes=[[0]], segmentResets=[[False]], interFrameNumOfSamples=10),
    dict(name="sawtooth_3frames", startingPoint=7, numOfFrames=3, numOfChirps=3, numOfSegments=[1, 2, 4],
         numOfSamples=[[100], [50, 25], [10, 20, 30, 40]], slopes=[[3], [-2, 4], [1, -1, 2, -2]],
         segmentResets=[[False], [False, False], [False, True, False, True]], interFrameNumOfSamples=17),
    dict(name="negative_start", startingPoint=-100, numOfFrames=2, numOfChirps=1, numOfSegments=[3],
         numOfSamples=[[64, 2, 64]], slopes=[[11, 0, -11]], segmentResets=[[False, False, False]],
         interFrameNumOfSamples=5),
]


def main():
    plfg = load_reference_plfg()
    out = []
    for c in CASES:
        kw = {k: v for k, v in c.items() if k != "name"}
        y = plfg(**kw)
        out.append(dict(c, output=[int(v) for v in y]))
    with open(os.path.join(HERE, "plfg_golden.json"), "w") as f:
        json.dump(out, f, separators=(",", ":"))
    print("wrote", len(out), "cases:", [(c["name"], len(c["output"])) for c in out])


if __name__ == "__main__":
    main()
