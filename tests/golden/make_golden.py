#!/usr/bin/env python
"""Generates the committed fixtures of tests/golden/.

Provenance (read before trusting a file):
  * fk_known_answers.json, panda_pairs.json -- NOT produced by code: hand-derived from data/panda/panda.urdf:31-223 and
    data/panda/panda.srdf:44-79; this script only writes the literals below to disk.
  * c1_scene{A,B}.npz, c2_self_20k.npz, c5_two_panda_4k.npz -- outputs of the CPU ORACLE (oracle/oracle.c, double
    precision) on the seeded inputs of SURVEY 8(d).  The reference itself cannot be built or imported in this container
    (DESIGN.md section 2), so these are regression fixtures of the restatement ("parity unpinned"), not reference outputs.
    They travel to the GPU box, where the CUDA path is compared against them without running the oracle.

Run from the repo root:  python tests/golden/make_golden.py
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

FK_KNOWN = {
    "source": "hand-derived from data/panda/panda.urdf:31-223 (SURVEY.md 8c (1))",
    "cases": [
        {"qpos": [0, 0, 0, 0, 0, 0, 0, 0, 0],
         "links": {"panda_link8": {"p": [0.088, 0.0, 0.926], "R_diag": [1.0, -1.0, -1.0]},
                   "panda_link7": {"p": [0.088, 0.0, 1.033]},
                   "panda_hand": {"p": [0.088, 0.0, 0.926], "R_row0": [0.707107, 0.707107, 0.0]},
                   "panda_leftfinger": {"p": [0.088, 0.0, 0.8676]},
                   "panda_rightfinger": {"p": [0.088, 0.0, 0.8676]}}},
        {"qpos": [0, 0.2, 0, -2.6, 0, 3.0, 0.8, 0, 0],
         "links": {"panda_link8": {"p": [0.457507, 0.0, 0.204750]},
                   "panda_hand": {"R_row0": [0.979962, -0.014310, 0.198669], "R_row2": [0.198648, -0.002901, -0.980067]},
                   "panda_leftfinger": {"p": [0.469110, 0.0, 0.147514]},
                   "panda_rightfinger": {"p": [0.469110, 0.0, 0.147514]}}},
    ],
}

_L = "panda_link"
PAIRS_19 = {
    "source": "55 unordered pairs -> 46 (URDF parent/child, src/fcl_model.cpp:282-290) -> 19 (36 SRDF disable_collisions, "
              "data/panda/panda.srdf:44-79); SURVEY.md 8 preamble",
    "pairs": [[_L + "0", _L + "5"], [_L + "0", _L + "6"], [_L + "0", _L + "7"], [_L + "0", "panda_hand"],
              [_L + "0", "panda_leftfinger"], [_L + "0", "panda_rightfinger"],
              [_L + "1", _L + "5"], [_L + "1", _L + "6"], [_L + "1", _L + "7"], [_L + "1", "panda_hand"],
              [_L + "1", "panda_leftfinger"], [_L + "1", "panda_rightfinger"],
              [_L + "2", _L + "5"], [_L + "2", _L + "7"], [_L + "2", "panda_hand"], [_L + "2", "panda_leftfinger"],
              [_L + "2", "panda_rightfinger"], [_L + "5", "panda_leftfinger"], [_L + "5", "panda_rightfinger"]],
}


def _pack(desc, q32):
    """oracle run -> dict of arrays (pair masks packed little-endian along the pair axis)"""
    from oracle import oracle as O

    qd = q32.astype(np.float64)
    v, ph = O.collide_batch(desc, qd, threads=os.cpu_count() or 1)
    names = np.array([f"{desc.pair_kinds[p]}|{desc.object_names[desc.pair_a[p]]}|{desc.object_names[desc.pair_b[p]]}"
                      for p in range(desc.n_pairs)])
    return {"q": q32, "verdict": np.packbits(v.astype(np.uint8), bitorder="little"),
            "pairs": np.packbits(ph.astype(np.uint8), axis=1, bitorder="little"), "pair_names": names,
            "n_pairs": np.int32(desc.n_pairs)}


def main():
    from tests import workloads as WL
    from tests.common import compile_desc, make_panda_world, random_q

    with open(os.path.join(HERE, "fk_known_answers.json"), "w") as f:
        json.dump(FK_KNOWN, f, indent=1)
    with open(os.path.join(HERE, "panda_pairs.json"), "w") as f:
        json.dump(PAIRS_19, f, indent=1)
    for scene in "AB":
        w = WL.c1_world(scene)
        np.savez_compressed(os.path.join(HERE, f"c1_scene{scene}.npz"), **_pack(compile_desc(w), WL.c1_q(10000).astype(np.float32)))
    w = make_panda_world()
    np.savez_compressed(os.path.join(HERE, "c2_self_20k.npz"), **_pack(compile_desc(w), random_q(20000, 1).astype(np.float32)))
    wl = WL.c5(4096)
    np.savez_compressed(os.path.join(HERE, "c5_two_panda_4k.npz"), **_pack(wl["desc"], wl["make_q"](0, 4096)))
    for fn in sorted(os.listdir(HERE)):
        print(fn, os.path.getsize(os.path.join(HERE, fn)))


if __name__ == "__main__":
    main()
