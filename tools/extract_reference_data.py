#!/usr/bin/env python
"""Derive the data tables the product bundles from a checkout of the reference repository:

    python tools/extract_reference_data.py [/root/reference]

  <pkg>/data/exo_v3_geometry.json   kinematic tree of Simulation/exo_v3.urdf (our own parser, urdf.parse_urdf)
  <pkg>/data/reference_motions.npz  columns 0-4 of Simulation/reference_motions/ref_motion_{0..7}.txt

These are measurements/geometry (data), not code; the benchmark configs in BASELINE.json are defined on them.
"""
import importlib
import json
import os
import sys

import numpy as np

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
PKG = "a-deep-reinforcement-learning-enabled-soft-exoskeleton-for-parkinson-s-patients_b200"


def main():
    ref = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
    urdf = importlib.import_module(PKG + ".urdf")
    motions = importlib.import_module(PKG + ".motions")
    data = os.path.join(REPO, PKG, "data")
    os.makedirs(data, exist_ok=True)
    tree = urdf.parse_urdf(os.path.join(ref, "Simulation", "exo_v3.urdf"))
    with open(os.path.join(data, "exo_v3_geometry.json"), "w") as f:
        json.dump(tree, f, indent=1)
    ms = motions.load_motions(os.path.join(ref, "Simulation", "reference_motions"))
    np.savez_compressed(os.path.join(data, "reference_motions.npz"), **{"motion_%d" % i: m for i, m in enumerate(ms)})
    print("links:", len(tree["links"]), "motions:", [m.shape[1] for m in ms])


if __name__ == "__main__":
    main()
