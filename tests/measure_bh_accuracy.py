#!/usr/bin/env python
"""Barnes-Hut accuracy table (BH criterion ii): RMS and max relative acceleration error against the fp64
direct sum for theta in {0.3, 0.5, 0.8}, ours (LBVH, group-box acceptance test) next to the oracle of the reference's
INTENDED octree algorithm (per-body test, src/cuda/force_barnes_hut.cu:154-196 with leaf splitting), same bodies, same
eps.  A measurement script, not a test (tests/test_gpu_sort_tree.py asserts the relation at theta = 0.5).
usage: python tests/measure_bh_accuracy.py [n]   -> one JSON line per (distribution, theta)"""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))  # run from anywhere


def main():
    import nbody_b200 as nb
    from nbody_b200 import ics
    from oracle import bindings as orc
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 50000
    for name, st in (("plummer", ics.plummer(n, seed=12)), ("uniform_sphere", ics.uniform_sphere(n, 1.0, 1.0 / n, seed=12))):
        truth = orc.direct_f64(st["pos_x"], st["pos_y"], st["pos_z"], st["mass"], 1.0, 1e-4)
        tn = np.sqrt(sum(t ** 2 for t in truth))

        def err(a):
            e = np.sqrt(sum((x.astype(np.float64) - t) ** 2 for x, t in zip(a, truth))) / tn
            return float(np.sqrt(np.mean(e ** 2))), float(e.max())

        bb = orc.bbox(st["pos_x"], st["pos_y"], st["pos_z"])
        keys, idx = orc.morton_keys(st["pos_x"], st["pos_y"], st["pos_z"], bb)
        _, sp = orc.sort_pairs(keys, idx)
        nodes, dropped = orc.octree_build(st["pos_x"], st["pos_y"], st["pos_z"], st["mass"], sp, bb, as_written=False)
        for theta in (0.3, 0.5, 0.8):
            h = nb.ParticleDataManager.fromState(st)
            d = nb.ParticleData()
            nb.ParticleDataManager.allocateDevice(d, n)
            nb.ParticleDataManager.copyToDevice(d, h)
            calc = nb.BarnesHutCalculator(theta)
            calc.setSofteningParameter(0.01)
            calc.computeForces(d)
            o = d.numpy()
            ours = err((o["acc_x"], o["acc_y"], o["acc_z"]))
            octf = orc.octree_forces(st["pos_x"], st["pos_y"], st["pos_z"], nodes, theta, 1.0, 0.01)
            oct_ = err(octf)
            print(json.dumps({"distribution": name, "bodies": n, "theta": theta, "ours_rms": ours[0], "ours_max": ours[1],
                              "octree_oracle_rms": oct_[0], "octree_oracle_max": oct_[1],
                              "ours_interactions_per_body": calc.getTree().interactions() / n}), flush=True)


if __name__ == "__main__":
    main()
