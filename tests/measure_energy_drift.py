#!/usr/bin/env python
"""SURVEY 8(d) config 2, energy-drift check: relative drift of the total energy over 1000 Velocity-Verlet steps
(dt 1e-3, eps 0.01, Plummer seed 42) at 65 536 and 262 144 bodies, ours and -- same initial conditions -- the
reference's own kernels (oracle/_ref/libnbody_ref.so), E in fp64 by our diagnostic for both.  A measurement
script, not a test (the 8192-body version of the same check is tests/test_gpu_vs_reference.py): run on a B200,
one JSON line per case.  usage: python tests/measure_energy_drift.py [--steps K] [--every S] [--seeds a,b,c] [n ...]
--seeds runs every size once per seed (an ENSEMBLE: the end-point drift of one run is one sample of a rounding walk)
and appends one summary line per size with the median and maximum over the seeds for both arms.
--every S also samples the energy every S steps on both arms and records the largest excursion along the trajectory
(the end-point value alone is one sample of a rounding-noise random walk)."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))  # run from anywhere


def main():
    import torch
    import nbody_b200 as nb
    from nbody_b200 import ics
    from oracle import ref_bindings as ref

    argv = sys.argv[1:]
    steps = 1000
    if "--steps" in argv:  # fewer steps for the sizes where 1000 do not fit the GPU budget (1 Mi: 0.43 s per step)
        k = argv.index("--steps")
        steps = int(argv[k + 1])
        del argv[k:k + 2]
    every = 0
    if "--every" in argv:
        k = argv.index("--every")
        every = int(argv[k + 1])
        del argv[k:k + 2]
    seeds = [42]
    if "--seeds" in argv:
        k = argv.index("--seeds")
        seeds = [int(x) for x in argv[k + 1].split(",")]
        del argv[k:k + 2]
    sizes = [int(a) for a in argv] or [65536, 262144]
    dt, eps = 1e-3, 0.01
    records = []
    for n, seed in [(n, sd) for n in sizes for sd in seeds]:
        st = ics.plummer(n, seed=seed)

        def device(state):
            h = nb.ParticleDataManager.fromState(state)
            d = nb.ParticleData()
            nb.ParticleDataManager.allocateDevice(d, h.count)
            nb.ParticleDataManager.copyToDevice(d, h)
            return d

        d = device(st)
        calc = nb.DirectForceCalculator()
        calc.setSofteningParameter(eps)
        calc.computeForces(d)
        integ = nb.Integrator()
        e0 = integ.computeTotalEnergy(d, 1.0, eps)
        torch.cuda.synchronize()
        t0 = time.time()
        samples_ours = []
        for k in range(steps):
            integ.integrate(d, calc, dt)
            if every and (k + 1) % every == 0 and k + 1 < steps:
                samples_ours.append(integ.computeTotalEnergy(d, 1.0, eps))
        e1 = integ.computeTotalEnergy(d, 1.0, eps)
        t_ours = time.time() - t0
        samples_ours.append(e1)
        rec = {"bodies": n, "seed": seed, "steps": steps, "dt": dt, "softening": eps, "E0": e0, "E_ours": e1,
               "drift_ours": abs(e1 - e0) / abs(e0), "seconds_ours": t_ours}
        if every:
            rec.update({"sample_every": every,
                        "max_excursion_ours": max(abs(e - e0) for e in samples_ours) / abs(e0)})
        if ref.available():
            r = {k: v.copy() for k, v in st.items()}
            t0 = time.time()
            if every:  # the same trajectory in segments: the state (accelerations included) round-trips through the host
                samples_ref, done = [], 0
                while done < steps:
                    seg = min(every, steps - done)
                    ref.run(r, 0, dt, 1.0, eps, steps=seg, initial_forces=(done == 0))
                    done += seg
                    samples_ref.append(integ.computeTotalEnergy(device(r), 1.0, eps))
                rec["max_excursion_reference"] = max(abs(e - e0) for e in samples_ref) / abs(e0)
            else:
                ref.run(r, 0, dt, 1.0, eps, steps=steps, initial_forces=True)
            rec["seconds_reference"] = time.time() - t0
            er = integ.computeTotalEnergy(device(r), 1.0, eps)
            rec.update({"E_reference": er, "drift_reference": abs(er - e0) / abs(e0)})
            # how far apart the two trajectories are after 1000 steps
            o = d.numpy()
            dx = np.sqrt((o["pos_x"] - r["pos_x"]) ** 2 + (o["pos_y"] - r["pos_y"]) ** 2 + (o["pos_z"] - r["pos_z"]) ** 2)
            rec["max_position_difference"] = float(dx.max())
        print(json.dumps(rec), flush=True)
        records.append(rec)
    if len(seeds) > 1:
        for n in sizes:
            rs = [r for r in records if r["bodies"] == n]
            summ = {"summary": True, "bodies": n, "seeds": [r["seed"] for r in rs], "steps": steps}
            for key in ("drift_ours", "drift_reference", "max_excursion_ours", "max_excursion_reference"):
                v = [r[key] for r in rs if key in r]
                if v:
                    summ["median_" + key], summ["max_" + key], summ["min_" + key] = (float(np.median(v)), float(max(v)),
                                                                                     float(min(v)))
            if all("drift_reference" in r for r in rs):
                summ["seeds_where_ours_is_lower"] = int(sum(r["drift_ours"] <= r["drift_reference"] for r in rs))
                if all("max_excursion_reference" in r for r in rs):
                    summ["seeds_where_our_excursion_is_lower"] = int(
                        sum(r["max_excursion_ours"] <= r["max_excursion_reference"] for r in rs))
            print(json.dumps(summ), flush=True)


if __name__ == "__main__":
    main()
