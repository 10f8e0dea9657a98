#!/usr/bin/env python
"""Generate the golden fixtures from the REFERENCE ITSELF (oracle/_ref/libnbody_ref.so = the
unmodified reference sources compiled by oracle/build_ref.sh).

  python tests/golden/make_golden.py host   # CPU container: the reference's host-callable helpers
  python tests/golden/make_golden.py gpu    # B200 box: the reference's CUDA kernels (run via gpurun,
                                            # writes gpurun_out/ref_gpu_*.npz; copy into tests/golden/)
Inputs are seeded; the fixtures hold inputs' seeds and the reference's outputs only.
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ref_bindings as ref  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def host():
    rng = np.random.default_rng(20251104)
    pts = rng.random((256, 3)).astype(np.float32)
    pts[0] = [0, 0, 0]
    pts[1] = [1, 1, 1]              # x*1024 = 1024 -> clamped to 1023
    pts[2] = [0.999999, 0.5, 0.25]
    morton = [ref.morton_code(float(x), float(y), float(z)) for x, y, z in pts]
    cpts = (rng.random((256, 3)) * 40 - 5).astype(np.float32)
    cells = [0.5, 1.0, 2.0, 0.37]
    cidx = [[ref.cell_index(float(x), float(y), float(z), c) for x, y, z in cpts] for c in cells]
    pairs = []
    for _ in range(64):
        p1, p2 = rng.normal(size=3).astype(np.float32), rng.normal(size=3).astype(np.float32)
        m2, G, eps = float(np.float32(rng.random() * 3)), float(np.float32(0.5 + rng.random())), float(np.float32(rng.random() * 0.1))
        f = ref.pair_force_cpu(p1, p2, 1.0, m2, G, eps)
        pairs.append(dict(p1=p1.tolist(), p2=p2.tolist(), m2=m2, G=G, eps=eps, f=[float(v) for v in f]))
    out = dict(source="oracle/_ref host-callable reference functions: computeMortonCode (force_barnes_hut.cu:31-38), "
                      "SpatialHashGrid::getCellIndex (force_spatial_hash.cu:14-17), computeGravitationalForceCPU "
                      "(force_direct.cu:109-117)",
               morton_points=pts.tolist(), morton=morton, cell_points=cpts.tolist(), cell_sizes=cells, cell_index=cidx,
               pair_forces=pairs)
    with open(os.path.join(HERE, "ref_host_helpers.json"), "w") as f:
        json.dump(out, f)
    print("wrote ref_host_helpers.json")


def gpu():
    from nbody_b200 import ics
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    n = 4096
    # direct + Verlet + energy on the Plummer config-1 workload
    st = ics.plummer(n, seed=42)
    s0 = {k: v.copy() for k, v in st.items()}
    ref.forces(s0, 0, 1.0, 0.01)
    s1 = {k: v.copy() for k, v in st.items()}
    _, en = ref.run(s1, 0, 1e-3, 1.0, 0.01, steps=10, initial_forces=True, want_energy=True)
    # Barnes-Hut front half (as shipped) on the reference's SPHERICAL shape
    sp = ics.uniform_sphere(n, 10.0, seed=42)
    bh = ref.bh_tree({k: v.copy() for k, v in sp.items()}, max_nodes=64)
    sbh = {k: v.copy() for k, v in sp.items()}
    ref.forces(sbh, 1, 1.0, 0.1, theta=0.5)
    # spatial hash on a uniform box
    bx = ics.uniform_box(n, -10, 10, seed=42)
    gr = ref.sh_grid({k: v.copy() for k, v in bx.items()}, 1.0)
    ssh = {k: v.copy() for k, v in bx.items()}
    ref.forces(ssh, 2, 1.0, 0.1, cell=1.0, cutoff=2.0)
    np.savez_compressed(
        os.path.join(ROOT, "gpurun_out", "ref_gpu_n4096.npz"),
        note=np.array("reference CUDA kernels (unmodified, sm_100) on B200; inputs: ics.plummer/uniform_sphere/uniform_box(4096, seed=42)"),
        direct_acc=np.stack([s0["acc_x"], s0["acc_y"], s0["acc_z"]]),
        verlet10=np.stack([s1[k] for k in ("pos_x", "pos_y", "pos_z", "vel_x", "vel_y", "vel_z", "acc_x", "acc_y", "acc_z",
                                           "acc_old_x", "acc_old_y", "acc_old_z")]),
        energies=np.array(en, np.float32),
        bh_bbox=bh["bbox"], bh_keys_unsorted=bh["keys_unsorted"], bh_keys=bh["keys"], bh_perm=bh["perm"],
        bh_node_count=np.array(bh["node_count"]), bh_nodes=bh["nodes"].view(np.uint8),
        bh_acc=np.stack([sbh["acc_x"], sbh["acc_y"], sbh["acc_z"]]),
        sh_bbox=gr["bbox"], sh_dims=gr["dims"], sh_particle_cell=gr["particle_cell"], sh_cell_start=gr["cell_start"],
        sh_cell_end=gr["cell_end"], sh_sorted_indices=gr["sorted_indices"],
        sh_acc=np.stack([ssh["acc_x"], ssh["acc_y"], ssh["acc_z"]]))
    print("wrote gpurun_out/ref_gpu_n4096.npz; bh nodes:", bh["node_count"])


if __name__ == "__main__":
    {"host": host, "gpu": gpu}[sys.argv[1]]()
