"""Generates tests/golden/golden_shape.npz + golden_shape.json from the UNMODIFIED
CurvedSolid/binary_shape.cpp (oracle/_ref/libref_shape.so; see oracle/refwrap/shape_ref.cpp).

Run in the build container only (needs /root/reference and `make -C oracle ref`; about 10 minutes):

    python tests/golden/make_golden_shape.py

Same record format as make_golden.py.  Two families:
  shape_t<N>      the program as shipped (301x301, radius 50, three droplets of radius 10, wall_gradient
                  -0.05), after N loop iterations; shape_t20001 is its complete run (binary_shape.cpp:22,664)
  shape_pert_t<N> the same program with its global constants and fields overwritten before init():
                  PERT below, fields from oracle_lib.shape_perturbed_fields
"""
import json
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import make_golden as mg  # noqa: E402  (record(), arrays, meta)
import oracle_lib as ol  # noqa: E402

PERT = [0.03, 0.05, 0.9, 0.1, 0.8, 1.4, 1.0 / 1.3]   # aconst, kconst, gammaconst, wall_gradient, tau_gas, tau_liq, omega_phi


def main():
    t0 = time.time()
    mask = None
    for steps in (1, 2, 201, 2001, 20001):
        r = ol.ref_shape_run(steps)
        tag = "shape_t%d" % steps
        mask = r["geometry"]
        for k in ("phase", "rho", "ux", "uy"):
            mg.record(tag, k, r[k], whole=(steps == 2001 and k == "phase"), stride=17)
        mg.record(tag, "f", r["f"], stride=211)
        mg.record(tag, "g", r["g"], stride=211)
        # what calculate_mass() prints (binary_shape.cpp:639-646): sum of phi over fluid nodes in index order
        mass = 0.0
        for v in r["phase"][mask == 1]:
            mass += v
        mg.meta[tag] = dict(nx=301, ny=301, radius=50, radius_droplet=10, steps=steps, mass=float(mass))
        print(tag, "%.0fs" % (time.time() - t0), flush=True)
    mg.record("shape_geometry", "mask", mask.astype(np.float64), whole=True)
    m, rho, phi, _, _ = ol.shape_reference_setup()
    rho2, phi2, ux2, uy2 = ol.shape_perturbed_fields(m, rho, phi)
    for steps in (1, 300):
        r = ol.ref_shape_run(steps, PERT, 0, phi2, rho2, ux2, uy2)
        tag = "shape_pert_t%d" % steps
        for k in ("phase", "rho", "ux", "uy"):
            mg.record(tag, k, r[k], stride=17)
        mg.record(tag, "f", r["f"], stride=211)
        mg.record(tag, "g", r["g"], stride=211)
        mg.meta[tag] = dict(nx=301, ny=301, steps=steps, params=PERT + [0.0, 0.0])
        print(tag, "%.0fs" % (time.time() - t0), flush=True)
    np.savez_compressed(os.path.join(HERE, "golden_shape.npz"), **mg.arrays)
    json.dump(mg.meta, open(os.path.join(HERE, "golden_shape.json"), "w"), indent=1, sort_keys=True)
    print("wrote golden_shape.npz (%d arrays), %.0fs" % (len(mg.arrays), time.time() - t0))


if __name__ == "__main__":
    main()
