"""Generates tests/golden/golden_trt.npz + golden_trt.json from the UNMODIFIED
DiffusionCylinder/cylinder.cpp (oracle/_ref/libref_cylinder.so; see oracle/refwrap/cylinder_ref.cpp).

Run in the build container only (needs /root/reference and `make -C oracle ref`):

    python tests/golden/make_golden_trt.py

Same record format as make_golden.py.  Two families:
  trt_t<N>      the program as shipped (129x129, radius 40, zero velocity), after N loop iterations;
                trt_t64001 is its complete run (N+1 iterations, cylinder.cpp:21,353)
  trt_adv_t<N>  the same program with its global constants and fields overwritten before init():
                omega 1.3, ce 0.25, conc_wall 0.7, hashed rho and velocity (oracle_lib.trt_advected_fields)
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

ADV = dict(omega=1.3, ce=0.25, conc_wall=0.7)


def main():
    t0 = time.time()
    mask = None
    for steps in (1, 2, 201, 4001, 64001):
        r = ol.ref_cylinder_run(steps)
        tag = "trt_t%d" % steps
        mask = r["geometry"]
        mg.record(tag, "rho", r["rho"], whole=(steps == 4001), stride=7)
        mg.record(tag, "f", r["f"], stride=31)
        mg.meta[tag] = dict(nx=129, ny=129, radius=40, steps=steps)
        print(tag, "%.0fs" % (time.time() - t0), flush=True)
    mg.record("trt_geometry", "mask", mask.astype(np.float64), whole=True)
    rho, ux, uy = ol.trt_advected_fields(mask, ADV["conc_wall"])
    for steps in (1, 500):
        r = ol.ref_cylinder_run(steps, ADV["omega"], ADV["ce"], ADV["conc_wall"], rho, ux, uy)
        tag = "trt_adv_t%d" % steps
        mg.record(tag, "rho", r["rho"], stride=7)
        mg.record(tag, "f", r["f"], stride=31)
        mg.meta[tag] = dict(nx=129, ny=129, radius=40, steps=steps, **ADV)
        print(tag, "%.0fs" % (time.time() - t0), flush=True)
    np.savez_compressed(os.path.join(HERE, "golden_trt.npz"), **mg.arrays)
    json.dump(mg.meta, open(os.path.join(HERE, "golden_trt.json"), "w"), indent=1, sort_keys=True)
    print("wrote golden_trt.npz (%d arrays), %.0fs" % (len(mg.arrays), time.time() - t0))


if __name__ == "__main__":
    main()
