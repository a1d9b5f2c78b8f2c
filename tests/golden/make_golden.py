"""Generates tests/golden/golden.npz + golden.json from the UNMODIFIED reference.

Run in the build container only (needs /root/reference and `make -C oracle ref`):

    python tests/golden/make_golden.py

Every entry is produced by the reference's own code compiled into oracle/_ref
(see oracle/Makefile, oracle/refwrap/).  For each snapshot the file keeps, per
field: the SHA-256 of the raw little-endian float64 bytes, the sum, min, max and
a strided sample of exact values (stride chosen so the fixture stays small);
small arrays are kept whole.  It also checks the text-output fingerprints that
BASELINE.md section 3 recorded from the stock programs, as a guard that the
wrappers did not change what the programs compute.
"""
import hashlib
import json
import os
import shutil
import sys
import tempfile
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import oracle_lib as ol  # noqa: E402

arrays = {}
meta = {}


def record(name, field, a, stride=None, whole=False):
    a = np.ascontiguousarray(a, dtype=np.float64).ravel()
    ent = dict(sha256=hashlib.sha256(a.tobytes()).hexdigest(), n=int(a.size), sum=float(np.sum(a)),
               min=float(np.min(a)), max=float(np.max(a)))
    key = "%s/%s" % (name, field)
    if whole:
        arrays[key] = a
        ent["whole"] = True
    else:
        ent["stride"] = int(stride)
        arrays[key] = a[::stride].copy()
    meta[key] = ent


def sha_prefix(path):
    return hashlib.sha256(open(path, "rb").read()).hexdigest()[:16]


def main():
    t00 = time.time()
    # ---- MRT (whole arrays; the programs run their full length) ----
    for variant, a1, a2, tag in ((0, 10, 10, "mrt_gs_10_10"), (0, 12, 8, "mrt_gs_12_8"), (1, 0, 0, "mrt_hermite")):
        d = tempfile.mkdtemp()
        os.mkdir(os.path.join(d, "tmp"))
        r = ol.ref_mrt_run(variant, a1, a2, outdir=d)
        for k in ("rho", "ux", "uy", "f"):
            record(tag, k, r[k], whole=True)
        meta[tag] = dict(nx=r["nx"], ny=r["ny"], steps=r["nsteps"] + 1, args=[a1, a2])
        if tag == "mrt_gs_10_10":
            assert sha_prefix(os.path.join(d, "tmp", "phase000800.dat")) == "07ca3fdcfdd7cb4c"
            assert sha_prefix(os.path.join(d, "tmp", "proba000800.dat")) == "2070cdc933cf2c1d"
        if tag == "mrt_hermite":
            assert sha_prefix(os.path.join(d, "tmp", "phase000500.dat")) == "efa99204039552d8"
            assert sha_prefix(os.path.join(d, "tmp", "proba000500.dat")) == "45c1dd18a5ef069e"
        shutil.rmtree(d)
    print("mrt done %.0fs" % (time.time() - t00), flush=True)

    # ---- FreeEnergy/binary.cpp ----
    for last in (0, 100, 2000, 20000):
        d = tempfile.mkdtemp()
        r = ol.ref_binary_run(last, noutput=100 if last == 20000 else 0, outdir=d)
        tag = "binary2d_t%d" % last
        whole = last in (0, 100)
        for k in ("phase", "rho", "ux", "uy"):
            record(tag, k, r[k], stride=7, whole=whole)
        for k in ("f", "g"):
            record(tag, k, r[k], stride=61)
        meta[tag] = dict(nx=r["nx"], ny=r["ny"], steps=last + 1)
        if last == 20000:
            assert sha_prefix(os.path.join(d, "phase020000.dat")) == "97765059e339cfbc"
            assert sha_prefix(os.path.join(d, "velocityx020000.dat")) == "adfd2c25b9864189"
            assert sha_prefix(os.path.join(d, "phase000100.dat")) == "836c201b829113f9"
        shutil.rmtree(d)
        print(tag, "done %.0fs" % (time.time() - t00), flush=True)

    # ---- Microchannel3D QuarterChannel (27x27x52) and a FullChannel-parameter case ----
    ref = ol.Ref3D(27, 27, 52, ol.PARAMS_QUARTER, "quarter")
    ref.set_bubble(6)
    ref.init()
    t = 0
    for upto in (1, 101, 1001):
        ref.step(upto - t)
        t = upto
        tag = "quarter_t%d" % (upto - 1)
        fl = ref.fields()
        for k, a in fl.items():
            record(tag, k, a, stride=37)
        f, g = ref.pops()
        nw = ol.nonwall_mask_3d(27, 27, 52)
        record(tag, "f_nonwall", f[nw], stride=997)
        record(tag, "g_nonwall", g[nw], stride=997)
        meta[tag] = dict(nx=27, ny=27, nz=52, steps=upto, width=6, params=ol.PARAMS_QUARTER)
        if upto == 1001:
            assert abs(np.sum(fl["phase"]) - 21447.109013716385) < 1e-6
            assert fl["phase"].reshape(52, 27, 27)[26, 5, 5] == -1.0126351056294705
            assert fl["uz"].reshape(52, 27, 27)[26, 1, 1] == 0.0047798239850062102
        print(tag, "done %.0fs" % (time.time() - t00), flush=True)
    ref.close()

    ref = ol.Ref3D(92, 52, 24, ol.PARAMS_FULL, "full")
    ref.set_bubble(8)
    ref.init()
    ref.step(21)
    tag = "fullslice_t20"
    for k, a in ref.fields().items():
        record(tag, k, a, stride=101)
    meta[tag] = dict(nx=92, ny=52, nz=24, steps=21, width=8, params=ol.PARAMS_FULL)
    ref.close()
    print(tag, "done %.0fs" % (time.time() - t00), flush=True)

    # the real FullChannel grid, first 4 steps (5.6 s/step on one core)
    ref = ol.Ref3D(92, 52, 1500, ol.PARAMS_FULL, "full")
    ref.set_bubble(8)
    ref.init()
    ref.step(4)
    tag = "full_t3"
    for k, a in ref.fields().items():
        record(tag, k, a, stride=9973)
    meta[tag] = dict(nx=92, ny=52, nz=1500, steps=4, width=8, params=ol.PARAMS_FULL)
    ref.close()
    print(tag, "done %.0fs" % (time.time() - t00), flush=True)

    # ---- ShanChen: with the shared exp (primary) and with glibc exp (secondary) ----
    for shared, tagp in ((True, "shanchen_sharedexp"), (False, "shanchen_libm")):
        for last in ((0, 1000, 30000) if shared else (1000, 30000)):
            d = tempfile.mkdtemp()
            r = ol.ref_shanchen_run(last, shared_exp=shared, outdir=d)
            tag = "%s_t%d" % (tagp, last)
            for k in ("rho", "u1", "u2"):
                record(tag, k, r[k], stride=5, whole=(last == 1000 and shared))
            record(tag, "f", r["f"], stride=53)
            meta[tag] = dict(nx=r["nx"], ny=r["ny"], steps=last + 1)
            if not shared and last == 30000:
                assert sha_prefix(os.path.join(d, "density30000.dat")) == "156dac3b3e219744"
                assert sha_prefix(os.path.join(d, "velx30000.dat")) == "a682be0c428b603f"
                assert sha_prefix(os.path.join(d, "density01000.dat")) == "277dd3baabde8f27"
            shutil.rmtree(d)
            print(tag, "done %.0fs" % (time.time() - t00), flush=True)

    np.savez_compressed(os.path.join(HERE, "golden.npz"), **arrays)
    json.dump(meta, open(os.path.join(HERE, "golden.json"), "w"), indent=1, sort_keys=True)
    print("wrote golden.npz (%d arrays), %.0fs" % (len(arrays), time.time() - t00))


if __name__ == "__main__":
    main()
