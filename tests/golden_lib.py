"""Access to the committed golden fixtures (tests/golden/golden.npz + golden.json),
generated from the unmodified reference by tests/golden/make_golden.py."""
import hashlib
import json
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
_npz = None
_meta = None


def _load():
    global _npz, _meta
    if _npz is None:
        # golden.npz/json (make_golden.py) plus the additive sets golden_<name>.npz/json (make_golden_<name>.py)
        import glob
        _npz, _meta = {}, {}
        for jf in sorted(glob.glob(os.path.join(HERE, "golden", "golden*.json"))):
            _meta.update(json.load(open(jf)))
            with np.load(jf[:-5] + ".npz") as z:
                for k in z.files:
                    _npz[k] = z[k]
    return _npz, _meta


def meta(tag):
    return _load()[1][tag]


def check(tag, field, a):
    """Returns the number of sampled entries that differ and whether the SHA-256 of the whole array matches."""
    npz, m = _load()
    key = "%s/%s" % (tag, field)
    ent = m[key]
    a = np.ascontiguousarray(a, dtype=np.float64).ravel()
    assert a.size == ent["n"], (key, a.size, ent["n"])
    ref = npz[key]
    got = a if ent.get("whole") else a[::ent["stride"]]
    ndiff = int(np.count_nonzero(~((got == ref) | (np.isnan(got) & np.isnan(ref)))))
    sha_ok = hashlib.sha256(a.tobytes()).hexdigest() == ent["sha256"]
    return ndiff, sha_ok


def samples(tag, field):
    npz, m = _load()
    key = "%s/%s" % (tag, field)
    return npz[key], m[key]


def rel_inf(tag, field, a, scale=None):
    """||a - golden||_inf / ||golden||_inf on the sampled entries (BASELINE.md section 5 metric)."""
    ref, ent = samples(tag, field)
    a = np.ascontiguousarray(a, dtype=np.float64).ravel()
    got = a if ent.get("whole") else a[::ent["stride"]]
    den = scale if scale is not None else max(abs(ent["min"]), abs(ent["max"]))
    return float(np.max(np.abs(got - ref)) / den) if den > 0 else float(np.max(np.abs(got - ref)))
