"""CPU restatement of ``PerturbModel.predict`` -- TEST INFRASTRUCTURE ONLY (never imported by the product).

Follows /root/reference/src/thermoextrap/models.py:1009-1046 (xarray form) and
legacy/reweight.py:89-101 (numpy form) line by line.  Pinned by: tests/golden/perturb.json, the output of
the reference's own ``legacy.PerturbModel.predict`` run from /root/reference on its FixtureData
(tests/golden/make_golden.py; the reference's test tests/test_beta.py:154-162 compares exactly these two
forms); ``dalpha = 0`` gives the plain sample mean; a Gaussian ``u`` gives the analytic reweighted mean.
"""

from __future__ import annotations

import numpy as np


def perturb_predict(x, u, dalpha):
    """``out[alpha, *val]``; ``x[rec, *val]``, ``u[rec]``, ``dalpha = alpha - alpha0``."""
    x = np.asarray(x, dtype=np.float64)
    u = np.asarray(u, dtype=np.float64)
    dalpha = np.atleast_1d(np.asarray(dalpha, dtype=np.float64))
    x2 = x.reshape(x.shape[0], -1)
    dbu = (-1.0) * np.tensordot(dalpha, u, axes=0)           # reweight.py:93-94
    diff = dbu - dbu.max(axis=1, keepdims=True)                 # :95
    expv = np.exp(diff)                                         # :96
    numer = np.dot(expv, x2) / float(x2.shape[0])               # :99
    denom = np.average(expv, axis=1)                            # :100
    return (numer / denom[:, None]).reshape(len(dalpha), *x.shape[1:])  # :101
