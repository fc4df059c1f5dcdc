"""CPU oracle for the thermoextrap hot path.  TEST INFRASTRUCTURE ONLY.

Nothing under ``oracle/`` is imported by the product package
(``thermoextrap_b200``).  Only ``tests/``, ``__graft_entry__.smoke()`` and
``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs may use it, and
there only as the checker / reported CPU baseline.

Parity status
-------------
The arithmetic of the reference's hot path lives in the third-party package
``cmomy`` (pinned 0.24.0, ``/root/reference/uv.lock:454-455``), which is NOT
vendored under ``/root/reference`` and is not installable here (no network).
The oracle therefore restates cmomy's *published* algorithm (one-pass
weighted central-comoment push, Chan/Pebay block merge) and is pinned against
what the reference's own tests pin for this path:

* raw moments == ``np.average(x * u**i)``  (``tests/test_data.py:7-38``,
  ``src/thermoextrap/legacy/utilities.py:12-39``)
* freq-weighted resample == gather-by-indices then reduce
  (``tests/test_data.py:94-112``)
* derivatives == the legacy quotient-rule implementation
  (``src/thermoextrap/legacy/utilities.py:61-111``; golden values generated
  by running that reference file itself, ``tests/golden/make_golden.py``)
* symbolic expressions == the reference's own ``beta.py`` / ``lnpi.py`` /
  ``models.py`` sympy classes imported in the build container
  (``tests/golden/make_golden.py`` -> ``tests/golden/exprs.json``)
* lnPi golden vectors (``tests/lnpi_data/sample_data.json`` -> subset
  committed as ``tests/golden/lnpi_sample.json``)
* the numpy bounded-integer stream (live numpy is the oracle).
"""
