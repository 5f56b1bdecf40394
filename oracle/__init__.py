"""oracle/ — TEST INFRASTRUCTURE, not product code.

CPU (PyTorch fp32) restatement of the reference hot path plus the tooling that pins it against the
reference itself.  Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import anything from here, and only as the checker / the timed CPU
baseline.  ``pytorch_segmentation_b200`` never imports it and has no CPU fallback.

Parity status: the reference ships no tests and no golden vectors, so parity is
*unpinned by the reference's own fixtures*.  It is pinned here instead by running the reference's
unmodified ``models/*.py`` / ``utils/utils.py`` / ``test.py:test`` (imported from /root/reference by
``oracle/reference_loader.py`` in the build container) and committing the resulting vectors under
``tests/golden/`` (generator: ``oracle/make_golden.py``).  The arithmetic leaf ``ConvNormAct`` and
the backbones live in the absent third-party package ``pytorch-modules>=0.3.0``
(requirements.txt:4); ``oracle/shim/pytorch_modules`` restates them from the reference's call
sites.
"""
