"""CPU oracle for the CORSAIR correspondence-and-registration hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``corsair_b200/`` imports this package; only
``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference`` legs
of ``bench.py`` may use it, and there only as the checker or as the timed CPU baseline.

* ``oracle.port``      – NumPy/scipy/sklearn restatement of the reference's Python functions
                         (each cites the reference file:line it follows).  It is validated
                         against the reference's own code imported from ``/root/reference``
                         (``oracle.ref_import``) by ``tests/test_oracle_vs_reference.py`` and
                         against committed fixtures in ``tests/golden/``.
* ``oracle.ransac_np`` – NumPy float64 restatement of Open3D's
                         ``registration_ransac_based_on_correspondence`` driven by explicit
                         minimal sets.  PARITY UNPINNED for this one call: Open3D (pinned by
                         ``docker/Dockerfile:77-80`` to commit 160209d0) is not in the
                         reference tree and not installable here; the reference has no tests
                         or golden vectors for it.
* ``oracle/ransac_c.c`` – the same restatement in C with OpenMP (timed CPU baseline and a
                         second, independent checker for the NumPy one).
"""
