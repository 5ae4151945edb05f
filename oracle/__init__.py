"""TEST INFRASTRUCTURE ONLY.

``oracle/`` holds (i) ``choreo_oracle.py``, a NumPy restatement of the reference's action hot
path, and (ii) ``_ref/`` (git-ignored), the unmodified reference built by ``build_ref.sh``.
Only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference`` legs of
``bench.py`` may import anything from here; the product package ``choreographies2_b200`` never does.
"""
