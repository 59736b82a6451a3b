"""oracle/ -- TEST INFRASTRUCTURE ONLY.

CPU restatement of the reference's per-frame localisation hot path:
the reference's own numpy code restated in ``ref_port.py`` and the Open3D 0.17.0 calls it
delegates to restated in ``o3d_core.c`` / ``o3d_reg.c`` / ``o3d_core.py`` (PARITY UNPINNED for the
Open3D part: Open3D is a pip dependency that is not vendored under /root/reference and is not
installable offline; nothing in the reference pins its outputs).

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl reference``
legs may import this package.  The product package never does.
"""
