"""CPU oracle for the WOFT per-frame hot path.  TEST INFRASTRUCTURE ONLY.

Nothing in ``woft_b200/`` may import this package.  It is imported by
``tests/``, by ``__graft_entry__.smoke()`` and by the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py``, and there only as the checker (or as
the thing timed on the host cores), never as part of the product path.

Contents
--------
``woft_oracle``     torch-CPU restatement of the reference's correlation volume,
                    pyramid, radius-r lookup and weighted homography estimators.
                    Pinned against the reference's own code: ``gen_golden.py``
                    imports ``/root/reference`` (build container only), runs it on
                    seeded inputs and writes ``tests/golden/*.npz``; the CPU tests
                    check this restatement against those vectors.
``kornia_restated`` restatement of the three ``kornia==0.5.11`` functions the
                    reference calls.  kornia is an un-vendored dependency that is
                    not installable here: this part is **parity unpinned** (it is
                    written from the published 0.5.x algorithm and could not be
                    diffed against the real package).
``ref_import``      stub modules + sys.path set-up to import the unmodified
                    reference from ``/root/reference`` (build container only; the
                    GPU box has no ``/root/reference``).
``gen_golden``      the script that generated ``tests/golden``.
"""
