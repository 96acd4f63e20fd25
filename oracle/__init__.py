"""CPU oracle for the LiDAR-camera fusion / evaluation path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``lidar_object_detection_b200/`` may
import this package.  The only legitimate callers are ``tests/``,
``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference``
legs of ``bench.py`` -- and there only as the checker or the timed CPU
baseline, never as the product path.

Layout
------
``kitti360_restated.py``  restatement of the *missing* third-party dependency
                          ``kitti360Scripts`` (autonomousvision/kitti360Scripts,
                          git submodule pinned at commit
                          40ee015de8277b2d907d32a30cdbd97072c7d7c3, absent from
                          /root/reference).  PARITY UNPINNED for these few
                          functions: the pinned source could not be diffed here.
``pipeline.py``           numpy / OpenCV restatement of the reference's own
                          functions on the hot path (each cites file:line).
                          PINNED: ``gen_golden.py`` ran the reference's
                          unmodified functions (imported from /root/reference
                          through stub modules, see ``reference_loader.py``)
                          on the KITTI360_sample frames and committed the
                          outputs under ``tests/golden/``; ``tests/`` check this
                          restatement against those vectors.
``mask_provider.py``      deterministic stand-in instance masks (YOLO is out of
                          scope: masks are inputs).
``oracle_c.c``            plain-C restatement of the floating-point-order
                          sensitive pieces (projection, box containment) with
                          explicit fma(), used to check depth bit-for-bit.
``reference_loader.py``   imports the reference scripts unmodified (only works
                          where /root/reference exists, i.e. in the build
                          container; never on the GPU box).
``gen_golden.py``         regenerates ``tests/golden/``.
"""
