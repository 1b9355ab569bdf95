"""TEST INFRASTRUCTURE ONLY.

CPU oracle for the SparseStorage hot path of PFCCLab/paddle_sparse.

* ``oracle.native``  — ctypes doorway onto (a) ``oracle/_ref/libps_ref_cpu.so``, the
  reference's own ``csrc/cpu/*.cpp`` compiled unmodified, and (b)
  ``oracle/_build/libps_oracle_port.so``, this repo's plain-C restatement.
* ``oracle.np_ref``  — NumPy restatement of the reference's Python op chains
  (``paddle_sparse/storage.py`` and friends) and of the ``paddle_scatter`` semantics
  they rely on.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU legs
(``cpu_baseline`` / ``--impl reference``) may import this package.  The product
package ``paddle_sparse_b200`` never does and has no CPU fallback.
"""
