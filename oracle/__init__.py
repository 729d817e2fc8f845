"""CPU oracle for the Think-and-Tell hot path.  TEST INFRASTRUCTURE ONLY.

PARITY UNPINNED: the reference holds no tests / golden vectors for this path and
TensorFlow is not installable in this image, so this restatement of the
reference's Keras graph cannot be checked against the reference running here.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline /
``--impl reference`` legs may import this package -- as the checker or as the timed CPU
baseline, never as part of the product path.
"""
