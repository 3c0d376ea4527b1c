"""CPU oracle for the candidate-generation hot path.  TEST INFRASTRUCTURE ONLY.

Nothing under ``orion_b200/`` may import this package.  It is used by
``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py`` as the *checker*, never as the
thing that is shipped or measured as the product.
"""
