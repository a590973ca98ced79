"""CPU oracle for the media hot path -- TEST INFRASTRUCTURE, not product code.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this package.  The product path
(``aiko_services_b200``) never does, and fails loudly without its CUDA library.

``oracle.image``  ctypes wrapper over ``liboracle.so`` (``image_oracle.c``)
``oracle.audio``  numpy float64 restatement of the reference's (dead-code) audio chain
"""
