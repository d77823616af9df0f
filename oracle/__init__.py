"""oracle/ -- TEST INFRASTRUCTURE ONLY.

CPU restatement of the reference's CFM.sample + Vocos path (``f5_oracle.py``),
shims for the third-party modules the reference imports but this image lacks
(``shims.py``) and a by-path loader for the unmodified reference files
(``ref_loader.py``, container only -- /root/reference does not exist on the GPU box).

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline /
``--impl reference`` legs may import anything from here, and only as the checker
or as the timed CPU baseline -- never as the product path.
"""
