"""TEST INFRASTRUCTURE ONLY.

CPU oracle for the paged-decode hot path (plain-C restatement in
``paged_attn_oracle.c``) and a loader for the reference's own CPU kernels built
into ``oracle/_ref`` (``ref_cpu.py``).

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline``
/ ``--impl reference`` legs may import this package.  The product package
``vllm_main_modify_b200`` never does.
"""
