"""TEST INFRASTRUCTURE ONLY.

CPU oracle for the decode hot path: a plain-C restatement of the reference's algorithms
(oracle.c -> liboracle.so, wrapped by oracle.cpu_oracle) and, where it can run, the reference's own
CPU kernels compiled from source (oracle/_ref, wrapped by oracle.ref_ops).

Nothing under iaas_sglang_b200/ imports this package.  Only tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs do, and only as the checker or as the CPU number
reported beside the GPU number.
"""
