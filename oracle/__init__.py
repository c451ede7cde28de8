"""ORACLE — TEST INFRASTRUCTURE ONLY.  Parity status: PINNED against the reference's own CUDA kernels (oracle/_ref; see oracle/orc_types.h).
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this."""
