"""Decomposition (``decomp``), multi-GPU fabric over CUDA IPC (``fabric``) and the library-collective solvers
(``collective``: NCCL baseline on GPUs, gloo-distributed CPU solver)."""
