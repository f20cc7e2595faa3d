"""Helpers of bench.py (measurement only, not product code): the workload definition and its algorithmic-byte
formulas, the kernel-level profiler, the CPU reference arm (oracle port), the torch-GPU restatement of the reference
path, the rollout and the spatially sharded measurements."""
