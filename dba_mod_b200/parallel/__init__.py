"""Multi-GPU plumbing: one process per GPU, symmetric memory, fused peer-memory collectives, NCCL baseline."""
