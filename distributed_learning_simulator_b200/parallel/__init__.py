"""Multi-GPU execution: one process per GPU under ``torch.distributed`` (NCCL for the
plumbing), peer-mapped parameter arenas, and the hand-written P2P reduce+broadcast."""
