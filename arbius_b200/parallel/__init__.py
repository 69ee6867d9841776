"""Multi-GPU runtime: process groups, the peer-memory heap and the fused compute+collective ops."""
from .groups import init_distributed, local_rank, subset_group
