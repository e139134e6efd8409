"""Domain decomposition, QAP placement and the process group (all CPU-side)."""
from .partition import rank_partition, node_partition, RankPartition, NodePartition, set_partition_ties  # noqa: F401
from .placement import qap_solve, place_node_aware  # noqa: F401
from .comm import init_from_env, rank, size, barrier, finalize  # noqa: F401
