"""``model.RAFT.raft`` (reference model/model.py:7 imports RAFT from here)."""
from zero_tig_b200.model.RAFT.raft import RAFT  # noqa: F401
