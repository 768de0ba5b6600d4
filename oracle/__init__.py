"""CPU fp32 oracle of the partitioned transformer block forward.

TEST INFRASTRUCTURE ONLY.  It stands in for a reference implementation that does not exist
(TensorHusker/Butterfly ships no source: ref:CLAUDE.md:13-15).  PARITY UNPINNED by the reference;
pinned against HuggingFace transformers instead (tests/test_oracle_hf.py) and frozen by the
golden fixtures in tests/golden/.  Importable only from tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py.
"""
from .model import OracleModel  # noqa: F401
from . import ops  # noqa: F401
