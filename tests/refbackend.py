"""The reference-driven learners backend lives with the other checkers in oracle/
(bench.py's CPU legs use it too); re-exported here for the tests."""
from oracle.refbackend import RefBackend, _Counter  # noqa: F401
