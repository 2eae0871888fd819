"""Alias package: ``from build_hypergraph import ...`` keeps working for code written
against the reference (reference build_hypergraph/__init__.py)."""
from hg_b200 import *            # noqa: F401,F403
from hg_b200 import __all__      # noqa: F401
