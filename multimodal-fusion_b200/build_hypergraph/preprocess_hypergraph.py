"""`python -m build_hypergraph.preprocess_hypergraph` -- the reference's CLI entry point."""
from hg_b200.preprocess_hypergraph import *      # noqa: F401,F403
from hg_b200.preprocess_hypergraph import main

if __name__ == "__main__":
    main()
