#!/usr/bin/env python
"""Drop-in for the reference's ``run_sim.py``: same flags, same log layout (see
tiresias_b200/cli/run_sim.py)."""
from tiresias_b200.cli.run_sim import main

if __name__ == "__main__":
    main()
