#!/usr/bin/env python
"""Drop-in for the reference's ``execute.py`` batch driver (see tiresias_b200/cli/execute.py)."""
from tiresias_b200.cli.execute import main

if __name__ == "__main__":
    main()
