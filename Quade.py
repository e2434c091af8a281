#!/usr/bin/env python
"""Drop-in for the reference's src/Quade.py command line: `Quade.py -c Conf.txt [-i]`."""
import sys

from quade_b200.quade import Quade, main  # noqa: F401

if __name__ == "__main__":
    sys.exit(main())
