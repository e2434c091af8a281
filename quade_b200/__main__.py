"""python -m quade_b200 -c Conf.txt [-i]   (src/Quade.py:290-293); under torchrun: one process per GPU."""
import sys

from .quade import main

if __name__ == "__main__":
    sys.exit(main())
