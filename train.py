#!/usr/bin/env python3
"""Drop-in for the reference's ``train.py``: same CLI, same log lines, same files."""
import sys

from b200ddp.train import main

if __name__ == "__main__":
    sys.exit(main())
