"""``python -m b200dp.launch -np N [-- cmd ...]`` -- see parallel/launcher.py."""
import sys

from .parallel.launcher import main

if __name__ == "__main__":
    sys.exit(main())
