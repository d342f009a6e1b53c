"""``python -m src.backends.b200_launch train --videos a.mp4 …`` / ``… eval --model m.pth --sfm-cache c.pkl …``:
the reference's own ``src/train.py`` / ``src/eval.py`` main(), unmodified, with B200Backend registered
(run from a directory where both this repo and the reference tree are on ``sys.path``, or let
``dropin.reference_root()`` find the tree)."""
import sys

from src.backends.b200_backend import _pkg

if __name__ == "__main__":
    sys.exit(_pkg.dropin.main())
