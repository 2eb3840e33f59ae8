#!/usr/bin/env python
"""Tutorial entry point `part2a`: parameter server: rank 0 gathers, averages and scatters every gradient.

Same command line as the reference script it replaces (ref: master/part2a/part2a.py + slave/part2a/part2a.py):

    python examples/part2a.py --master-ip 10.10.1.1 --num-nodes 4 --rank 0

One file serves every rank (the reference keeps separate master/ and slave/ copies; rank 0 is the
master).  All other options of `python -m cs744_pytorch_distributed_tutorial_b200.cli` are accepted
as well, e.g. `--per-rank-batch 256`, `--iters 40`, `--device cpu`.
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from cs744_pytorch_distributed_tutorial_b200.cli import main  # noqa: E402

if __name__ == "__main__":
    sys.exit(main(["--mode", "part2a"] + sys.argv[1:]))
