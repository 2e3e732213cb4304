"""Part 1 entry point: single-process trainer (ref Part 1/main.py).

    python -m cs744_ddp_b200.parts.part1 [--synthetic ...]
"""
from .common import main as _main


def main(argv=None):
    return _main("1", argv)


if __name__ == "__main__":
    main()
