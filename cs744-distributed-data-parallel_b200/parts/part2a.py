"""Part 2a entry point: rank-0 gather -> mean -> scatter averaging (ref Part 2a/main.py).

    python -m cs744_ddp_b200.parts.part2a --master IP --num-nodes N --rank R [--epochs E]
"""
from .common import main as _main


def main(argv=None):
    return _main("2a", argv)


if __name__ == "__main__":
    main()
