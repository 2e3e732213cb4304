"""Part 2b entry point: per-parameter all-reduce averaging (ref Part 2b/main.py).

    python -m cs744_ddp_b200.parts.part2b --master IP --num-nodes N --rank R [--epochs E]
"""
from .common import main as _main


def main(argv=None):
    return _main("2b", argv)


if __name__ == "__main__":
    main()
