"""Part 3 entry point: bucketed all-reduce overlapped with backward (ref Part 3/main.py).

    python -m cs744_ddp_b200.parts.part3 --master IP --num-nodes N --rank R [--epochs E]
"""
from .common import main as _main


def main(argv=None):
    return _main("3", argv)


if __name__ == "__main__":
    main()
