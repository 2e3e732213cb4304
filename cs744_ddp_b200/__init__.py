"""Importable alias for the framework package.

The framework lives in ``cs744-distributed-data-parallel_b200/`` (the directory name the
project brief asks for).  A hyphen is not legal in a Python identifier, so this thin alias
points its ``__path__`` at that directory and executes the real ``__init__``: every
sub-module is importable as ``cs744_ddp_b200.<module>``.
"""
import pathlib as _pathlib

_REAL = _pathlib.Path(__file__).resolve().parent.parent / "cs744-distributed-data-parallel_b200"
__path__ = [str(_REAL)]
__file__ = str(_REAL / "__init__.py")
exec(compile((_REAL / "__init__.py").read_text(), __file__, "exec"))
