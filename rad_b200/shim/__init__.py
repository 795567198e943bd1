"""``install()`` puts the ``usearch`` shim package at the front of ``sys.path``."""
import os
import sys


def install() -> str:
    here = os.path.dirname(os.path.abspath(__file__))
    if here not in sys.path:
        sys.path.insert(0, here)
    return here
