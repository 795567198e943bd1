"""Where the reference's UNMODIFIED pure-Python `rad` package can be imported from: the reference tree itself in the
build container, or the copy oracle/build_ref.sh stages under the git-ignored oracle/_ref/ (what travels to the GPU
box).  Test infrastructure only."""
import os

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def rad_root():
    """Directory to put on sys.path so that `import rad.traverser` finds the reference's modules, or None."""
    for root in ("/root/reference", os.path.join(_ROOT, "oracle", "_ref")):
        if os.path.isfile(os.path.join(root, "rad", "traverser.py")):
            return root
    return None
