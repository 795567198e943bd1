"""Shim package: makes ``from usearch.index import Index`` resolve to rad_b200's GPU-backed drop-in.

Put this directory's parent (``rad_b200/shim``) at the front of ``sys.path`` (or call ``rad_b200.shim.install()``);
tests, ``scripts/start_hnsw_server.py`` and the notebook of keiserlab/rad then run unmodified."""
from rad_b200.index import VERSION_MAJOR, VERSION_MINOR, VERSION_PATCH  # noqa: F401

__version__ = f"{VERSION_MAJOR}.{VERSION_MINOR}.{VERSION_PATCH}"
