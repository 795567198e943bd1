"""RAD's own service layer (UNMODIFIED rad/hnsw_service.py and rad/hnsw_server.py, imported from the reference tree) on
top of rad_b200.index.Index, beside the same services on top of the reference's usearch Index, on the same `.usearch`
file.  This is the duck-typed contract of SURVEY.md §8b: RAD reads the graph through `get_neighbors`,
`get_top_level_nodes`, `len`, `max_level`, `levels_stats`, ... and nothing else (rad/hnsw_service.py:219-232, 400-412;
rad/hnsw_server.py:148-161, 196, 471-483, 593-602).

CPU only (graph reads never touch CUDA), over the reference-built golden file.  The same services over an index the
GPU BUILT are exercised on the GPU box by tests/test_rad_gpu_e2e.py."""
import multiprocessing
import os
import socket
import sys
import time

import numpy as np
import pytest

from tests.refpath import rad_root

REFERENCE = rad_root()
pytestmark = pytest.mark.skipif(REFERENCE is None, reason="the reference's rad package is not present on this machine")


@pytest.fixture(scope="module")
def rad_modules():
    sys.path.insert(0, REFERENCE)
    try:
        import rad.hnsw_service as svc  # noqa: PLC0415  (the reference's module, unmodified)
        yield svc
    finally:
        sys.path.remove(REFERENCE)


@pytest.fixture(scope="module")
def indexes(golden_dir):
    """(ours, reference-or-None) viewing the same reference-written file, the way scripts/start_hnsw_server.py:69 does."""
    from rad_b200.index import Index
    path = os.path.join(golden_dir, "ref_n1000_m8.usearch")
    ours = Index(path=path, view=True, exclude_vectors=True)
    ref = None
    try:
        from oracle import oracle as O
        if O.ref_available():
            ref = O.ref_index_module().Index(path=path, view=True, exclude_vectors=True)
    except Exception:  # noqa: BLE001 -- without the compiled reference the golden arrays are the yardstick
        ref = None
    return ours, ref


def _pairs(seq):
    return [int(x) for x in seq]


def test_local_service_answers_equal_the_reference(rad_modules, indexes, golden_dir):
    ours, ref = indexes
    gold = np.load(os.path.join(golden_dir, "ref_n1000_m8.npz"))
    levels = gold["levels"]
    service = rad_modules.LocalHNSWService(ours)
    ref_service = rad_modules.LocalHNSWService(ref) if ref is not None else None
    try:
        assert service.is_healthy() is True
        info = service.get_service_info()
        assert info["service_type"] == "LocalHNSWService" and info["status"] == "running"
        top = service.get_top_level_nodes()
        assert top and len(top) % 2 == 0
        expected_top = [int(s) for s in np.nonzero(levels == levels.max())[0]]
        assert [int(x) for x in top[0::2]] == expected_top
        if ref_service is not None:
            assert list(top) == list(ref_service.get_top_level_nodes())
        # every node of the upper levels and a slice of level 0, through the service's request/response queues
        probes = [(int(s), int(l)) for s in np.nonzero(levels > 0)[0] for l in range(int(levels[s]) + 1)]
        probes += [(s, 0) for s in range(0, len(levels), 7)]
        for slot, level in probes:
            # the service answers [node_id, smiles, ...] (smiles = '' without a database: hnsw_service.py:270-283)
            got = list(service.get_neighbors(slot, level))
            assert len(got) % 2 == 0
            if ref_service is not None:
                assert got == list(ref_service.get_neighbors(slot, level)), (slot, level)
            assert [int(x) for x in got[0::2]] == _pairs(ours.get_neighbors(slot, level))[0::2]
    finally:
        service.shutdown()
        if ref_service is not None:
            ref_service.shutdown()


def test_direct_reads_equal_the_reference_everywhere(indexes):
    ours, ref = indexes
    if ref is None:
        pytest.skip("compiled reference binding not available")
    assert len(ours) == len(ref) and ours.max_level == ref.max_level
    assert ours.connectivity == ref.connectivity and ours.ndim == ref.ndim and str(ours.dtype) == str(ref.dtype)
    assert ours.multi == ref.multi and ours.capacity >= len(ours)
    assert _pairs(ours.get_top_level_nodes()) == _pairs(ref.get_top_level_nodes())
    for slot in range(len(ours)):
        assert _pairs(ours.get_neighbors(slot, 0)) == _pairs(ref.get_neighbors(slot, 0))
    mine, theirs = ours.levels_stats, ref.levels_stats
    assert len(mine) == len(theirs)
    for a, b in zip(mine, theirs):
        assert (a.nodes, a.edges, a.max_edges, a.allocated_bytes) == (b.nodes, b.edges, b.max_edges, b.allocated_bytes)
    # the error RAD's server relies on (rad/hnsw_server.py:481-487): out-of-range level -> an exception, same type
    for bad in ((0, 99), (len(ours) + 5, 0)):
        with pytest.raises(Exception) as e1:
            ours.get_neighbors(*bad)
        with pytest.raises(Exception) as e2:
            ref.get_neighbors(*bad)
        assert type(e1.value) is type(e2.value)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _serve(path, port, workdir):
    os.chdir(workdir)  # the server writes hnsw_top_nodes_cache_*.json into its working directory (rad/hnsw_server.py:171)
    sys.path.insert(0, REFERENCE)
    from rad.hnsw_server import run_hnsw_server  # noqa: PLC0415
    from rad_b200.index import Index  # noqa: PLC0415
    run_hnsw_server(hnsw=Index(path=path, view=True, exclude_vectors=True), host="127.0.0.1", port=port, workers=1,
                    debug=False)


def test_http_server_and_remote_client(rad_modules, indexes, golden_dir, tmp_path):
    """rad/hnsw_server.py (FastAPI) serving rad_b200's Index in a forked child, rad's RemoteHNSWService as the client
    (the reference's tests/test_remote_hnsw.py:54-120 flow)."""
    pytest.importorskip("fastapi")
    pytest.importorskip("uvicorn")
    pytest.importorskip("requests")
    ours, _ = indexes
    port = _free_port()
    path = os.path.join(golden_dir, "ref_n1000_m8.usearch")
    ctx = multiprocessing.get_context("fork")
    proc = ctx.Process(target=_serve, args=(path, port, str(tmp_path)), daemon=True)
    proc.start()
    try:
        client = None
        deadline = time.time() + 30
        while time.time() < deadline:
            try:
                client = rad_modules.RemoteHNSWService(f"http://127.0.0.1:{port}")
                if client.is_healthy():
                    break
            except Exception:  # noqa: BLE001 -- server still starting
                client = None
            time.sleep(0.3)
        assert client is not None and client.is_healthy()
        top = client.get_top_level_nodes()
        assert [int(x) for x in top[0::2]] == _pairs(ours.get_top_level_nodes())[0::2]
        for slot in (0, 1, 17, 500, 999):
            got = client.get_neighbors(slot, 0)
            assert [int(x) for x in got[0::2]] == _pairs(ours.get_neighbors(slot, 0))[0::2]
        info = client.get_service_info()
        assert "client_info" in info and "remote_service_info" in info
        client.shutdown()
    finally:
        proc.terminate()
        proc.join(10)
