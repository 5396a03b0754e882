"""Host-side logic of the partitioned solve that needs no GPU: the RCB partitioner (host C++ inside libfeab.so),
the export blob layout, and the blob exchange over torch.distributed with the gloo backend at world_size 2."""
import ctypes as C
import os
import socket

import numpy as np
import pytest

from fea_visualizer_b200 import feab, meshes


@pytest.mark.parametrize("parts", [1, 2, 3, 4, 8])
def test_rcb_is_balanced_deterministic_and_compact(parts):
    xyz, tri = meshes.icosphere(24)
    p1 = feab.partition_rcb(xyz, parts)
    p2 = feab.partition_rcb(xyz.copy(), parts)
    assert (p1 == p2).all() and p1.min() == 0 and p1.max() == parts - 1
    cnt = np.bincount(p1, minlength=parts)
    assert cnt.max() - cnt.min() <= parts
    if parts > 1:
        cut = (p1[tri[:, 0]] != p1[tri[:, 1]]) | (p1[tri[:, 1]] != p1[tri[:, 2]])
        assert cut.mean() < 0.15           # compact pieces: few triangles straddle a cut
    with pytest.raises(feab.FeabError):
        feab.partition_rcb(xyz, 0)


def test_export_struct_matches_header():
    hdr = open(os.path.join(os.path.dirname(feab.HERE), "include", "feab.h")).read()
    assert "unsigned char ipc_handle[64]" in hdr and "uint64_t off_vec[5]" in hdr
    assert C.sizeof(feab.FeabMgExport) == 64 + 8 + 8 + 40 + 32 + 16


def _worker(rank, world, port, q):
    import torch.distributed as dist
    from fea_visualizer_b200.multigpu import gather_blobs
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    e = feab.FeabMgExport()
    e.rank, e.world, e.n_own, e.base_ptr = rank, world, 1000 + rank, 0xABC0 + rank
    blobs = gather_blobs(bytes(e), dist)
    got = [feab.FeabMgExport.from_buffer_copy(b) for b in blobs]
    q.put((rank, [(g.rank, g.world, g.n_own, g.base_ptr) for g in got]))
    dist.barrier()
    dist.destroy_process_group()


def test_blob_exchange_over_gloo_world_size_2():
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in ps:
        p.start()
    res = dict(q.get(timeout=120) for _ in range(2))
    for p in ps:
        p.join(timeout=60)
        assert p.exitcode == 0
    want = [(0, 2, 1000, 0xABC0), (1, 2, 1001, 0xABC1)]
    assert res[0] == want and res[1] == want
