"""CPU tests for the host-side mesh generators / ingest."""
import numpy as np
import pytest

from fea_visualizer_b200 import meshes


@pytest.mark.parametrize("nu", [1, 2, 3, 8, 17])
def test_icosphere_counts_and_topology(nu):
    xyz, tri = meshes.icosphere(nu)
    V, T = 10 * nu * nu + 2, 20 * nu * nu
    assert xyz.shape == (V, 3) and tri.shape == (T, 3) and xyz.dtype == np.float32 and tri.dtype == np.uint32
    assert np.allclose(np.linalg.norm(xyz.astype(np.float64), axis=1), 1.0, atol=1e-6)
    e = np.concatenate([tri[:, [0, 1]], tri[:, [1, 2]], tri[:, [2, 0]]]).astype(np.int64)
    # closed, consistently oriented 2-manifold: every directed edge once, its reverse once
    key = e[:, 0] * V + e[:, 1]
    assert len(np.unique(key)) == len(key)
    rkey = e[:, 1] * V + e[:, 0]
    assert set(key.tolist()) == set(rkey.tolist())
    assert V - len(key) // 2 + T == 2                      # Euler characteristic of the sphere
    assert meshes.boundary_flags(V, tri).sum() == 0
    # outward winding
    p = xyz.astype(np.float64)
    nrm = np.cross(p[tri[:, 1]] - p[tri[:, 0]], p[tri[:, 2]] - p[tri[:, 0]])
    assert (np.einsum("ij,ij->i", nrm, p[tri].mean(axis=1)) > 0).all()
    assert len(np.unique(tri)) == V


def test_icosphere_sizes_of_the_baseline_configs():
    for nu, V in ((32, 10242), (60, 36002), (316, 998562), (1265, 16002252), (2530, 64009002)):
        assert 10 * nu * nu + 2 == V


def test_grid_and_disk_boundaries():
    xyz, tri = meshes.grid(12, 9)
    ob = meshes.boundary_flags(len(xyz), tri)
    assert ob.sum() == 2 * (12 + 9) and (xyz[:, 1] == 0).all()
    xyz, tri = meshes.disk(6)
    ob = meshes.boundary_flags(len(xyz), tri)
    assert len(xyz) == 1 + 3 * 6 * 7 and len(tri) == 6 * 36 and ob.sum() == 36


def test_obj_ingest(tmp_path):
    p = tmp_path / "q.obj"
    p.write_text("v 0 0 0\nv 1 0 0\nv 1 1 0\nv 0 1 0\nvn 0 0 1\nf 1//1 2//1 3//1 4//1\n")
    xyz, tri = meshes.load_obj(str(p))
    assert xyz.shape == (4, 3) and tri.tolist() == [[0, 1, 2], [0, 2, 3]]
    assert meshes.boundary_flags(4, tri).sum() == 4
    p.write_text("v 0 0 0\nv 1 0 0\nv 1 1 0\nf 1 2 3\n")
    with pytest.raises(RuntimeError):      # the reference rejects OBJ files without normals (Surface.cpp:73-74)
        meshes.load_obj(str(p))


def test_seed_nodes():
    s = meshes.seed_nodes(10242)
    assert len(s) == 50 and s[0] == 0 and s[1] == 2654435761 % 10242 and s.max() < 10242
