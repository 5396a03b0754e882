"""GPU test of feas_detect_boundary_noalloc (include/feab_surface.h), the one entry point of libfeab that allocates no
GPU memory at all.  It exists for host programs that cannot call cudaMalloc themselves and cannot let a library do it
for them either: a render loop that cudaMalloc's device-wide synchronisation would stall, a plugin inside an
application whose own pool already owns the whole GPU, code that must stay out of the CUDA allocator after start-up.
Such a caller hands over device buffers it already has; the edge table the algorithm needs sits in a fixed 24 MiB array
inside the library, shared by every call.

What is checked, all through the C ABI: the flags and the boundary count equal feas_detect_boundary's (the allocating
entry point; integer work, so the bar is bit-exact) on several calls IN A ROW -- large mesh, then a small one, then
the large one again, so a call that failed to wipe what its predecessor left in the shared table would be caught; that
meshes beyond the fixed maximum and out-of-range indices are refused with a message; that free device memory does not
move across a call.  The "caller" here is torch, which plays the application that owns the device buffers.

The file name sorts last on purpose: pytest -x stops at the first failure, so this test must not run ahead of (and
hide) the parity tests.  It passes alone in a fresh process and after the rest of the suite.
"""
import ctypes as C

import numpy as np
import pytest

from fea_visualizer_b200 import feab, meshes
from fea_visualizer_b200 import surface as S

pytestmark = pytest.mark.gpu


def caller_buffers(tri, num_vertices):
    """The caller's own device memory: the triangle array and a flag buffer pre-filled with junk."""
    import torch
    t = torch.from_numpy(np.ascontiguousarray(tri, dtype=np.uint32).view(np.int32).reshape(-1)).to("cuda:0")
    out = torch.full((num_vertices,), 0x5A, dtype=torch.uint8, device="cuda:0")
    torch.cuda.synchronize()
    return t, out


def with_allocation(xyz, tri):
    """feas_detect_boundary: the existing entry point, which cudaMallocs its edge table per call."""
    D = S.DeviceSurface()
    D.set_mesh(xyz, None, tri)
    ob, nb, _ = D.detect_boundary()
    D.close()
    return ob, nb


def without_allocation(xyz, tri):
    t, out = caller_buffers(tri, len(xyz))
    nb = S.detect_boundary_noalloc(0, 0, len(xyz), len(tri), t.data_ptr(), out.data_ptr())
    return out.cpu().numpy(), nb


def open_sphere(nu):
    """An icosphere with every 7th triangle removed: plenty of boundary edges, irregular pattern."""
    xyz, tri = meshes.icosphere(nu)[:2]
    tri = np.ascontiguousarray(tri[np.arange(len(tri)) % 7 != 0])
    return xyz, tri


def test_same_flags_as_the_allocating_entry_point_on_calls_in_a_row():
    big = open_sphere(130)                      # 289 715 triangles: uses the whole 2^21-slot table
    assert 6 * len(big[1]) > 2 ** 20 and len(big[1]) <= S.NOALLOC_MAX_TRIANGLES
    gx, gt = meshes.grid(40, 30)[:2]            # small open mesh: uses a corner of the table the big one filled
    closed = meshes.icosphere(6)[:2]            # closed: no boundary at all
    dx, dt = meshes.disk(12)[:2]
    sequence = [big, (gx, gt), big, closed, (dx, dt), (gx, gt)]
    expected = [with_allocation(x, t) for x, t in sequence]
    assert expected[0][1] > 0 and expected[1][1] > 0 and expected[3][1] == 0
    for k, ((x, t), (ob_ref, nb_ref)) in enumerate(zip(sequence, expected)):
        ob, nb = without_allocation(x, t)
        assert nb == nb_ref, (k, nb, nb_ref)
        assert np.array_equal(ob, ob_ref), (k, int((ob != ob_ref).sum()))
        assert set(np.unique(ob)) <= {0, 1}     # the junk the caller left in its buffer is gone


def test_degenerate_edges_and_empty_triangle_list():
    xyz, tri = meshes.grid(9, 7)[:2]
    tri = np.array(tri, dtype=np.uint32)
    tri[3] = [tri[3][0], tri[3][0], tri[3][1]]  # an edge (a, a): counts twice, like the reference's edges[a][a] += 2
    ob_ref, nb_ref = with_allocation(xyz, tri)
    for _ in range(2):
        ob, nb = without_allocation(xyz, tri)
        assert nb == nb_ref and np.array_equal(ob, ob_ref)
    _, out = caller_buffers(tri, len(xyz))
    assert S.detect_boundary_noalloc(0, 0, len(xyz), 0, 0, out.data_ptr()) == 0
    assert int(out.sum().item()) == 0


def test_beyond_the_fixed_table_is_refused_and_the_next_call_still_works():
    xyz, tri = meshes.grid(9, 7)[:2]
    t, out = caller_buffers(tri, len(xyz))
    with pytest.raises(feab.FeabError) as e:
        S.detect_boundary_noalloc(0, 0, 10, S.NOALLOC_MAX_TRIANGLES + 1, t.data_ptr(), out.data_ptr())
    assert e.value.code == -1 and "FEAS_NOALLOC_MAX_TRIANGLES" in str(e.value)
    assert int((out != 0x5A).sum().item()) == 0                     # refused before anything was touched
    bad = np.array(tri, dtype=np.uint32)
    bad[5, 1] = len(xyz)                                            # one index past the end
    tb, _ = caller_buffers(bad, len(xyz))
    with pytest.raises(feab.FeabError) as e:
        S.detect_boundary_noalloc(0, 0, len(xyz), len(bad), tb.data_ptr(), out.data_ptr())
    assert e.value.code == -1 and "index" in str(e.value)
    ob_ref, nb_ref = with_allocation(xyz, tri)                      # the error flag of that call does not stick
    ob, nb = without_allocation(xyz, tri)
    assert nb == nb_ref and np.array_equal(ob, ob_ref)


def test_free_device_memory_does_not_move_across_a_call():
    import torch
    xyz, tri = open_sphere(60)
    t, out = caller_buffers(tri, len(xyz))
    S.detect_boundary_noalloc(0, 0, len(xyz), len(tri), t.data_ptr(), out.data_ptr())   # first call: one-time set-up
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info(0)[0]
    for _ in range(3):
        S.detect_boundary_noalloc(0, 0, len(xyz), len(tri), t.data_ptr(), out.data_ptr())
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info(0)[0] == free0
