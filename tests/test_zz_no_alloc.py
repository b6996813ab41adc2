"""orb_assign_features_to_grid_static, the allocation-free entry point (static __device__ scratch array instead of a workspace):
its grid must be the one orb_assign_features_to_grid builds, on calls in a row (the array keeps the previous call's contents),
with host and with device pointers, and sizes beyond the array's maximum must be refused.  Collected last on purpose."""
import ctypes as C
import importlib
import numpy as np
import pytest

capi = importlib.import_module("orb_slam2-chinese-annotation_b200.capi")
synth = importlib.import_module("orb_slam2-chinese-annotation_b200.synth")
views = importlib.import_module("orb_slam2-chinese-annotation_b200.views")
matcher = importlib.import_module("orb_slam2-chinese-annotation_b200.matcher")
pytestmark = pytest.mark.gpu

CELLS = 64 * 48


def _frame(n, idx):
    d = synth.make_frame_arrays(n, idx)
    d["x"][: n // 10] = np.float32(d["bounds"][2]) - 0.2   # round() lands on column 64 -> dropped
    return views.Frame(d["x"], d["y"], d["octave"], d["angle"], d["desc"], d["sf"], d["bounds"], d["u_right"])


def _static_host(f):
    cs = np.full(CELLS + 1, -7, np.int32); ci = np.full(max(f.N, 1), -7, np.int32); n = np.full(1, -7, np.int32)
    g = f.view().grid
    matcher.assign_features_to_grid_static(f.x, f.y, f.N, g, cs, ci, n)
    return cs, ci, int(n[0])


def test_static_grid_equals_the_workspace_entry_point_on_calls_in_a_row():
    gm = matcher.ORBmatcher(0.8, True, max_n=8192)
    # a large frame first, smaller ones after it: stale items of the earlier call stay in the array behind the new ones
    for idx, n in [(0, 2000), (1, 777), (2, 1), (3, 4096), (4, 8192), (5, 300)]:
        f = _frame(n, idx)
        ws, wi = gm.assign_features_to_grid(f)
        for _ in range(2):
            cs, ci, tot = _static_host(f)
            assert tot == len(wi) and np.array_equal(cs, ws) and np.array_equal(ci[:tot], wi), (idx, n)
            assert np.all(ci[tot:] == -7)   # nothing of an earlier call is copied out
    gm.close()


def test_static_grid_with_device_pointers_and_a_stream():
    import torch
    gm = matcher.ORBmatcher(0.8, True, max_n=4096)
    st = torch.cuda.Stream()
    for idx, n in [(6, 1500), (7, 64)]:
        f = _frame(n, idx)
        ws, wi = gm.assign_features_to_grid(f)
        dx = torch.from_numpy(f.x).cuda(); dy = torch.from_numpy(f.y).cuda()
        dcs = torch.full((CELLS + 1,), -7, dtype=torch.int32, device="cuda")
        dci = torch.full((n,), -7, dtype=torch.int32, device="cuda")
        dn = torch.full((1,), -7, dtype=torch.int32, device="cuda")
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated()
        matcher.assign_features_to_grid_static(dx.data_ptr(), dy.data_ptr(), n, f.view().grid, dcs.data_ptr(), dci.data_ptr(),
                                               dn.data_ptr(), stream=st.cuda_stream)
        assert torch.cuda.memory_allocated() == before
        tot = int(dn.item())
        assert tot == len(wi) and np.array_equal(dcs.cpu().numpy(), ws) and np.array_equal(dci.cpu().numpy()[:tot], wi)
    gm.close()


def test_static_grid_empty_frame_and_limits():
    f = _frame(10, 8)
    g = f.view().grid
    cs = np.full(CELLS + 1, -7, np.int32); ci = np.full(1, -7, np.int32); n = np.full(1, -7, np.int32)
    matcher.assign_features_to_grid_static(None, None, 0, g, cs, ci, n)
    assert n[0] == 0 and not cs.any()
    big = np.zeros(matcher.GRID_STATIC_MAX_N + 1, np.float32)
    out = np.zeros(len(big), np.int32)
    with pytest.raises(capi.OrbError) as e:
        matcher.assign_features_to_grid_static(big, big, len(big), g, cs, out, n)
    assert e.value.status == capi.ORB_ERR_CAPACITY and "maximum of 8192" in str(e.value)
    with pytest.raises(capi.OrbError) as e:
        matcher.assign_features_to_grid_static(big, big, 5, g, None, out, n)
    assert e.value.status == capi.ORB_ERR_BAD_ARG
    # the refused calls left the entry point usable
    cs2, ci2, tot = _static_host(f)
    assert 0 < tot <= 10 and cs2[-1] == tot and sorted(ci2[:tot]) == sorted(set(ci2[:tot]))
