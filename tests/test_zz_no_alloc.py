"""nerfb_pack_vol_static: the allocation-free variant of nerfb_pack_vol (scratch in a file-scope `static __device__` array
of the library, 2^24 voxels at most) against the existing entry point with a caller-provided workspace.  No counterpart in
the reference.  Collected last on purpose."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _lib():
    from nerf_siren_b200 import _lib
    return _lib.lib


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _volume(n, seed, frac_empty):
    g = torch.Generator(device=DEV).manual_seed(seed)
    v = torch.rand(n, 4, device=DEV, generator=g)
    v[:, 3] = v[:, 3] * 40 - 40 * frac_empty          # sigma <= 0 -> alpha == 0 -> voxel dropped
    return v.contiguous()


def _reference(lib, v, neg_step):
    """nerfb_pack_vol with a torch-allocated workspace (the existing entry point)."""
    n = v.shape[0]
    nbytes = int(lib.nerfb_pack_vol_workspace_bytes(n))
    ws = torch.empty(nbytes + 1024, dtype=torch.uint8, device=DEV)
    ws_ptr = (ws.data_ptr() + 255) & ~255
    n_valid = torch.zeros(1, dtype=torch.int64, device=DEV)
    assert lib.nerfb_pack_vol(v.data_ptr(), n, neg_step, 0, 0, n_valid.data_ptr(), ws_ptr, nbytes, _stream()) == 0
    k = int(n_valid.item())
    pairs = torch.zeros(max(k, 1), 2, dtype=torch.int64, device=DEV).to(torch.int32)     # 8 bytes per pair, as uint32 x 2
    assert lib.nerfb_pack_vol(v.data_ptr(), n, neg_step, pairs.data_ptr(), k, n_valid.data_ptr(), ws_ptr, nbytes, _stream()) == 0
    torch.cuda.synchronize()
    return k, pairs[:k].clone()


def _static(lib, v, neg_step, pairs, n_valid):
    """Two calls of the new entry (count, then write) into PRE-allocated outputs: the entry itself allocates nothing."""
    n = v.shape[0]
    assert lib.nerfb_pack_vol_static(v.data_ptr(), n, neg_step, 0, 0, n_valid.data_ptr(), _stream()) == 0
    k = int(n_valid.item())
    assert k <= pairs.shape[0]
    assert lib.nerfb_pack_vol_static(v.data_ptr(), n, neg_step, pairs.data_ptr(), k, n_valid.data_ptr(), _stream()) == 0
    torch.cuda.synchronize()
    return k


def test_pack_vol_static_equals_pack_vol_on_two_calls_in_a_row():
    """Integer / byte work: the bar is bit-exact.  Two DIFFERENT volumes back to back (the second smaller than the first, so
    stale counts / offsets of call one sit behind the part call two uses), then the first again: the shared scratch keeps
    the previous call's contents and nothing may depend on them.  Device memory in use must not grow across the calls."""
    lib = _lib()
    neg_step = -0.01
    vols = [_volume(300_000 + 77, 1, 0.6), _volume(64 * 64 * 64, 2, 0.3), _volume(300_000 + 77, 1, 0.6)]
    refs = [_reference(lib, v, neg_step) for v in vols]
    pairs = torch.zeros(max(k for k, _ in refs) + 1, 2, dtype=torch.int64, device=DEV).to(torch.int32)
    n_valid = torch.zeros(1, dtype=torch.int64, device=DEV)
    for v, (k_ref, p_ref) in zip(vols, refs):
        pairs.fill_(-1)
        torch.cuda.synchronize()
        free_before, _ = torch.cuda.mem_get_info()        # brackets ONLY the two calls of the new entry (torch ops in between
        k = _static(lib, v, neg_step, pairs, n_valid)      # the checks below may make the caching allocator grow)
        free_after, _ = torch.cuda.mem_get_info()
        assert free_after >= free_before, "the allocation-free entry point changed the device's free memory"
        assert k == k_ref and k > 0
        assert torch.equal(pairs[:k], p_ref)
        assert bool((pairs[k:] == -1).all()), "wrote past n_valid"


def test_pack_vol_static_limits_and_edges():
    lib = _lib()
    mx = int(lib.nerfb_pack_vol_static_max_voxels())
    assert mx == 1 << 24
    n_valid = torch.full((1,), -5, dtype=torch.int64, device=DEV)
    v = _volume(1024, 3, 0.5)
    # beyond the built-in scratch: refused (NERFB_E_WORKSPACE = -3) before anything is launched or read
    assert lib.nerfb_pack_vol_static(v.data_ptr(), mx + 1, -0.01, 0, 0, n_valid.data_ptr(), _stream()) == -3
    torch.cuda.synchronize()
    assert int(n_valid.item()) == -5
    # empty volume, and an all-empty one
    assert lib.nerfb_pack_vol_static(v.data_ptr(), 0, -0.01, 0, 0, n_valid.data_ptr(), _stream()) == 0
    torch.cuda.synchronize()
    assert int(n_valid.item()) == 0
    v[:, 3] = -1.0
    assert lib.nerfb_pack_vol_static(v.data_ptr(), 1024, -0.01, 0, 0, n_valid.data_ptr(), _stream()) == 0
    torch.cuda.synchronize()
    assert int(n_valid.item()) == 0
    # the existing entry point agrees on the empty volume (it used to sum a stale count of its workspace)
    ws = torch.full((4096,), 0x7F, dtype=torch.uint8, device=DEV)
    ws_ptr = (ws.data_ptr() + 255) & ~255
    n_valid.fill_(-7)
    assert lib.nerfb_pack_vol(v.data_ptr(), 0, -0.01, 0, 0, n_valid.data_ptr(), ws_ptr, 2048, _stream()) == 0
    torch.cuda.synchronize()
    assert int(n_valid.item()) == 0
    # bad arguments
    assert lib.nerfb_pack_vol_static(0, 16, -0.01, 0, 0, n_valid.data_ptr(), _stream()) == -1
    assert lib.nerfb_pack_vol_static(v.data_ptr(), 16, -0.01, 0, 4, n_valid.data_ptr(), _stream()) == -1     # capacity without pairs


def test_pack_vol_static_at_the_maximum_size():
    """2^24 voxels (256 MB of input): the whole scratch array is in use; count and a sample of the pairs against the
    existing entry point."""
    lib = _lib()
    n = 1 << 24
    v = _volume(n, 4, 0.9)
    k_ref, p_ref = _reference(lib, v, -0.02)
    pairs = torch.zeros(k_ref + 1, 2, dtype=torch.int64, device=DEV).to(torch.int32)
    n_valid = torch.zeros(1, dtype=torch.int64, device=DEV)
    k = _static(lib, v, -0.02, pairs, n_valid)
    assert k == k_ref
    assert torch.equal(pairs[:k], p_ref)
