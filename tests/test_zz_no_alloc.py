"""mxb_permute_static: the token -> expert permute with its scratch in a file-scope __device__ array (no handle, no
cudaMalloc; include/mxmoe_b200.h).  Collected LAST (file name) on purpose: the entry point has no counterpart in the
reference and a failure here must not hide the parity tests under `pytest -x`.

Checked through the C-ABI against the existing entry point mxb_permute on the same inputs — integer work, so the bar is
bit-exact — on two calls in a row (the array keeps the first call's contents), at sizes that take the two-kernel form
which is the one that uses the scratch (T * k > 2048), plus the single-block form and the empty batch."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ids(T, k, E, seed):
    rng = np.random.default_rng(seed)
    return np.stack([rng.permutation(E)[:k] for _ in range(T)]).astype(np.int32) if T else np.zeros((0, k), np.int32)


def _both(h, ops, ids_np, E):
    T, k = ids_np.shape
    ids = torch.from_numpy(ids_np).cuda()
    ref = h.permute(ids, E)  # existing entry point (workspace cudaMalloc'ed by mxb_create)
    offs = torch.full((E + 1,), -7, dtype=torch.int32, device="cuda")
    src = torch.full((max(T * k, 1),), -7, dtype=torch.int32, device="cuda")
    slot = torch.full((max(T, 1), k), -7, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    ops.permute_static(ids, E, offs, src, slot)
    assert torch.cuda.memory_allocated() == before
    torch.cuda.synchronize()
    assert torch.equal(offs, ref[0])
    assert torch.equal(src[:T * k], ref[1][:T * k])
    assert torch.equal(slot[:T], ref[2][:T])
    return offs.cpu().numpy()


@pytest.mark.parametrize("T,k,E", [(8192, 4, 60), (2048, 8, 64), (5000, 6, 64)])
def test_permute_static_matches_permute_twice(T, k, E):
    from mxmoe_b200 import ops
    h = ops.Handle(0)
    a = _both(h, ops, _ids(T, k, E, 1), E)
    b = _both(h, ops, _ids(T, k, E, 2), E)  # second call in a row: the scratch holds the first call's histograms
    assert a[-1] == T * k and b[-1] == T * k
    assert not np.array_equal(a, b)  # the two inputs route differently, so a stale histogram would show


def test_permute_static_sizes_in_sequence():
    """large -> small -> empty -> large on the one shared array, with a skewed routing (one hot expert)"""
    from mxmoe_b200 import ops
    h = ops.Handle(0)
    E = 64
    hot = _ids(6000, 6, E, 3)
    hot[::2, 0] = 5  # (a token may then name expert 5 twice: the permute is a stable sort of the flattened ids either way)
    for ids in (hot, _ids(64, 6, E, 4), _ids(0, 6, E, 5), _ids(16384, 2, E, 6), _ids(1, 6, E, 7)):
        _both(h, ops, ids, E)


def test_permute_static_rejects_beyond_the_static_array():
    from mxmoe_b200 import ops
    lib = ops.load_library()
    T, k, E = 65536, 2, 1024  # 64 blocks x 1024 experts + the ticket = 65537 ints > 256 KiB
    ids = torch.zeros(T, k, dtype=torch.int32, device="cuda")
    offs = torch.zeros(E + 1, dtype=torch.int32, device="cuda")
    src = torch.zeros(T * k, dtype=torch.int32, device="cuda")
    slot = torch.zeros(T, k, dtype=torch.int32, device="cuda")
    with pytest.raises(ops.MxbError, match="static array holds"):
        ops.permute_static(ids, E, offs, src, slot)
    assert lib.mxb_permute_static(64, None, 0, 1, 4, offs.data_ptr(), None, None, None) != 0  # device outside [0, 64)
    # one block fewer fits, and the library still works afterwards
    T = 65536 - 2048
    ops.permute_static(ids[:T], E, offs, src, slot)
    torch.cuda.synchronize()
    assert int(offs[1]) == T * k  # every id is expert 0
