"""Expert-parallel host logic on CPU: 2 ranks over gloo.  The split-size plan (mxmoe_b200.ep.ep_plan) and the
(source, expert)-segment ordering are exercised with the oracle's permute/combine and a stand-in expert
function; the result must equal the single-process oracle."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import oracle as O


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _expert_fn(rows: np.ndarray, e: int) -> np.ndarray:
    return (rows.astype(np.float32) * (1.0 + 0.25 * e) + e).astype(np.float16)


def _worker(rank, world, port, E, k, T, Hd, q):
    from mxmoe_b200.ep import ep_plan, expert_bounds
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rng = np.random.default_rng(100 + rank)
        x = rng.standard_normal((T, Hd)).astype(np.float16)
        ids = np.stack([rng.permutation(E)[:k] for _ in range(T)]).astype(np.int32)
        w = rng.random((T, k)).astype(np.float32)
        offs, src, slot = O.permute(ids, E)
        bounds = expert_bounds(E, world)
        my_counts = torch.from_numpy(np.diff(offs).astype(np.int32))
        allc = torch.zeros(world, E, dtype=torch.int32)
        dist.all_gather_into_tensor(allc.view(-1), my_counts)
        send, recv, seg = ep_plan(allc.numpy(), bounds, rank)
        assert sum(send) == T * k
        rows = torch.from_numpy(x[src])
        got = torch.zeros(sum(recv), Hd, dtype=torch.float16)
        dist.all_to_all_single(got, rows, output_split_sizes=recv, input_split_sizes=send)
        # segments arrive source-major, local-expert-minor
        seg_offs = np.concatenate([[0], np.cumsum(seg)])
        EL = bounds[rank + 1] - bounds[rank]
        outr = np.zeros((sum(recv), Hd), np.float16)
        g = got.numpy()
        for s in range(world):
            for i in range(EL):
                a, b = seg_offs[s * EL + i], seg_offs[s * EL + i + 1]
                outr[a:b] = _expert_fn(g[a:b], bounds[rank] + i)
        back = torch.zeros(T * k, Hd, dtype=torch.float16)
        dist.all_to_all_single(back, torch.from_numpy(outr), output_split_sizes=send, input_split_sizes=recv)
        y = O.combine(back.numpy(), slot, w)
        # single-process reference
        ref_rows = np.zeros((T * k, Hd), np.float16)
        xs = x[src]
        for e in range(E):
            ref_rows[offs[e]:offs[e + 1]] = _expert_fn(xs[offs[e]:offs[e + 1]], e)
        ref = O.combine(ref_rows, slot, w)
        q.put((rank, bool(np.array_equal(y.view(np.uint16), ref.view(np.uint16)))))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("E,k", [(6, 2), (7, 3)])
def test_ep_dispatch_combine_two_ranks_gloo(E, k):
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, E, k, 33, 16, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
    assert sorted(res) == [(0, True), (1, True)]


def test_ep_plan_shapes():
    from mxmoe_b200.ep import ep_plan, expert_bounds
    b = expert_bounds(60, 8)
    assert b[0] == 0 and b[-1] == 60 and all(7 <= b[i + 1] - b[i] <= 8 for i in range(8))
    counts = np.arange(8 * 60).reshape(8, 60)
    send, recv, seg = ep_plan(counts, b, 3)
    assert sum(send) == counts[3].sum() and sum(recv) == counts[:, b[3]:b[4]].sum() and len(seg) == 8 * (b[4] - b[3])


def test_balanced_expert_order_is_a_permutation_and_balances():
    """Placement from a popularity profile: every rank keeps its slice SIZE (expert_bounds), the most loaded rank gets
    closer to the mean than with contiguous placement, and the result is a permutation (the relabelling is exact)."""
    import numpy as np
    from mxmoe_b200 import workload as Wl
    from mxmoe_b200.ep import balanced_expert_order, expert_bounds
    freq = np.array(Wl.synthetic_gate_trace("qwen2_moe", [0], "fixture")["layer-0"]["access_freq"], dtype=np.float64)
    E = len(freq)
    for W in (2, 4, 8):
        order = balanced_expert_order(freq, W)
        assert sorted(order) == list(range(E))
        b = expert_bounds(E, W)
        cont = max(freq[b[r]:b[r + 1]].sum() for r in range(W))
        bal = max(freq[order[b[r]:b[r + 1]]].sum() for r in range(W))
        assert bal <= cont + 1e-12
    assert balanced_expert_order([1.0] * 6, 3) == [0, 3, 1, 4, 2, 5]
