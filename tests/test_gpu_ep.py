"""Expert-parallel layer on 2 GPUs (NCCL) against the single-GPU layer with the same weights: every rank's output
rows must be bit-identical (all per-row math is independent of how rows are grouped).  Skipped with < 2 GPUs."""
import os
import socket

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _lin(w, a, g=-1, sym=True):
    return {"w_bits": w, "w_gsize": g, "w_sym": sym, "w_clip": [1.0, 1.0], "a_bits": a, "a_gsize": g, "a_sym": True, "a_clip": [1.0, 1.0]}


def _worker(rank, world, port, mixed_f16, exchange, E, q):
    import torch.distributed as dist
    from mxmoe_b200 import ops
    from mxmoe_b200.config import qmodel_cfg_from_dict
    from mxmoe_b200.ep import EPMoELayer
    from mxmoe_b200.layer import MoELayer
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        k, Hd, I, ns, T = 2, 256, 128, 2, 300  # T*k > 512 rows: both layers take the tcgen05 path for the w4a16 expert
        cfgs = [(_lin(8, 8), _lin(4, 4)), (_lin(4, 4), _lin(8, 8)), (_lin(4, 4), _lin(4, 4))]
        if mixed_f16:
            cfgs.append((_lin(16, 16), _lin(4, 16, 128, False)))
        ex = {}
        for e in range(E + 1):
            gu, dn = cfgs[e % len(cfgs)]
            ex[str(e)] = {"gate": gu, "up": dict(gu), "down": dn}
        ql = qmodel_cfg_from_dict({"0": {"experts": ex}}).layers["0"]
        g = torch.Generator().manual_seed(5)
        weights = {}
        for e in range(E + 1):
            Ie = I * ns if e == E else I
            weights[e] = (((torch.rand(2 * Ie, Hd, generator=g) * 2 - 1) * 0.12).half(), ((torch.rand(Hd, Ie, generator=g) * 2 - 1) * 0.18).half())
        h = ops.Handle(rank)
        ep = EPMoELayer(h, E, k, Hd, I, ns, ql, max_tokens=T, hadamard_down=False, weights=weights, exchange=exchange,
                        expert_order=[(3 * i + 1) % E for i in range(E)] if E == 7 else None)  # E = 7: a scattered placement
        full = MoELayer(h, E, k, Hd, I, ns, ql, max_tokens=T, weights=weights)
        rng = np.random.default_rng(10 + rank)
        x = torch.from_numpy((rng.random((T, Hd)) * 2 - 1).astype(np.float16)).cuda()
        ids = torch.from_numpy(np.stack([rng.permutation(E)[:k] for _ in range(T)]).astype(np.int32)).cuda()
        w = torch.from_numpy(rng.random((T, k)).astype(np.float32)).cuda()
        ok = True
        for _ in range(2):
            y_ep = ep.forward(x, ids, w).clone()
            y_full = full.forward(x, ids, w).clone()
            torch.cuda.synchronize()
            ok = ok and torch.equal(y_ep, y_full) and bool(torch.isfinite(y_ep.float()).all())
        q.put((rank, ok))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("mixed_f16,exchange,E", [(False, "push", 6), (True, "push", 7), (False, "nccl", 7), (True, "nccl", 6)])
def test_ep_two_gpus_matches_single_gpu(mixed_f16, exchange, E):
    """exchange="push": counts / dispatch / combine written straight into the peers' symmetric-memory buffers by our
    kernels (mxb_ep_push_counts, mxb_ep_plan, mxb_quant_act_push, down GroupGEMM with C on the peer + c_row);
    exchange="nccl": all_to_all_single with host-side split sizes.  E = 7 splits unevenly (3 + 4 experts)."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, mixed_f16, exchange, E, q)) for r in range(world)]
    for p in procs:
        p.start()
    import queue
    import time
    res, t0 = [], time.time()
    while len(res) < world and time.time() - t0 < 150:
        try:
            res.append(q.get(timeout=2))
        except queue.Empty:
            if any(p.exitcode not in (None, 0) for p in procs):  # a rank died: do not wait for the other one's collective
                break
    if len(res) < world:
        for p in procs:
            if p.is_alive():
                p.kill()
        pytest.fail(f"ranks finished: {res}; exit codes {[p.exitcode for p in procs]}")
    for p in procs:
        p.join(timeout=60)
    assert sorted(res) == [(0, True), (1, True)]
