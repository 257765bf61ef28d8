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


def _worker(rank, world, port, mixed_f16, exchange, E, q, hadamard=False, shape=None):
    import torch.distributed as dist
    from mxmoe_b200 import ops
    from mxmoe_b200.config import qmodel_cfg_from_dict
    from mxmoe_b200.ep import EPMoELayer
    from mxmoe_b200.layer import MoELayer
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        if shape == "c2":
            # BASELINE configs[1] dimensions and qconfig (what SCALE measures), fewer tokens per rank: Qwen1.5-MoE layer 11,
            # 60 experts top-4 + the fused shared expert, mixed w4a4 + w8a8 at 5.0 bits, online Hadamard on the down input
            # (I = 1408 = 176 x 8, shared 5632 = 176 x 32), the bench's own weights (seeded per expert) and routing profile
            import bench
            from mxmoe_b200 import config as C
            sh = C.MODEL_SHAPES["qwen2_moe"]
            E, k, Hd, I, ns, T = sh["E"], sh["topk"], sh["H"], sh["I"], sh["ns"], 1024
            ql = bench.qlayer_for("qwen2_moe", 11, "mixed5")
            weights, wscale = None, 0.05
            ids_h, w_h = bench.routing("qwen2_moe", T, seed=rank, device=None)
        else:
            k, Hd, I, ns, T = 2, 256, (1408 if hadamard else 128), 2, 300  # T*k > 512 rows: both layers take the tcgen05 path for the w4a16 expert
            cfgs = [(_lin(8, 8), _lin(4, 4)), (_lin(4, 4), _lin(8, 8)), (_lin(4, 4), _lin(4, 4))]
            if mixed_f16:
                cfgs.append((_lin(16, 16), _lin(4, 16, 128, False)))
            ex = {}
            for e in range(E + 1):
                gu, dn = cfgs[e % len(cfgs)]
                ex[str(e)] = {"gate": gu, "up": dict(gu), "down": dn}
            ql = qmodel_cfg_from_dict({"0": {"experts": ex}}).layers["0"]
            g = torch.Generator().manual_seed(5)
            weights, wscale = {}, 1.0
            for e in range(E + 1):
                Ie = I * ns if e == E else I
                weights[e] = (((torch.rand(2 * Ie, Hd, generator=g) * 2 - 1) * 0.12).half(),
                              ((torch.rand(Hd, Ie, generator=g) * 2 - 1) * (0.18 if I == 128 else 0.05)).half())
            ids_h = w_h = None
        h = ops.Handle(rank)
        ep = EPMoELayer(h, E, k, Hd, I, ns, ql, max_tokens=T, hadamard_down=hadamard, weights=weights, exchange=exchange,
                        weight_scale=wscale,
                        expert_order=[(3 * i + 1) % E for i in range(E)] if E == 7 else None)  # E = 7: a scattered placement
        full = MoELayer(h, E, k, Hd, I, ns, ql, max_tokens=T, hadamard_down=hadamard, weights=weights, weight_scale=wscale)
        rng = np.random.default_rng(10 + rank)
        x = torch.from_numpy((rng.random((T, Hd)) * 2 - 1).astype(np.float16)).cuda()
        if ids_h is None:
            ids = torch.from_numpy(np.stack([rng.permutation(E)[:k] for _ in range(T)]).astype(np.int32)).cuda()
            w = torch.from_numpy(rng.random((T, k)).astype(np.float32)).cuda()
        else:
            ids, w = ids_h.cuda(), w_h.cuda()
        ok = True
        for _ in range(2):
            y_ep = ep.forward(x, ids, w).clone()
            y_full = full.forward(x, ids, w).clone()
            torch.cuda.synchronize()
            ok = ok and torch.equal(y_ep, y_full) and bool(torch.isfinite(y_ep.float()).all()) and bool(y_ep.float().abs().max() > 0)
        q.put((rank, ok))
    finally:
        dist.destroy_process_group()


def _run_world(world, args, timeout=240):
    import queue
    import time
    import torch.multiprocessing as mp
    port = _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port) + tuple(args[:3]) + (q,) + tuple(args[3:])) for r in range(world)]
    for p in procs:
        p.start()
    res, t0 = [], time.time()
    while len(res) < world and time.time() - t0 < timeout:
        try:
            res.append(q.get(timeout=2))
        except queue.Empty:
            if any(p.exitcode not in (None, 0) for p in procs):  # a rank died: do not wait for the others' collective
                break
    if len(res) < world:
        for p in procs:
            if p.is_alive():
                p.kill()
        pytest.fail(f"ranks finished: {res}; exit codes {[p.exitcode for p in procs]}")
    for p in procs:
        p.join(timeout=60)
    assert sorted(res) == [(r, True) for r in range(world)]


@pytest.mark.parametrize("mixed_f16,exchange,E", [(False, "push", 6), (True, "push", 7), (False, "nccl", 7), (True, "nccl", 6)])
def test_ep_two_gpus_matches_single_gpu(mixed_f16, exchange, E):
    """exchange="push": counts / dispatch / combine written straight into the peers' symmetric-memory buffers by our
    kernels (mxb_ep_push_counts, mxb_ep_plan, mxb_quant_act_push, down GroupGEMM with C on the peer + c_row);
    exchange="nccl": all_to_all_single with host-side split sizes.  E = 7 splits unevenly (3 + 4 experts)."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    _run_world(2, (mixed_f16, exchange, E))


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("exchange,E,mixed_f16", [("push", 9, False), ("push", 11, True), ("nccl", 9, False)])
def test_ep_with_online_hadamard_matches_single_gpu(world, exchange, E, mixed_f16):
    """The exchange with the online Hadamard on the down input (I = 1408 = 176 x 8; shared expert 2816 = 176 x 16) at
    W = 2 / 4 / 8: at W = 8 seven eighths of the rows are remote, every rank returns to 8 sources (back_row / c_row per
    (source, expert) segment) and E = 9 / 11 leave ranks with one or two experts (uneven split)."""
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    _run_world(world, (mixed_f16, exchange, E, True))


@pytest.mark.parametrize("world", [2, 4, 8])
def test_ep_at_c2_dimensions_matches_single_gpu(world):
    """What SCALE runs, checked: H = 2048, I = 1408, 60 experts top-4 + shared expert (N = 11264), mixed w4a4 + w8a8,
    +Hadamard, push exchange, the bench's weights and routing; 1024 tokens per rank.  Every rank's output must be
    bit-identical to the single-GPU layer over the same tokens (that layer against the oracle: test_gpu_layer_configs)."""
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    _run_world(world, (False, "push", 60, True, "c2"), timeout=420)
