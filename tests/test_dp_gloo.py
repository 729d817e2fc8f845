"""World-size-2 `gloo` test (CPU) of the data-parallel protocol the product uses,
restated on the oracle: BatchNorm batch sums all-reduced (forward and, through autograd, backward), loss scaled by
1/(B_global*T), gradient buckets SUM-all-reduced, L2 gradients added locally afterwards, the embedding's
un-aggregated clip norm reduced as a sum of squares, then the same per-variable clip + Adam on every rank.
The result must equal ONE process stepping on the global batch."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import keras_semantics as K
from tests.util import syn


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _case(n_enc):
    cfg = K.tiny_config(n_enc=n_enc).no_dropout()
    groups = [syn.small_groups(cfg.N, cfg.R, seed=1 + e, overlap=True) for e in range(n_enc)]
    P = K.init_params(cfg, groups, seed=3)
    rng = np.random.default_rng(11)
    for k in ("att_b1", "att_b2", "inter_b", "out_b", "lstm_b"):
        P[k] = (P[k] + 0.05 * rng.standard_normal(P[k].shape)).astype(np.float32)
    (x, cap, a0, c0), tgt = syn.make_batch(8, cfg.T, cfg.V, cfg.U, cfg.N, seed=5)
    return cfg, groups, P, (x, cap, a0, c0, tgt)


def _shard(batch, rank, world, n_enc):
    """Rank k takes rows [k*B, (k+1)*B) of the global batch; ms2 keeps the half-A / half-B layout per rank."""
    n = batch[0].shape[0]
    if n_enc == 1:
        idx = np.arange(rank * n // world, (rank + 1) * n // world)
    else:
        h, per = n // 2, n // 2 // world
        idx = np.concatenate([np.arange(rank * per, (rank + 1) * per), h + np.arange(rank * per, (rank + 1) * per)])
    return tuple(np.asarray(a)[idx] for a in batch)


def _flat(g):
    return {k: (np.concatenate([a.reshape(-1) for a in v]) if isinstance(v, list) else np.asarray(v).reshape(-1))
            for k, v in g.items()}


def _worker(rank, world, port, n_enc, out):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(1)
    import torch.distributed.nn.functional as dnn
    cfg, groups, P, batch = _case(n_enc)
    local = _shard(batch, rank, world, n_enc)
    K.BN_SUM_REDUCE = lambda t: dnn.all_reduce(t, op=dist.ReduceOp.SUM)
    # CE part only: L2 gradients are added after the exchange (identical on all ranks)
    cfg0 = K.Config(**{**cfg.__dict__, "input_reg": 0.0, "attn_reg": 0.0, "lstm_reg": 0.0, "output_reg": 0.0})
    met, grads, emb_sq, stats, _ = K.loss_and_grads(cfg0, P, groups, local, dtype=torch.float64)
    K.BN_SUM_REDUCE = None
    red = {}
    for k, g in _flat(grads).items():
        t = torch.from_numpy(np.ascontiguousarray(g)) / world        # local mean -> share of the global mean
        dist.all_reduce(t)
        red[k] = t.numpy()
    sq = torch.tensor([emb_sq / world ** 2], dtype=torch.float64); dist.all_reduce(sq)     # un-aggregated slices carry the 1/world scale
    loss = torch.tensor([met["loss"] / world], dtype=torch.float64); dist.all_reduce(loss)
    if rank == 0:
        out.put((red, float(sq), float(loss), {k: np.asarray(v) for k, v in stats.items()}))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("n_enc", [1, 2])
def test_two_rank_protocol_equals_global_batch(n_enc):
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n_enc, out)) for r in range(2)]
    for p in procs:
        p.start()
    red, sq, loss, stats = out.get(timeout=240)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    cfg, groups, P, batch = _case(n_enc)
    cfg0 = K.Config(**{**cfg.__dict__, "input_reg": 0.0, "attn_reg": 0.0, "lstm_reg": 0.0, "output_reg": 0.0})
    met, grads, emb_sq, stats1, _ = K.loss_and_grads(cfg0, P, groups, batch, dtype=torch.float64)
    assert abs(loss - met["loss"]) < 1e-10
    assert abs(sq - emb_sq) <= 1e-9 * max(emb_sq, 1e-30)
    for k, g in _flat(grads).items():
        err = np.abs(red[k] - g).max()
        assert err <= 1e-9 * max(np.abs(g).max(), 1e-12) + 1e-14, (k, err)
    for k in stats1:
        assert np.allclose(stats[k], stats1[k], rtol=1e-10, atol=1e-12), k
