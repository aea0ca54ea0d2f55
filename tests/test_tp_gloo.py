"""World-size-2 `gloo` test (CPU) of the tensor-parallel sharding of the decode hot path.

One decode layer (qkv -> paged attention over the rank's kv heads -> o_proj + all-reduce -> gate_up ->
silu*up -> down_proj + all-reduce) is computed by two ranks on their shards with the oracle's attention and
plain fp32 matmuls, and must reproduce the unsharded layer.  This is the host logic `DecodeRunner` relies on
at TP > 1: head-wise qkv / cache split, merged-column gate|up split, row-parallel partial sums.
"""
import os
import socket
import sys
from pathlib import Path

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

from oracle import oracle as O  # noqa: E402
from vllm_b200 import tp_shard as T  # noqa: E402

LAY = dict(num_q_heads=8, num_kv_heads=2, head_size=16, intermediate=96)
HIDDEN, BS, LENS = 64, 16, [37, 5, 64]


def _case():
    g = torch.Generator().manual_seed(3)
    Hq, Hkv, D, I = LAY["num_q_heads"], LAY["num_kv_heads"], LAY["head_size"], LAY["intermediate"]
    S = len(LENS)
    maxb = (max(LENS) + BS - 1) // BS
    nb = S * maxb + 2
    w = {
        "qkv": torch.randn((Hq + 2 * Hkv) * D, HIDDEN, generator=g) * 0.1,
        "o": torch.randn(HIDDEN, Hq * D, generator=g) * 0.1,
        "gate_up": torch.randn(2 * I, HIDDEN, generator=g) * 0.1,
        "down": torch.randn(HIDDEN, I, generator=g) * 0.1,
    }
    x = torch.randn(S, HIDDEN, generator=g)
    kv = torch.randn(2, nb, BS, Hkv, D, generator=g)
    bt = (torch.randperm(nb - 1, generator=g)[: S * maxb] + 1).reshape(S, maxb).to(torch.int32)
    sl = torch.tensor(LENS, dtype=torch.int32)
    return w, x, kv, bt, sl


def _layer(w, x, kv, bt, sl, Hq, Hkv, D, reduce):
    """One decode layer on (possibly sharded) weights; `reduce` sums row-parallel partial outputs."""
    qkv = x @ w["qkv"].t()
    q = qkv[:, : Hq * D].reshape(-1, Hq, D)
    attn = O.paged_attention_decode(q, kv[0], kv[1], bt, sl, D ** -0.5)
    h = reduce(attn.reshape(x.shape[0], Hq * D).float() @ w["o"].t())
    gu = h @ w["gate_up"].t()
    inter = gu.shape[1] // 2
    act = torch.nn.functional.silu(gu[:, :inter]) * gu[:, inter:]
    return reduce(act @ w["down"].t())


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        lay = T.TPLayout(tp_size=world, **LAY)
        w, x, kv, bt, sl = _case()
        ws = {
            "qkv": T.shard_qkv(w["qkv"], lay, rank),
            "o": T.shard_row_parallel(w["o"], world, rank),
            "gate_up": T.shard_gate_up(w["gate_up"], lay, rank),
            "down": T.shard_row_parallel(w["down"], world, rank),
        }
        kvs = T.shard_kv_cache(kv, lay, rank)

        def all_reduce(t):
            t = t.contiguous()
            dist.all_reduce(t)
            return t

        out = _layer(ws, x, kvs, bt, sl, lay.q_heads_per_rank, lay.kv_heads_per_rank, lay.head_size, all_reduce)
        # every rank must hold the same reduced result; rank 0 reports it
        gathered = [torch.empty_like(out) for _ in range(world)]
        dist.all_gather(gathered, out)
        same = all(torch.equal(gathered[0], t) for t in gathered)
        if rank == 0:
            q.put((out, same))
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_tp2_layer_matches_unsharded():
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    out, same = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert same, "ranks disagree after the all-reduce"
    w, x, kv, bt, sl = _case()
    ref = _layer(w, x, kv, bt, sl, LAY["num_q_heads"], LAY["num_kv_heads"], LAY["head_size"], lambda t: t)
    torch.testing.assert_close(out, ref, atol=1e-4, rtol=1e-4)


def test_shard_shapes_and_errors():
    lay = T.TPLayout(tp_size=2, **LAY)
    w, _, kv, _, _ = _case()
    D = LAY["head_size"]
    assert T.shard_qkv(w["qkv"], lay, 1).shape == ((4 + 2 * 1) * D, HIDDEN)
    assert T.shard_gate_up(w["gate_up"], lay, 0).shape == (96, HIDDEN)
    assert T.shard_row_parallel(w["down"], 2, 1).shape == (HIDDEN, 48)
    assert T.shard_kv_cache(kv, lay, 1).shape[3] == 1
    # the two ranks' q heads partition the full set, in order
    assert list(lay.q_head_range(0)) + list(lay.q_head_range(1)) == list(range(8))
    # vocab rows: last shard may be short
    v = torch.zeros(11, 4)
    assert T.shard_vocab_rows(v, 2, 0).shape[0] == 6 and T.shard_vocab_rows(v, 2, 1).shape[0] == 5
    with pytest.raises(ValueError):
        T.TPLayout(tp_size=4, **LAY)  # 2 kv heads cannot be split 4 ways (replication is not built)
    with pytest.raises(ValueError):
        T.shard_qkv(w["qkv"][:-1], lay, 0)
