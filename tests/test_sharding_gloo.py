"""The N>1 path on CPU: two processes over torch.distributed (gloo). Open-loop ME shards by mini-GOP with no
data-path collective: each rank works out, with the product's host logic
(svtme_shard_owner, picture_plan), which pictures it owns and which reference pictures it must upload; the
ranks' shards must tile the stream, and the per-picture results (here from the oracle, the checker) must not
depend on the rank that produced them. The only collectives are the ones bench.py uses: a barrier and a MAX."""
import ctypes as C
import os
import socket
import zlib

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

WORLD = 2
W, H, HIER, N_MGOPS, PRESET = 128, 64, 3, 4, 8
MGOP, N_FRAMES = 1 << HIER, N_MGOPS * (1 << HIER) + 1


def shard_plan(rank, world):
    """-> (owned pocs in coding order, pocs to upload)"""
    import svtme_b200 as me
    from svt_av1_psy_b200 import gop
    owner = me.host_lib().svtme_shard_owner
    owner.argtypes = [C.c_int64, C.c_int32, C.c_int32, C.c_int32]
    mrp = gop.mrp_for_preset(PRESET)
    owned, upload = [], set()
    for m in range(N_MGOPS):
        first = m * MGOP
        if owner(first + 1, MGOP, 0, world) != rank:
            continue
        for poc in gop.coding_order(first, HIER, N_FRAMES):
            assert owner(poc, MGOP, 0, world) == rank  # a mini-GOP is never split
            plan = gop.picture_plan(poc, HIER, mrp["l0"], mrp["l1"], N_FRAMES)
            owned.append(poc)
            upload.update([poc] + plan["l0"] + plan["l1"])
    return owned, sorted(upload)


def picture_digest(poc):
    """oracle ME of one picture of the synthetic stream -> crc32 of its MeSbResults arrays"""
    import sys
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from svt_av1_psy_b200 import gop
    from svt_av1_psy_b200.synth import PanningNoise
    from util import orc_me, params_from_encoder_inputs
    clip = PanningNoise(W, H, seed=5, vx=2, vy=1, max_frames=N_FRAMES)
    mrp = gop.mrp_for_preset(PRESET)
    plan = gop.picture_plan(poc, HIER, mrp["l0"], mrp["l1"], N_FRAMES)
    refs = [[clip.frame(p) for p in plan["l0"]], [clip.frame(p) for p in plan["l1"]]]
    prm = params_from_encoder_inputs(clip.frame(poc), refs, [plan["l0"], plan["l1"]], poc, PRESET, temporal_layer=plan["tl"],
                                     hier_levels=3, is_ref=plan["is_ref"], mrp=mrp)
    res, _, _ = orc_me(clip.frame(poc), refs, prm, debug=False)
    crc = 0
    for f in ("total_me_candidate_index", "me_mv_array", "me_candidate_array", "me_64x64_distortion"):
        crc = zlib.crc32(np.ascontiguousarray(getattr(res, f)).tobytes(), crc)
    return crc


def worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    owned, upload = shard_plan(rank, world)
    digests = {poc: picture_digest(poc) for poc in owned[:3]}  # a few pictures per rank keep the test short
    gathered = [None] * world
    dist.all_gather_object(gathered, (owned, upload, digests))
    t = torch.tensor([float(rank + 1)], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)  # the max-over-ranks reduction of bench.py
    dist.barrier()
    if rank == 0:
        torch.save({"gathered": gathered, "max": float(t.item())}, os.path.join(out_dir, "out.pt"))
    dist.destroy_process_group()


def free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_two_rank_sharding(tmp_path, orc):
    from oracle import pyoracle as po
    po.oracle_lib()
    mp.spawn(worker, args=(WORLD, free_port(), str(tmp_path)), nprocs=WORLD, join=True)
    out = torch.load(os.path.join(str(tmp_path), "out.pt"), weights_only=False)
    assert out["max"] == float(WORLD)
    (own0, up0, dig0), (own1, up1, dig1) = out["gathered"]
    all_pocs = sorted(own0 + own1)
    assert all_pocs == list(range(1, N_MGOPS * MGOP + 1)) and not set(own0) & set(own1)  # the shards tile the stream
    assert len(own0) == len(own1)  # weak scaling: equal work per rank
    # a rank uploads its own pictures plus the few neighbouring base pictures it references, nothing else
    assert set(own0) <= set(up0) and set(own1) <= set(up1)
    assert len(set(up0) - set(own0)) <= 2 * N_MGOPS and len(set(up1) - set(own1)) <= 2 * N_MGOPS
    # results do not depend on the rank: recompute here what each rank reported
    for dig in (dig0, dig1):
        assert len(dig) == 3
        for poc, crc in dig.items():
            assert picture_digest(poc) == crc, poc
