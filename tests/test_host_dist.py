"""CPU: host-side logic -- config keys, structures, image sharding and the optional num_pos
all-reduce over a world_size-2 gloo group."""
import os

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import detectron2_centernet_b200 as P
from detectron2_centernet_b200 import dist as ctdist
from oracle import ctdet_oracle as O
from detectron2_centernet_b200 import synthetic as S

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_config_keys_and_base_inheritance():
    cfg = P.get_cfg(os.path.join(ROOT, "configs", "ctdet_dla_34_1x.yaml"))
    c = cfg["MODEL"]["CENTERNET"]
    assert c["FOCAL_LOSS_ALPHA"] == [1]                      # overridden by the DLA config
    assert c["HM_WEIGHT"] == 1 and c["WH_WEIGHT"] == 0.1 and c["OFF_WEIGHT"] == 1
    assert c["SCORE_THRESH_TEST"] == 0.05 and c["TOPK_CANDIDATES_TEST"] == 100 and c["DOWN_RATIO"] == 4
    assert cfg["TEST"]["DETECTIONS_PER_IMAGE"] == 100
    assert cfg["MODEL"]["BACKBONE"]["NAME"] == "build_dla34_backbone"      # from _BASE_
    assert P.get_cfg()["MODEL"]["CENTERNET"]["FOCAL_LOSS_ALPHA"] == [0.25]  # defaults.py:484
    m = P.CenterNetHotPath(cfg, num_classes=9)
    assert m.num_classes == 9 and m.focal_loss_alpha == [1] and m.max_detections_per_image == 100


def test_instances_and_boxes_contract():
    inst = P.Instances((480, 640))
    inst.pred_boxes = P.Boxes(torch.tensor([[0., 0., 10., 10.], [5., 5., 5., 9.], [-3., 2., 700., 500.]]))
    inst.scores = torch.tensor([0.9, 0.8, 0.7])
    inst.pred_classes = torch.tensor([1, 2, 3], dtype=torch.int32)
    assert len(inst) == 3 and inst.image_size == (480, 640) and inst.has("scores")
    with pytest.raises(AssertionError):
        inst.bad = torch.zeros(2)
    sub = inst[inst.scores > 0.75]
    assert len(sub) == 2 and sub.pred_classes.tolist() == [1, 2]
    assert inst.pred_boxes.nonempty().tolist() == [True, False, True]
    b = inst.pred_boxes.clone()
    b.clip((480, 640))
    assert b.tensor[2].tolist() == [0., 2., 640., 480.]
    b.scale(0.5, 2.0)
    assert b.tensor[0].tolist() == [0., 0., 5., 20.]
    with pytest.raises(AttributeError):
        inst.nothing


def test_shard_range_covers_everything():
    for n in (0, 1, 7, 32, 512):
        for world in (1, 2, 3, 8):
            spans = [ctdist.shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    B, C, H, W = 6, 5, 32, 32
    bl, cl = S.make_batch_objects(77, B, 128, n_max=12, num_classes=C)          # same global batch on every rank
    lo, hi = ctdist.shard_range(B, rank, world)
    gt = O.gen_heatmap_batch(bl[lo:hi], cl[lo:hi], H, W, C)                       # this rank's images only
    t = {k: torch.from_numpy(v) for k, v in gt.items()}
    local = ctdist.local_normalisers(t)
    synced = ctdist.sync_normalisers(t)
    q.put((rank, lo, hi, local.tolist(), synced.tolist()))
    dist.destroy_process_group()


def test_sync_normalisers_world2_gloo():
    world, port = 2, 29711
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    B, C, H, W = 6, 5, 32, 32
    bl, cl = S.make_batch_objects(77, B, 128, n_max=12, num_classes=C)
    gt = O.gen_heatmap_batch(bl, cl, H, W, C)
    want = [float((gt["hm"] == 1).sum()), float(gt["reg_mask"].sum())]
    assert [r[1:3] for r in res] == [(0, 3), (3, 6)]                              # contiguous image blocks
    for r in res:
        assert r[4] == want                                                       # all-reduced = global
    assert [a + b for a, b in zip(res[0][3], res[1][3])] == want                  # locals add up
    assert res[0][3] != want                                                      # and differ from the global
