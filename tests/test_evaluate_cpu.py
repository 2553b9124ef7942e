"""Host logic of the msc+flip driver: scores vs the reference's golden vector, image sharding, and the only
collective of the path (histogram all-reduce) on a world_size-2 gloo group.  CPU only."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import weclip_oracle as O
from weclip_b200 import evaluate


def test_scores_match_reference_golden(golden_dir):
    g = np.load(os.path.join(golden_dir, "scores.npz"))
    sc = evaluate.scores(g["hist"])
    np.testing.assert_allclose([sc["pAcc"], sc["mAcc"], sc["miou"]], g["summary"], rtol=1e-12)
    np.testing.assert_allclose([sc["iou"][i] for i in range(21)], g["iou"], rtol=1e-12)


def test_scores_edge_cases():
    h = np.zeros((21, 21))
    h[3, 3] = 10                      # a single class present: mIoU is that class' IoU, absent classes ignored
    assert evaluate.scores(h)["miou"] == 1.0
    h[3, 5] = 10
    assert abs(evaluate.scores(h)["miou"] - 0.5) < 1e-12


def test_shards_partition_the_image_list():
    for n, world in [(0, 2), (1, 8), (1449, 8), (10, 3)]:
        seen = sorted(i for r in range(world) for i in evaluate.shard_indices(n, r, world))
        assert seen == list(range(n))
        sizes = [len(evaluate.shard_indices(n, r, world)) for r in range(world)]
        assert max(sizes) - min(sizes) <= 1


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, gts, preds, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    ev = evaluate.MscFlipEvaluator(model=None, num_classes=21, device="cpu")
    for i in evaluate.shard_indices(len(gts), rank, world):
        for j in range(3):
            ev.hist[j] += torch.from_numpy(O.fast_hist(gts[i], preds[(i + j) % len(preds)], 21))
        ev.n_images += 1
    ev.all_reduce()
    if rank == 0:
        torch.save(ev.hist, out)
    dist.destroy_process_group()


def test_histogram_allreduce_world2(tmp_path):
    rng = np.random.default_rng(0)
    gts = [rng.integers(0, 23, (17, 19)).astype(np.int16) for _ in range(5)]
    preds = [rng.integers(0, 21, (17, 19)).astype(np.int16) for _ in range(5)]
    out = str(tmp_path / "hist.pt")
    mp.spawn(_worker, args=(2, _free_port(), gts, preds, out), nprocs=2, join=True)
    got = torch.load(out).numpy()
    for j in range(3):
        want = sum(O.fast_hist(gts[i], preds[(i + j) % 5], 21) for i in range(5))
        assert np.array_equal(got[j], want)
