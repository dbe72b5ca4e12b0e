import os

import torch

from distributed_deep_learning_on_personal_computers_b200 import TrainConfig
from distributed_deep_learning_on_personal_computers_b200.utils import data, metrics
from distributed_deep_learning_on_personal_computers_b200.config import parse_cli


def test_directory_loader_reference_layout(tmp_path):
    tiles = data.synthetic_tiles(7, tile=32)
    data.write_directory(str(tmp_path), tiles)
    train, hold = data.load_directory(str(tmp_path), holdout=3)
    assert len(train) == 4 and len(hold) == 3                     # last-N hold-out (R:672-673)
    assert torch.equal(train.images, tiles.images[:4]) and torch.equal(hold.labels, tiles.labels[4:])
    assert train.images.dtype == torch.uint8 and train.labels.dtype == torch.uint8


def test_sampler_shards_are_disjoint_and_epoch_dependent():
    s0 = data.RankSampler(10, 0, 2, seed=1)
    s1 = data.RankSampler(10, 1, 2, seed=1)
    a, b = s0.indices(0), s1.indices(0)
    assert len(a) == len(b) == 5 and not set(a) & set(b)
    assert s0.indices(1) != a
    compat = data.RankSampler(10, 1, 2, shuffle=False, shard=False)
    assert compat.indices(0) == list(range(10))                    # reference behaviour (Q2)


def test_pinned_loader_yields_all_batches():
    tiles = data.synthetic_tiles(9, tile=32)
    ld = data.PinnedBatchLoader(tiles, 2, torch.device("cpu"))
    order = list(range(9))
    seen = [(img.clone(), lab.clone()) for img, lab in ld.batches(order)]
    assert len(seen) == 4
    assert torch.equal(seen[3][0], tiles.images[6:8])
    assert ld.bytes_per_batch == 2 * 32 * 32 * 4


def test_sample_dump_and_runlog(tmp_path):
    cfg = TrainConfig(out_dir=str(tmp_path), world_size=1)
    log = metrics.RunLog(cfg)
    log.header(123)
    log.epoch(0, 1.5, 0.4, 2.0, 10.0, 1.0)
    text = open(cfg.log_file()).read()
    assert os.path.basename(cfg.log_file()) == "otus_q201.txt"     # otus_{model_bytes}.txt (R:715)
    assert "global_batch=1" in text and "epoch=0 mean_loss=1.500000" in text
    tiles = data.synthetic_tiles(2, tile=32)
    files = metrics.dump_samples(str(tmp_path / "s"), tiles.images, tiles.labels, tiles.labels)
    assert sorted(os.path.basename(f) for f in files) == sorted(
        [f"{k} {i}.png" for i in range(2) for k in ("Model", "Label", "Image")])
    from PIL import Image
    assert Image.open(files[0]).size == (32, 32)


def test_cli_defaults_are_the_reference_constants():
    cfg = parse_cli([])
    assert (cfg.sync_every, cfg.batch_size, cfg.base_channels, cfg.epochs) == (50, 1, 32, 100)
    assert cfg.codec == "q201" and cfg.compress and cfg.lr == 1e-3 and cfg.master_port == 5005
    cfg = parse_cli(["--codec", "float16", "--base-channels", "64", "--compress", "false"])
    assert cfg.codec == "q201" and cfg.base_channels == 64 and cfg.compress is False


def test_step_readback_returns_every_step_one_late():
    rb = metrics.StepReadback(torch.device("cpu"), depth=2)
    assert rb.pop() is None
    got = []
    for i in range(5):
        rb.push(torch.tensor([float(i), 10.0 * i]))
        if rb.in_flight() > 1:
            got.append(rb.pop())
    got += rb.drain()
    assert got == [(float(i), 10.0 * i) for i in range(5)] and rb.in_flight() == 0


def test_train_step_read_matches_train_step(tmp_path):
    from distributed_deep_learning_on_personal_computers_b200.trainer import Trainer
    mk = lambda: Trainer(TrainConfig(base_channels=8, tile=32, batch_size=2, train_tiles=4, holdout_tiles=0,
                                     sync_every=1, world_size=1, codec="fp32", device="cpu", sample_images=0,
                                     out_dir=str(tmp_path), seed=3))
    a, b = mk(), mk()
    tiles = data.synthetic_tiles(4, tile=32, seed=1)
    direct, lagged = [], []
    for i in range(3):
        img, lab = tiles.images[(i % 2) * 2:(i % 2) * 2 + 2], tiles.labels[(i % 2) * 2:(i % 2) * 2 + 2]
        direct.append(tuple(a.train_step(img, lab).tolist()))
        got = b.train_step_read(img, lab)
        assert (got is None) == (i == 0)
        if got is not None:
            lagged.append(got)
    lagged += b.flush_reads()
    assert len(lagged) == 3
    for x, y in zip(direct, lagged):
        assert abs(x[0] - y[0]) <= 1e-4 * abs(x[0]) and x[1] == y[1]
