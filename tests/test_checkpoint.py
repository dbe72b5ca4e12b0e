
import torch

from distributed_deep_learning_on_personal_computers_b200 import TrainConfig
from distributed_deep_learning_on_personal_computers_b200.trainer import Trainer


def _cfg(d, **kw):
    base = dict(base_channels=8, tile=64, train_tiles=8, holdout_tiles=2, batch_size=2, sync_every=1,
                epochs=1, world_size=1, codec="fp32", device="cpu", out_dir=d, sample_images=0,
                shuffle=False)
    base.update(kw)
    return TrainConfig(**base)


def test_save_resume_gives_identical_next_step(tmp_path):
    a = Trainer(_cfg(str(tmp_path / "a")))
    a.train_epoch()
    a.epoch += 1
    path = a.save(str(tmp_path / "ck.pt"))
    ck = torch.load(path, weights_only=False)
    assert set(ck) >= {"model", "optimizer", "epoch", "step", "config"}
    assert len(ck["model"]) == 166 and all(v.is_contiguous() for v in ck["model"].values())
    b = Trainer(_cfg(str(tmp_path / "b"), resume=path))
    assert b.epoch == 1 and b.global_step == a.global_step
    sa, sb = a.train_epoch(), b.train_epoch()
    assert sa["mean_loss"] == sb["mean_loss"]
    assert torch.equal(a.model.flat["param"], b.model.flat["param"])


def test_optimizer_state_loads_into_torch_adam(tmp_path):
    a = Trainer(_cfg(str(tmp_path / "a")))
    a.train_epoch()
    sd = a.optimizer.state_dict()
    from distributed_deep_learning_on_personal_computers_b200 import UNet
    ref = UNet(out_classes=6, base_channels=8)
    ref.load_state_dict(a.model.portable_state_dict())
    opt = torch.optim.Adam(ref.parameters())
    opt.load_state_dict({"state": sd["state"], "param_groups": sd["param_groups"]})
    # one more identical step on both sides gives the same parameters
    g = torch.randn_like(a.model.flat["param"]) * 1e-3
    a.model.flat["grad"].copy_(g)
    for (n, p), (_, q) in zip(ref.named_parameters(), a.model.named_parameters()):
        p.grad = q.grad.detach().clone().contiguous()
    opt.step()
    a.optimizer.step(grad_scale=1.0)
    for (n, p), (_, q) in zip(ref.named_parameters(), a.model.named_parameters()):
        torch.testing.assert_close(q.detach(), p.detach(), rtol=1e-5, atol=1e-7, msg=n)


def test_sgd_and_bilinear_run(tmp_path):
    t = Trainer(_cfg(str(tmp_path / "c"), optimizer="sgd", lr=0.05, momentum=0.9,
                     up_sample_mode="bilinear", epochs=2))
    l0 = t.train_epoch()["mean_loss"]
    t.epoch += 1
    l1 = t.train_epoch()["mean_loss"]
    assert l1 < l0


def test_torch_engine_matches_native_engine_on_cpu(tmp_path):
    a = Trainer(_cfg(str(tmp_path / "a"), engine="native"))
    b = Trainer(_cfg(str(tmp_path / "b"), engine="torch"))
    sa, sb = a.train_epoch(), b.train_epoch()
    assert abs(sa["mean_loss"] - sb["mean_loss"]) < 1e-3


def test_resume_auto_picks_the_newest_checkpoint_or_starts_fresh(tmp_path):
    d = str(tmp_path / "run")
    fresh = Trainer(_cfg(d, resume="auto"))                  # nothing there yet: a fresh start, not an error
    assert fresh.resumed_from is None and fresh.epoch == 0
    fresh.fit()                                              # one epoch -> <out_dir>/checkpoint.pt
    again = Trainer(_cfg(d, resume="auto", epochs=2))
    assert again.resumed_from.endswith("checkpoint.pt") and again.epoch == 1
    assert torch.equal(again.model.flat["param"], fresh.model.flat["param"])
    import os, time
    time.sleep(0.05)
    again.train_epoch()
    again.save(os.path.join(d, "abort.pt"))                  # what Trainer.check_peers leaves behind on PeerLost
    third = Trainer(_cfg(d, resume="auto", epochs=3))
    assert third.resumed_from.endswith("abort.pt")
    assert torch.equal(third.model.flat["param"], again.model.flat["param"])
