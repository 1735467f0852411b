"""Trainer shim (npf_gwwaveform_b200/utils/train.py) against the contract of the reference's utils/train.py:
checkpoint file names and formats, history records, ExponentialLR schedule, eval.csv, reload without retraining."""
import json
import os

import numpy as np
import pytest
import torch

from util_models import build_product, chirp_batch

pytestmark = pytest.mark.gpu


class Chirps(torch.utils.data.Dataset):
    """(x_i [N,1], y_i [N,1]) samples, like the reference's GP datasets (utils/data/gaussian_process.py)"""

    def __init__(self, n, N, seed):
        b = chirp_batch(n, 1, N, seed)
        self.x, self.y = b["X_trgt"], b["Y_trgt"]

    def __len__(self):
        return self.x.shape[0]

    def __getitem__(self, i):
        return self.x[i], self.y[i]


def _collate(seed):
    from npf_gwwaveform_b200.utils.datasplit import CntxtTrgtGetter, GetRandomIndcs, cntxt_trgt_collate, get_all_indcs

    return cntxt_trgt_collate(CntxtTrgtGetter(contexts_getter=GetRandomIndcs(a=12, b=12, seed=seed),
                                              targets_getter=get_all_indcs))


def test_gamma():
    from npf_gwwaveform_b200.utils.train import get_exponential_decay_gamma

    g = get_exponential_decay_gamma(100, 50)
    assert abs(g ** 50 - 1 / 100) < 1e-12          # utils/helpers.py:35-46


def test_train_models_contract(tmp_path):
    import npf_gwwaveform_b200 as npf
    from npf_gwwaveform_b200.utils.train import train_models

    torch.manual_seed(0)
    train, valid, test = Chirps(32, 64, 1), Chirps(8, 64, 2), Chirps(10, 64, 3)
    ck = str(tmp_path) + "/"
    kw = dict(iterator_train__collate_fn=_collate(5), iterator_valid__collate_fn=_collate(6))
    trainers = train_models({"chirps": train}, {"CNP": build_product("CNP", 32)}, npf.CNPFLoss,
                            test_datasets={"chirps": test}, valid_datasets={"chirps": valid}, chckpnt_dirname=ck,
                            is_retrain=True, max_epochs=6, batch_size=8, lr=1e-3, decay_lr=10, seed=123, **kw)
    tr = trainers["chirps/CNP/run_0"]
    run = os.path.join(ck, "chirps/CNP/run_0")
    for f in ("params.pt", "optimizer.pt", "history.json", "eval.csv", "model_summary.txt"):
        assert os.path.exists(os.path.join(run, f)), f
    # history: one record per epoch, losses finite and decreasing overall, lr follows ExponentialLR
    assert len(tr.history) >= 1 and len(json.load(open(os.path.join(run, "history.json")))) == len(tr.history)
    h = json.load(open(os.path.join(run, "history.json")))
    assert all(np.isfinite(r["train_loss"]) and np.isfinite(r["valid_loss"]) for r in h)
    # params.pt: reference-format state_dict that loads strictly into a fresh model
    sd = torch.load(os.path.join(run, "params.pt"), map_location="cpu")
    fresh = build_product("CNP", 32)
    fresh.load_state_dict(sd, strict=True)
    # optimizer.pt: loads into torch.optim.Adam over the same parameters
    opt = torch.optim.Adam(fresh.parameters(), lr=1e-3)
    opt.load_state_dict(torch.load(os.path.join(run, "optimizer.pt"), map_location="cpu", weights_only=False))
    st = opt.state[next(iter(fresh.parameters()))]
    assert float(st["step"]) > 0 and st["exp_avg"].abs().sum() > 0
    # eval.csv: one log-likelihood per test task
    ll = np.loadtxt(os.path.join(run, "eval.csv"), delimiter=",")
    assert ll.shape == (10,) and np.isfinite(ll).all()
    assert abs(tr.test_loglike_ - ll.mean()) < 1e-4 * abs(ll.mean())
    # training helped on the training distribution
    assert tr.history[-1]["train_loss"] < tr.history[0]["train_loss"]
    # second call without retraining loads the checkpoint and reproduces the evaluation (is_reeval forces a rerun)
    again = train_models({"chirps": train}, {"CNP": build_product("CNP", 32)}, npf.CNPFLoss,
                         test_datasets={"chirps": test}, chckpnt_dirname=ck, is_retrain=False, is_reeval=True,
                         batch_size=8, **kw)["chirps/CNP/run_0"]
    ll2 = np.loadtxt(os.path.join(run, "eval.csv"), delimiter=",")
    np.testing.assert_allclose(ll2, ll, rtol=1e-4, atol=1e-4)
    assert len(again.history) == len(tr.history)


def test_lr_schedule_and_graph_replay():
    import npf_gwwaveform_b200 as npf
    from npf_gwwaveform_b200.utils.train import NeuralNetShim, get_exponential_decay_gamma

    torch.manual_seed(0)
    net = NeuralNetShim(build_product("AttnCNP", 128), npf.CNPFLoss, lr=1e-3, max_epochs=4, batch_size=8, decay_lr=16,
                        iterator_train__collate_fn=_collate(7))
    net.fit(Chirps(32, 96, 4))
    g = get_exponential_decay_gamma(16, 4)
    assert [round(r["lr"] / 1e-3, 6) for r in net.history] == [round(g ** e, 6) for e in range(4)]
    assert getattr(net, "_graphs", {}), "steps with a repeating batch shape replay a CUDA graph"
    assert net.history[-1]["train_loss"] < net.history[0]["train_loss"]
