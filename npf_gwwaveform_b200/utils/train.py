"""Trainer shim with the contract of the reference's training driver.

The reference trains through skorch 0.8 (``utils/train.py:34-349``: ``NeuralNet(model, criterion, ...).fit(dataset)``
with Adam lr 1e-3, optional ``ExponentialLR``, a ``Checkpoint`` callback writing ``params.pt`` / ``optimizer.pt`` /
``history.json`` into ``<chckpnt_dirname><data>/<model>/run_<k>/``, ``model_summary.txt``, and ``eval.csv`` with the
per-task test log-likelihoods from ``utils/evaluate.py:9-28``).  skorch is not installable here (no network) and its
Python epoch loop is exactly the launch overhead the B200 path removes, so this module provides the SAME surface on top
of ``FlatAdamTrainer``:

* ``NeuralNetShim`` -- ``module_``, ``criterion_``, ``history``, ``initialize()``, ``fit()``, ``train_step()``,
  ``validation_step()``, ``get_loss()`` (sets ``criterion_.train(training)`` first, utils/train.py:342-349),
  ``save_params()`` / ``load_params()`` with the reference's file names and formats (``params.pt`` = ``state_dict``,
  ``optimizer.pt`` = a ``torch.optim.Adam.state_dict()`` layout, ``history.json`` = list of epoch records);
* ``train_models`` -- same arguments and return value (``{"<data>/<model>/run_<k>": trainer}``);
* ``eval_loglike``, ``get_exponential_decay_gamma`` (utils/helpers.py:35-46).

Batches are split into context / target sets on the device (``utils/datasplit.py``); forward + loss + backward of a step
run as one CUDA-graph replay when the batch shapes repeat.
"""
import json
import os
import time
from copy import deepcopy

import numpy as np
import torch

from ..training import FlatAdamTrainer
from .datasplit import CntxtTrgtGetter, GetRandomIndcs, cntxt_trgt_collate, get_all_indcs

__all__ = ["NeuralNetShim", "train_models", "eval_loglike", "get_exponential_decay_gamma"]

EVAL_FILENAME = "eval.csv"
MOD_SUMM_FILENAME = "model_summary.txt"


def get_exponential_decay_gamma(scheduling_factor, max_epochs):
    """gamma such that the learning rate has decayed by ``scheduling_factor`` after ``max_epochs``
    (utils/helpers.py:35-46)."""
    return (1 / scheduling_factor) ** (1 / max_epochs)


def _default_collate():
    return cntxt_trgt_collate(CntxtTrgtGetter(contexts_getter=GetRandomIndcs(a=0.0, b=50), targets_getter=get_all_indcs))


class NeuralNetShim:
    """What ``utils/train.py`` uses of ``skorch.NeuralNet``, driven by the flat-buffer Adam trainer."""

    def __init__(self, module, criterion, lr=1e-3, max_epochs=100, batch_size=16, device="cuda",
                 iterator_train__collate_fn=None, iterator_valid__collate_fn=None, iterator_train__shuffle=True,
                 iterator_valid__batch_size=None, decay_lr=None, seed=None, use_cuda_graph=True, verbose=0, **kwargs):
        self.module = module
        self.criterion = criterion
        self.lr, self.max_epochs, self.batch_size, self.device = lr, max_epochs, batch_size, torch.device(device)
        self.collate_train = iterator_train__collate_fn
        self.collate_valid = iterator_valid__collate_fn or iterator_train__collate_fn
        self.shuffle = iterator_train__shuffle
        self.valid_batch_size = iterator_valid__batch_size or 2 * batch_size
        self.decay_lr, self.seed, self.use_cuda_graph, self.verbose = decay_lr, seed, use_cuda_graph, verbose
        self.history = []
        # captures kept at once (the default collate draws 51 context sizes: every one gets its graph, they share one
        # memory pool; least recently used is dropped beyond this)
        self.max_graphs = 64
        self.initialized_ = False

    # ------------------------------------------------------------------ skorch surface
    def initialize(self):
        if self.initialized_:
            return self
        self.module_ = (self.module() if isinstance(self.module, type) or not isinstance(self.module, torch.nn.Module)
                        else self.module).to(self.device)
        self.criterion_ = self.criterion() if isinstance(self.criterion, type) else self.criterion
        if self.collate_train is None:
            self.collate_train = self.collate_valid = _default_collate()
        self.trainer_ = FlatAdamTrainer(self.module_, self._train_criterion, lr=self.lr)
        self._graph_shapes = None
        self.initialized_ = True
        return self

    def _train_criterion(self, y_pred, y_true):
        return self.get_loss(y_pred, y_true, training=True)

    def get_loss(self, y_pred, y_true, X=None, training=False):
        """criterion_.train(training) first, so the loss picks its train / eval estimator (utils/train.py:342-349)"""
        if isinstance(self.criterion_, torch.nn.Module):
            self.criterion_.train(training)
        return self.criterion_(y_pred, y_true.to(self.device))

    def get_iterator(self, dataset, training=False):
        return torch.utils.data.DataLoader(
            dataset, batch_size=self.batch_size if training else self.valid_batch_size,
            shuffle=self.shuffle if training else False, collate_fn=self.collate_train if training else self.collate_valid)

    def train_step(self, Xi, yi=None):
        """one optimisation step on a collated batch (dict with X_cntxt, Y_cntxt, X_trgt, Y_trgt)"""
        self.module_.train()
        shapes = tuple((k, tuple(v.shape)) for k, v in sorted(Xi.items()))
        if self.use_cuda_graph and self.trainer_._order is not None and self.trainer_.step_count >= 2:
            if self._graph_shapes != shapes:       # the context size changes from batch to batch: one capture per shape
                self._graphs = getattr(self, "_graphs", {})
                if shapes not in self._graphs:
                    if len(self._graphs) >= self.max_graphs:            # least recently used capture makes room
                        self._graphs.pop(next(iter(self._graphs)))
                    self.trainer_.capture(Xi)
                    self._graphs[shapes] = self.trainer_.export_capture()
                else:
                    self._graphs[shapes] = self._graphs.pop(shapes)     # move to the most-recent end
                    self.trainer_.import_capture(self._graphs[shapes])
                self._graph_shapes = shapes
            if self._graph_shapes == shapes:
                return {"loss": self.trainer_.step_graphed(Xi).clone()}     # the graph's loss tensor is reused
        return {"loss": self.trainer_.step(Xi)}

    def validation_step(self, Xi, yi=None):
        self.module_.eval()
        with torch.no_grad():
            y_pred = self.module_(**Xi)
            loss = self.get_loss(y_pred, Xi["Y_trgt"] if yi is None else yi, training=False)
        return {"loss": loss, "y_pred": y_pred}

    def fit(self, dataset_train, dataset_valid=None, epochs=None):
        self.initialize()
        if self.seed is not None:
            torch.manual_seed(self.seed)
            np.random.seed(self.seed)
        gamma = get_exponential_decay_gamma(self.decay_lr, self.max_epochs) if self.decay_lr is not None else 1.0
        best = {"train": float("inf"), "valid": float("inf")}
        for h in self.history:
            best["train"] = min(best["train"], h["train_loss"])
            if h.get("valid_loss") is not None:
                best["valid"] = min(best["valid"], h["valid_loss"])
        for _ in range(epochs or self.max_epochs):
            epoch = len(self.history) + 1
            self.trainer_.lr = self.lr * gamma ** (epoch - 1)           # ExponentialLR stepped once per epoch
            t0 = time.time()
            losses, sizes = [], []
            for Xi, yi in self.get_iterator(dataset_train, training=True):
                losses.append(self.train_step(Xi, yi)["loss"])
                sizes.append(Xi["X_trgt"].shape[0])
            rec = {"epoch": epoch, "train_loss": _weighted_mean(losses, sizes), "valid_loss": None,
                   "lr": self.trainer_.lr, "dur": None}
            if dataset_valid is not None:
                vl, vs = [], []
                for Xi, yi in self.get_iterator(dataset_valid, training=False):
                    vl.append(self.validation_step(Xi, yi)["loss"])
                    vs.append(Xi["X_trgt"].shape[0])
                rec["valid_loss"] = _weighted_mean(vl, vs)
            rec["train_loss_best"] = rec["train_loss"] < best["train"]
            best["train"] = min(best["train"], rec["train_loss"])
            rec["valid_loss_best"] = rec["valid_loss"] is not None and rec["valid_loss"] < best["valid"]
            if rec["valid_loss"] is not None:
                best["valid"] = min(best["valid"], rec["valid_loss"])
            rec["dur"] = time.time() - t0
            self.history.append(rec)
            if self.verbose:
                print(rec, flush=True)
            if getattr(self, "checkpoint_dir_", None) and (rec["valid_loss_best"] or dataset_valid is None):
                self.save_params(self.checkpoint_dir_)       # Checkpoint(monitor="valid_loss_best" | None)
        return self

    # ------------------------------------------------------------------ checkpoints (skorch Checkpoint file names)
    def optimizer_state_dict(self):
        """the flat Adam buffers in ``torch.optim.Adam.state_dict()`` layout (parameter order of the module)"""
        tr = self.trainer_
        state = {}
        for i, p in enumerate(tr.params):
            o, n = tr.offsets[i]
            state[i] = {"step": torch.tensor(float(tr.step_count)), "exp_avg": tr.exp_avg[o:o + n].view_as(p).clone(),
                        "exp_avg_sq": tr.exp_avg_sq[o:o + n].view_as(p).clone()}
        group = {"lr": tr.lr, "betas": tuple(tr.betas), "eps": tr.eps, "weight_decay": 0, "amsgrad": False,
                 "maximize": False, "foreach": None, "capturable": False, "differentiable": False, "fused": None,
                 "params": list(range(len(tr.params)))}
        return {"state": state, "param_groups": [group]}

    def load_optimizer_state_dict(self, sd):
        tr = self.trainer_
        for i, p in enumerate(tr.params):
            if i in sd["state"]:
                o, n = tr.offsets[i]
                tr.exp_avg[o:o + n].copy_(sd["state"][i]["exp_avg"].reshape(-1))
                tr.exp_avg_sq[o:o + n].copy_(sd["state"][i]["exp_avg_sq"].reshape(-1))
                tr.step_count = int(float(sd["state"][i]["step"]))
        if sd.get("param_groups"):
            tr.lr = sd["param_groups"][0]["lr"]

    def save_params(self, dirname):
        os.makedirs(dirname, exist_ok=True)
        torch.save({k: v.detach().cpu() for k, v in self.module_.state_dict().items()}, os.path.join(dirname, "params.pt"))
        osd = self.optimizer_state_dict()
        for st in osd["state"].values():
            st["exp_avg"], st["exp_avg_sq"] = st["exp_avg"].cpu(), st["exp_avg_sq"].cpu()
        torch.save(osd, os.path.join(dirname, "optimizer.pt"))
        with open(os.path.join(dirname, "history.json"), "w") as f:
            json.dump(self.history, f, indent=1)

    def load_params(self, dirname=None, checkpoint=None):
        dirname = dirname or checkpoint
        self.initialize()
        sd = torch.load(os.path.join(dirname, "params.pt"), map_location="cpu")
        with torch.no_grad():                       # in place: the parameters are views of the flat buffer
            own = self.module_.state_dict()
            assert set(own) == set(sd), "params.pt does not match the module (strict)"
            for k, v in sd.items():
                own[k].copy_(v)
        opt = os.path.join(dirname, "optimizer.pt")
        if os.path.exists(opt):
            self.load_optimizer_state_dict(torch.load(opt, map_location="cpu", weights_only=False))
        hist = os.path.join(dirname, "history.json")
        if os.path.exists(hist):
            with open(hist) as f:
                self.history = json.load(f)
        return self


def _weighted_mean(losses, sizes):
    if not losses:
        return float("nan")
    vals = torch.stack([l.detach().float().reshape(()) for l in losses]).cpu().numpy()      # one sync per epoch
    return float(np.average(vals, weights=np.asarray(sizes, dtype=np.float64)))


def eval_loglike(trainer, dataset, seed=123, **kwargs):
    """per-task log-likelihood (= minus the eval-mode loss with reduction=None), in dataset order
    (utils/evaluate.py:9-28)"""
    torch.manual_seed(seed)
    np.random.seed(seed)
    import random

    random.seed(seed)
    getter = getattr(getattr(trainer.collate_valid, "get_cntxt_trgt", None), "contexts_getter", None)
    if hasattr(getter, "reseed"):
        getter.reseed(seed)                       # same context sets on every evaluation
    old = trainer.criterion_.reduction
    trainer.criterion_.reduction = None
    out = []
    try:
        for Xi, yi in trainer.get_iterator(dataset, training=False):
            out.append(-trainer.validation_step(Xi, yi)["loss"])
    finally:
        trainer.criterion_.reduction = old
    return torch.cat(out, dim=0).detach().cpu().numpy()


def train_models(datasets, models, criterion, test_datasets=dict(), valid_datasets=dict(), chckpnt_dirname=None,
                 is_continue_train=False, is_retrain=False, runs=1, starting_run=0, device=None, max_epochs=100,
                 batch_size=16, lr=1e-3, decay_lr=None, is_reeval=False, seed=None, datasets_kwargs=dict(),
                 models_kwargs=dict(), **kwargs):
    """Train or load ``models`` on ``datasets`` (arguments and return value of utils/train.py:34-310).

    ``models`` maps a name to a model factory (or ``{data_name: {model_name: factory}}``); every run is checkpointed
    in ``<chckpnt_dirname><data>/<model>/run_<k>/`` with the reference's file names; test log-likelihoods go to
    ``eval.csv``.  Returns ``{"<data>/<model>/run_<k>": trainer}`` with modules moved back to the CPU."""
    trainers = {}
    device = device or "cuda"
    for data_name, data_train in datasets.items():
        current_models = models[data_name] if isinstance(list(models.values())[0], dict) else models
        data_test, data_valid = test_datasets.get(data_name), valid_datasets.get(data_name)
        for model_name, model in current_models.items():
            for run in range(starting_run, starting_run + runs):
                suffix = f"{data_name}/{model_name}/run_{run}"
                print(f"\n--- {'Training' if is_retrain else 'Loading'} {suffix} ---\n", flush=True)
                ckpt = None if chckpnt_dirname is None else chckpnt_dirname + suffix
                kw = dict(kwargs)
                kw.update(datasets_kwargs.get(data_name, {}))
                kw.update(models_kwargs.get(model_name, {}))
                trainer = NeuralNetShim(deepcopy(model) if isinstance(model, torch.nn.Module) else model, criterion,
                                        lr=lr, max_epochs=max_epochs, batch_size=batch_size, device=device,
                                        decay_lr=decay_lr, seed=None if seed is None else seed + run, **kw)
                trainer.initialize()
                trainer.checkpoint_dir_ = ckpt
                if is_continue_train:
                    assert ckpt is not None
                    trainer.load_params(ckpt)
                if is_retrain:
                    trainer.fit(data_train, data_valid)
                    if ckpt is not None:
                        with open(os.path.join(ckpt, MOD_SUMM_FILENAME), "w") as f:
                            f.write(str(trainer.module_))
                if ckpt is not None:
                    trainer.load_params(ckpt)
                test_loglike = None
                if data_test is not None:
                    eval_file = None if ckpt is None else os.path.join(ckpt, EVAL_FILENAME)
                    if eval_file and os.path.exists(eval_file) and not (is_retrain or is_reeval):
                        ll = np.loadtxt(eval_file, delimiter=",")
                    else:
                        ll = eval_loglike(trainer, data_test)
                        if eval_file:
                            np.savetxt(eval_file, ll, delimiter=",")
                    test_loglike = float(ll.mean(axis=0))
                trainer.test_loglike_ = test_loglike
                print(suffix, "| epochs:", len(trainer.history), "| train loss:",
                      None if not trainer.history else round(trainer.history[-1]["train_loss"], 4),
                      "| test log likelihood:", None if test_loglike is None else round(test_loglike, 4), flush=True)
                # the reference hands back CPU modules (utils/train.py:335-338).  Moving the module re-creates every
                # parameter's storage, so the flat Adam buffers and the captured graphs no longer alias the weights: drop
                # them; a later fit() / initialize() rebuilds the trainer on whatever device the module is moved to
                trainer.module_.cpu()
                if getattr(trainer, "trainer_", None) is not None:
                    trainer.trainer_.close()
                trainer.trainer_ = None
                trainer._graphs, trainer._graph_shapes = {}, None
                trainer.initialized_ = False
                trainer.module = trainer.module_
                torch.cuda.empty_cache()
                trainers[suffix] = trainer
    return trainers
