"""`train_classification.py` on the B200 engine: the fine-tune loop around `engine.ClassifierEngine`.

What is mirrored from the reference (classification/train_classification.py):
  * the command line, flag for flag with the same dests and defaults            (get_args, :7706-7982)
  * model construction from --pretraining / --checkpoint / --frozen              (build, :5643-5655)
  * fine-tune regimes none / full / head+1 / head+2                              (finetune.py:29-92)
  * the loss: 2 classes -> BCEWithLogits(logit1-logit0, pos_weight=neg/pos), else weighted CE   (:6086-6104, :3347-3374)
  * AdamW over every trainable parameter with ONE weight decay                   (:5751-5768)
  * per-epoch cosine / plateau schedulers, early stopping, monitor resolution    (:3943-3971, :3297-3345, :3903-3915)
  * evaluation: probabilities -> auroc / auprc / f1 / precision / recall / balanced accuracy / mcc / loss, Youden tau
    on the validation split                                                       (test, :4653-5494; metrics/*)
  * best-checkpoint payload keys + `<stem>.pth` pointer + `<stem>_last.pth`      (:7037-7111, :7550-7626, :3932-3941)

What is NOT mirrored: experiment YAMLs (--exp-config/--override), data packs / manifests / perturbations, morphology
strata, per-frame CSV exports, TensorBoard, provenance blocks - 8 k lines of experiment bookkeeping that never touch the
device path (SURVEY.md section 8 marks them out of scope).  Their flags parse (so launch scripts keep working) and are
either ignored or rejected with a clear message.  Data comes from `--dataset synthetic` (seeded fp32 batches, the bench
workload) or from loaders handed to `run(args, loaders=...)` yielding `(images, labels[, meta])` like PackDataset does.

One process per GPU (torchrun); gradients are averaged by the engine's bucketed NCCL all-reduce (dp.py), the epoch loss
by one all-reduce per epoch instead of the reference's all-reduce + barrier per step (:4547-4551).
"""
from __future__ import annotations

import argparse
import json
import math
import os
import random
import shutil
import sys
import time
from pathlib import Path

import numpy as np
import torch
import torch.distributed as dist

from . import metrics as M
from . import utils
from .engine import ClassifierEngine, PlateauLR, finetune_cosine_factor

# ---------------------------------------------------------------------------------------------------------------------
# command line (reference get_args, train_classification.py:7706-7982): (flags, kwargs) rows
_STORE = [
    (("--exp-config",), dict(type=str, dest="exp_config")),
    (("--override",), dict(action="append", nargs="+", dest="overrides", default=None, metavar="KEY=VALUE")),
    (("--model-key",), dict(type=str, dest="model_key")),
    (("--architecture",), dict(type=str, choices=["vit_b"], default="vit_b", dest="arch")),
    (("--pretraining",), dict(type=str, default=None, choices=["Hyperkvasir", "ImageNet_class", "ImageNet_self", "random"])),
    (("--ss-framework",), dict(type=str, choices=["mae"])),
    (("--checkpoint",), dict(type=str, dest="ckpt", default=None)),
    (("--parent-checkpoint",), dict(type=str, dest="parent_checkpoint", default=None)),
    (("--finetune-mode",), dict(type=str, default="full", choices=["none", "full", "head+1", "head+2"])),
    (("--frozen",), dict(action="store_true", default=False)),
    (("--dataset",), dict(type=str, default=None)),
    (("--train-pack",), dict(type=str, default=None)),
    (("--val-pack",), dict(type=str, default=None)),
    (("--test-pack",), dict(type=str, default=None)),
    (("--train-split",), dict(type=str, default="train")),
    (("--val-split",), dict(type=str, default="val")),
    (("--test-split",), dict(type=str, default="test")),
    (("--morphology-eval",), dict(nargs="+", dest="morphology_eval", default=None)),
    (("--dataset-percent",), dict(type=float, default=None)),
    (("--dataset-seed",), dict(type=int, default=None)),
    (("--pack-root",), dict(type=str, default=str(Path("data_packs")))),
    (("--roots",), dict(type=str, default=str(Path("data") / "roots.json"))),
    (("--perturbation-key",), dict(type=str, default="ssl4polyp")),
    (("--perturbation-splits",), dict(nargs="*", default=["test"])),
    (("--image-size",), dict(type=int, default=224)),
    (("--class-weights",), dict(type=str, default=None)),
    (("--epochs",), dict(type=int, default=50)),
    (("--batch-size",), dict(type=int, default=16)),
    (("--learning-rate",), dict(type=float, default=1e-4, dest="lr")),
    (("--weight-decay",), dict(type=float, default=0.05)),
    (("--scheduler",), dict(type=str, choices=["none", "cosine", "plateau"], default="none")),
    (("--max-train-steps",), dict(type=int, default=None)),
    (("--warmup-epochs",), dict(type=int, default=0)),
    (("--min-lr",), dict(type=float, default=1e-6)),
    (("--scheduler-patience",), dict(type=int, default=2)),
    (("--scheduler-factor",), dict(type=float, default=0.5)),
    (("--early-stop-monitor",), dict(type=str, default="val_loss")),
    (("--early-stop-patience",), dict(type=int, default=0)),
    (("--early-stop-mode",), dict(type=str, default="auto", choices=["min", "max", "auto"])),
    (("--early-stop-min-delta",), dict(type=float, default=0.0)),
    (("--early-stop-min-epochs",), dict(type=int, default=0)),
    (("--threshold-policy",), dict(type=str, default="auto", choices=["auto", "youden", "none", "f1-morph"])),
    (("--workers",), dict(type=int, default=8)),
    (("--prefetch-factor",), dict(type=int, default=2)),
    (("--pin-memory",), dict(dest="pin_memory", action="store_true", default=True)),
    (("--no-pin-memory",), dict(dest="pin_memory", action="store_false")),
    (("--persistent-workers",), dict(dest="persistent_workers", action="store_true", default=True)),
    (("--no-persistent-workers",), dict(dest="persistent_workers", action="store_false")),
    (("--limit-train-batches",), dict(type=int, default=None)),
    (("--limit-val-batches",), dict(type=int, default=None)),
    (("--limit-test-batches",), dict(type=int, default=None)),
    (("--precision",), dict(choices=["amp", "fp32"], default="amp")),
    (("--log-interval",), dict(type=int, default=10)),
    (("--output-dir",), dict(type=str, default=str(Path("checkpoints") / "classification"), dest="output_dir")),
    (("--seed",), dict(type=int, default=42)),
    (("--seeds",), dict(type=int, nargs="+", default=None)),
]
# additions of this implementation (not in the reference): the synthetic data source's size and class count
_EXTRA = [
    (("--synthetic-samples",), dict(type=int, default=256, help="images per synthetic split (with --dataset synthetic)")),
    (("--num-classes",), dict(type=int, default=2, help="classes of the synthetic source")),
]


def build_parser() -> argparse.ArgumentParser:
    parser = argparse.ArgumentParser(description="Fine-tune pretrained model for classification")
    for flags, kw in _STORE + _EXTRA:
        parser.add_argument(*flags, **kw)
    return parser


def get_args(argv=None):
    argv = sys.argv[1:] if argv is None else list(argv)
    args = build_parser().parse_args(argv)
    # :7974-7980 - remembers whether --seed was typed so that --seeds cannot silently override it
    args._seed_explicit = any(a == "--seed" or a.startswith("--seed=") for a in argv)
    return args


def _active_seed(args) -> int:
    """:1589-1612 - the first entry of --seeds unless --seed was given explicitly."""
    if getattr(args, "seeds", None) and not getattr(args, "_seed_explicit", False):
        return int(args.seeds[0])
    return int(args.seed)


def set_determinism(seed: int) -> None:
    """:1615-1629"""
    random.seed(seed)
    np.random.seed(seed)
    torch.manual_seed(seed)
    if torch.cuda.is_available():
        torch.cuda.manual_seed_all(seed)


# ---------------------------------------------------------------------------------------------------------------------
# fine-tune regimes (reference classification/finetune.py:29-92)
_MODES = ("none", "full", "head+1", "head+2")


def normalise_finetune_mode(raw, *, default: str = "full") -> str:
    mode = str(default if raw is None else raw).strip().lower() or str(default).strip().lower()
    if mode not in _MODES:
        raise ValueError(f"Unsupported fine-tuning mode '{raw}'. Expected one of {sorted(_MODES)}.")
    return mode


def configure_finetune_parameters(model: torch.nn.Module, mode: str) -> None:
    """`full`: everything trains.  Otherwise only `lin_head` (+ the last 1 / 2 entries of `model.blocks` when that is
    a ModuleList - timm 0.4.12 holds VisionTransformer.blocks in nn.Sequential, which the reference's isinstance test
    skips, finetune.py:76-77; preserved knowingly so both arms train the same set)."""
    mode = normalise_finetune_mode(mode)
    every = mode == "full"
    for p in model.parameters():
        p.requires_grad_(every)
        p.grad = None
    if not every:
        heads = [h for h in (getattr(model, "lin_head", None), getattr(model, "head", None)) if isinstance(h, torch.nn.Module)]
        for h in heads:
            for p in h.parameters():
                p.requires_grad_(True)
        blocks = getattr(model, "blocks", None)
        tail = {"head+1": 1, "head+2": 2}.get(mode, 0)
        if tail and isinstance(blocks, (torch.nn.ModuleList, list, tuple)) and len(blocks) > 0:
            for blk in list(blocks)[-tail:]:
                for p in blk.parameters():
                    p.requires_grad_(True)
    if hasattr(model, "frozen"):
        model.frozen = mode == "none"


# ---------------------------------------------------------------------------------------------------------------------
# monitor / early stopping (reference :3297-3345, :3903-3915)
def resolve_monitor_mode(monitor, mode) -> str:
    m = (mode or "auto").lower()
    if m not in ("min", "max", "auto"):
        raise ValueError(f"Unsupported early-stop mode: {mode!r}")
    if m != "auto":
        return m
    name = (monitor or "").lower()
    return "min" if name.endswith("loss") or name.startswith("loss") else "max"


def resolve_monitor_key(raw) -> str:
    if not raw:
        return "loss"
    key = raw.lower()
    return key[4:] if key.startswith("val_") else key


def improved(current: float, best, *, mode: str, min_delta: float) -> bool:
    if best is None or math.isnan(best):
        return True
    if math.isnan(current):
        return False
    return current < best - min_delta if mode == "min" else current > best + min_delta


def should_stop(no_improve: int, patience: int, epochs_done: int, min_epochs: int) -> bool:
    return patience > 0 and epochs_done >= max(min_epochs, 0) and no_improve >= patience


# ---------------------------------------------------------------------------------------------------------------------
# data: a seeded synthetic source, or caller-supplied loaders
class SyntheticFrames:
    """`n` fp32 224x224 frames with seeded labels; a class-dependent mean shift makes the task learnable so the loop's
    metrics move.  Iterating yields `(images, labels)` host tensors in `batch`-sized chunks; `labels_list` mirrors
    PackDataset's attribute that `build` reads class counts from (:5610-5622).  `shard=(rank, world)` splits the samples
    like `DistributedSampler` (training only; evaluation runs the whole split on every rank)."""

    def __init__(self, n: int, batch: int, n_class: int = 2, image_size: int = 224, seed: int = 0, shard=(0, 1), shuffle=False):
        g = torch.Generator().manual_seed(seed)
        self.labels = torch.randint(0, n_class, (n,), generator=g)
        if n >= n_class:
            self.labels[:n_class] = torch.arange(n_class)        # every class present (AUROC defined)
        self.labels_list = self.labels.tolist()
        self.n, self.batch, self.n_class, self.size, self.seed = n, batch, n_class, image_size, seed
        self.shard, self.shuffle, self.epoch = shard, shuffle, 0

    def set_epoch(self, epoch: int) -> None:
        self.epoch = epoch

    def _frame_batch(self, idx: torch.Tensor):
        out = torch.empty(len(idx), 3, self.size, self.size)
        for j, i in enumerate(idx.tolist()):
            g = torch.Generator().manual_seed(self.seed * 1000003 + i)
            out[j] = torch.randn(3, self.size, self.size, generator=g) + 0.35 * (float(self.labels[i]) - 0.5 * (self.n_class - 1))
        return out, self.labels[idx]

    def _rank_order(self) -> torch.Tensor:
        """This rank's sample indices for the current epoch: `DistributedSampler` semantics (the reference's training
        sampler) - pad the (shuffled) order to a multiple of the world size by wrapping around and take every world-th
        sample, so EVERY rank sees the same number of samples and therefore the same number of equally-sized batches
        (a rank with one step more would pair its bucket all-reduce with the other ranks' epoch-end collective)."""
        order = torch.arange(self.n)
        if self.shuffle:
            order = torch.randperm(self.n, generator=torch.Generator().manual_seed(self.seed + 7919 * self.epoch))
        r, w = self.shard
        if w > 1:
            total = (self.n + w - 1) // w * w
            if total > self.n:
                order = torch.cat([order, order[:total - self.n]])
            order = order[r::w]
        return order

    def __len__(self):
        r, w = self.shard
        per_rank = (self.n + w - 1) // w
        return (per_rank + self.batch - 1) // self.batch

    def __iter__(self):
        order = self._rank_order()
        for start in range(0, len(order), self.batch):
            yield self._frame_batch(order[start:start + self.batch])


def _class_weights(args, labels, n_class):
    """:5613-5631 - N / (n_class * count) from the training labels unless --class-weights overrides."""
    counts = np.bincount(np.asarray(labels, dtype=np.int64), minlength=n_class)
    weights = [(len(labels) / (n_class * c)) if c > 0 else 0.0 for c in counts]
    if args.class_weights is not None:
        weights = [float(w) for w in args.class_weights.split(",")]
        if len(weights) != n_class:
            raise ValueError("Number of class weights must match number of classes")
    return weights, counts.tolist()


def _pos_weight(weights, counts) -> float:
    """:6089-6098"""
    neg = float(counts[0]) if len(counts) >= 1 else None
    pos = float(counts[1]) if len(counts) >= 2 else None
    if pos and pos > 0:
        return neg / pos if neg is not None else 1.0
    if len(weights) >= 2 and weights[0] > 0 and weights[1] > 0:
        return float(weights[1] / weights[0])
    return 1.0


def build_model(args, n_class: int):
    """:5643-5655"""
    if args.pretraining in ("Hyperkvasir", "ImageNet_self"):
        assert args.ckpt is not None and os.path.exists(args.ckpt)
        return utils.get_MAE_backbone(args.ckpt, True, n_class, args.frozen, None)
    if args.pretraining == "ImageNet_class":
        return utils.get_ImageNet_or_random_ViT(True, n_class, args.frozen, None, ImageNet_weights=True)
    if args.pretraining == "random":
        return utils.get_ImageNet_or_random_ViT(True, n_class, args.frozen, None, ImageNet_weights=False)
    raise ValueError("--pretraining is required (Hyperkvasir | ImageNet_class | ImageNet_self | random)")


def probabilities(logits: torch.Tensor) -> torch.Tensor:
    """:3398-3416 - two classes: sigmoid(logit1 - logit0) stacked as (neg, pos); otherwise softmax."""
    if logits.ndim == 2 and logits.size(1) == 2:
        pos = torch.sigmoid(logits[:, 1] - logits[:, 0])
        return torch.stack((1.0 - pos, pos), dim=1)
    return torch.softmax(logits, dim=1)


def supervised_loss(logits, targets, n_class, pos_weight=None, class_weights=None):
    """:3361-3374 with the loss objects of :6086-6104."""
    if n_class == 2:
        return torch.nn.functional.binary_cross_entropy_with_logits(logits[:, 1] - logits[:, 0], targets.float(), pos_weight=pos_weight)
    return torch.nn.functional.cross_entropy(logits, targets, weight=class_weights)


def _limit(value):
    return value if value is not None and value > 0 else None


@torch.no_grad()
def evaluate(model, loader, device, n_class, *, pos_weight=None, class_weights=None, tau=None, limit=None, frame_transform=None):
    """One pass in eval mode (reference `test`, :4653-5494, minus exports/perturbations/morphology): every logit stays
    on the device until the end of the split, then one D2H copy and the host metrics.  A batch whose first element is a list
    of uint8 HWC frames goes through `input_pipeline.EvalTransform` (Resize bilinear -> ToTensor -> Normalize, bit-identical to
    classification/data/transforms.py:233-258 with augmentations off)."""
    was_training = model.training
    model.eval()
    logits, targets = [], []
    for b, batch in enumerate(loader):
        if limit is not None and b >= limit:
            break
        data, target = batch[0], batch[1]
        if isinstance(data, (list, tuple)):             # decoded uint8 HWC frames: the val / test transform runs on the device
            if frame_transform is None:
                from .input_pipeline import EvalTransform
                frame_transform = EvalTransform(224, device=device)
            data = frame_transform(data)
        logits.append(model(data.to(device, non_blocking=True)).float())
        targets.append(target.to(device, non_blocking=True))
    model.train(was_training)
    if not logits:
        return {"loss": float("nan"), "auroc": float("nan"), "count": 0}
    lg, tg = torch.cat(logits), torch.cat(targets)
    probs = probabilities(lg)
    res = {"loss": float(supervised_loss(lg, tg, n_class, pos_weight, class_weights).item()), "count": int(tg.numel())}
    p, y = probs.cpu(), tg.cpu()
    res.update(auroc=M.auroc(p, y, n_class), auprc=M.auprc(p, y, n_class), f1=M.mean_f1(p, y, n_class, tau),
               precision=M.mean_precision(p, y, n_class, tau), recall=M.mean_recall(p, y, n_class, tau),
               balanced_accuracy=M.balanced_accuracy(p, y, n_class, tau), tau=tau,
               class_counts=np.bincount(y.numpy(), minlength=n_class).tolist())
    if n_class == 2:
        res["mcc"] = M.mcc(p, y, tau)
        res["prevalence"] = float(y.float().mean())
    res["_probabilities"], res["_targets"] = p, y
    return res


def train_epoch(engine, loader, device, epoch, args, *, lr, global_step=0, log=print, rank=0, world=1, frame_transform=None):
    """Reference `train_epoch` (:4446-4649): one pass over the loader, a step per batch.  Full-size batches replay the
    captured step; a ragged tail runs eagerly.  The per-step loss stays on the device (the reference all-reduces it and
    hits a barrier every step, :4547-4551); the epoch mean is reduced once at the end.  A batch whose first element is a
    list of decoded uint8 HWC frames goes through `input_pipeline.TrainAugment`: the stage-'train' transform of
    classification/data/transforms.py:233-258 (Resize, ColorJitter, GaussianBlur, flips, rotation, ToTensor, Normalize)
    with the reference's own RNG draws on the host and the pixel work on the device."""
    engine.model.train()
    limit = _limit(args.limit_train_batches)
    max_steps = _limit(args.max_train_steps)
    per_rank = max(args.batch_size // world, 1)
    total, steps, t0 = None, 0, time.time()
    for b, batch in enumerate(loader):
        if limit is not None and b >= limit:
            break
        if max_steps is not None and global_step + steps >= max_steps:
            break
        if isinstance(batch[0], (list, tuple)):        # decoded frames: augment + normalise on the device
            if frame_transform is None:
                frame_transform = getattr(engine, "_train_augment", None)
                if frame_transform is None:
                    from .input_pipeline import TrainAugment
                    frame_transform = engine._train_augment = TrainAugment(224, device=device)
            data = frame_transform(batch[0])
        else:
            data = batch[0].to(device, non_blocking=True)
        target = batch[1].to(device, non_blocking=True)
        loss = engine.step(data, target, lr=lr, graph=data.shape[0] == per_rank)
        total = loss.clone() if total is None else total + loss
        steps += 1
        if rank == 0 and args.log_interval > 0 and (b + 1) % args.log_interval == 0:
            log(f"Train Epoch: {epoch} [{b + 1}/{len(loader)}]\tLoss: {float(loss):.6f}\tlr: {lr:.3e}\t{time.time() - t0:.1f}s")
    if total is None:
        return float("nan"), 0
    mean = total / steps
    if world > 1:
        dist.all_reduce(mean)
        mean = mean / world
    if engine.nonfinite_count():
        raise FloatingPointError("non-finite gradient during fine-tuning")
    return float(mean), steps


# ---------------------------------------------------------------------------------------------------------------------
# checkpoints (reference :3878-3941, :7037-7111, :7550-7628)
def selection_tag(monitor) -> str:
    """`val_loss` -> `valLoss`, `val_auroc` -> `valAUROC`, empty -> `best` (:3878-3900)."""
    import re
    tokens = [t for t in re.split(r"[^0-9A-Za-z]+", monitor or "") if t]
    special = {"auroc": "AUROC", "auc": "AUC", "auprc": "AUPRC"}
    parts = []
    for i, tok in enumerate(tokens):
        low = tok.lower()
        if low in special:
            parts.append(special[low])
        elif low == "loss":
            parts.append("loss" if i == 0 else "Loss")
        else:
            parts.append(low if i == 0 else tok.capitalize())
    return "".join(parts) or "best"


def update_pointer(pointer: Path, target: Path) -> None:
    """`<stem>.pth` -> newest checkpoint: a relative symlink, a copy where symlinks are refused (:3932-3941)."""
    pointer.parent.mkdir(parents=True, exist_ok=True)
    try:
        if pointer.is_symlink() or pointer.exists():
            pointer.unlink()
        pointer.symlink_to(target.name)
    except OSError:
        shutil.copy2(target, pointer)


def _payload(engine, model, epoch, loss, monitor_name, monitor_value, val, test):
    g = lambda d, k: (None if d is None or d.get(k) is None else float(d[k]))
    return {"epoch": int(epoch), "model_state_dict": model.state_dict(), "optimizer_state_dict": engine.optimizer_state_dict(),
            "scaler_state_dict": {},                  # no GradScaler: bf16 activations with fp32 accumulation need none
            "loss": None if loss is None else float(loss), "val_perf": monitor_value, "val_loss": g(val, "loss"),
            "val_auroc": g(val, "auroc"), "test_perf": g(test, "auroc"), "test_loss": g(test, "loss"), "test_auroc": g(test, "auroc"),
            "monitor_value": monitor_value, "monitor_metric": monitor_name, "test_monitor_value": g(test, "auroc"),
            "py_state": random.getstate(), "np_state": np.random.get_state(), "torch_state": torch.get_rng_state()}


def _public(metrics_dict):
    return None if metrics_dict is None else {k: v for k, v in metrics_dict.items() if not k.startswith("_")}


# ---------------------------------------------------------------------------------------------------------------------
def run(args, loaders=None, log=print):
    """The body of reference `train` (:6012-7703).  `loaders`: optional {"train","val","test"} iterables yielding
    (images fp32 [B,3,224,224], labels int64 [B], ...); the train one needs `labels_list`.  Returns a summary dict."""
    if not torch.cuda.is_available():
        raise RuntimeError("ssl4polyp_b200.train_classification needs a CUDA device (sm_100a); there is no CPU fallback")
    for flag in ("exp_config", "overrides", "train_pack", "val_pack", "test_pack", "parent_checkpoint", "morphology_eval"):
        if getattr(args, flag, None) and loaders is None:
            raise SystemExit(f"--{flag.replace('_', '-')}: experiment configs / data packs are outside the accelerated path; "
                             f"pass the reference's own DataLoaders to run(args, loaders=...) or use --dataset synthetic")
    world, rank, local = (int(os.environ.get(k, d)) for k, d in (("WORLD_SIZE", "1"), ("RANK", "0"), ("LOCAL_RANK", "0")))
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=device)
    seed = _active_seed(args)
    set_determinism(seed)
    args.finetune_mode = "none" if args.frozen else normalise_finetune_mode(args.finetune_mode)
    per_rank = max(args.batch_size // world, 1)                        # --batch-size is GLOBAL (:5504)

    if loaders is None:
        if (args.dataset or "").lower() != "synthetic":
            raise SystemExit("--dataset synthetic is the only built-in source (no data packs on this path)")
        n, k = args.synthetic_samples, args.num_classes
        loaders = {"train": SyntheticFrames(n, per_rank, k, args.image_size, seed, (rank, world), shuffle=True),
                   "val": SyntheticFrames(max(n // 4, k), per_rank, k, args.image_size, seed + 1),
                   "test": SyntheticFrames(max(n // 4, k), per_rank, k, args.image_size, seed + 2)}
    train_loader, val_loader, test_loader = loaders.get("train"), loaders.get("val"), loaders.get("test")
    labels = list(getattr(train_loader, "labels_list", None) or getattr(getattr(train_loader, "dataset", None), "labels_list", None) or [])
    if not labels:
        raise ValueError("No labelled datasets available; ensure at least one split provides labels.")
    n_class = len(set(labels))
    weights, counts = _class_weights(args, labels, n_class)
    pos_w = torch.tensor(_pos_weight(weights, counts), device=device) if n_class == 2 else None
    cls_w = torch.tensor(weights, device=device, dtype=torch.float32) if n_class != 2 else None

    model = build_model(args, n_class).to(device).train()
    configure_finetune_parameters(model, args.finetune_mode)
    engine = ClassifierEngine(model, lr=args.lr, weight_decay=args.weight_decay, pos_weight=pos_w, class_weights=cls_w)

    out_dir = Path(args.output_dir).expanduser()
    stem = out_dir / f"{args.pretraining}_{args.dataset or 'custom'}_{args.finetune_mode.replace('+', 'p')}_s{seed}"
    log_path = stem.parent / (stem.name + ".log")
    if rank == 0:
        out_dir.mkdir(parents=True, exist_ok=True)
        (stem.parent / (stem.name + ".args.json")).write_text(json.dumps({k: v for k, v in vars(args).items() if not k.startswith("_")}, indent=2, default=str))

    def say(line):
        if rank == 0:
            log(line)
            with open(log_path, "a") as fh:
                fh.write(line + "\n")

    monitor_name = args.early_stop_monitor
    monitor_key, monitor_mode = resolve_monitor_key(monitor_name), resolve_monitor_mode(monitor_name, args.early_stop_mode)
    tag = selection_tag(monitor_name)
    plateau = PlateauLR(args.lr, args.scheduler_factor, args.scheduler_patience, args.min_lr) if args.scheduler == "plateau" else None
    lr = args.lr * (finetune_cosine_factor(0, args.epochs, args.warmup_epochs) if args.scheduler == "cosine" else 1.0)
    policy = args.threshold_policy
    use_youden = n_class == 2 and policy in ("auto", "youden")          # f1-morph needs morphology labels: not built
    if policy == "f1-morph":
        raise SystemExit("--threshold-policy f1-morph needs morphology strata (not built); use auto | youden | none")

    best, best_epoch, best_path, no_improve, global_step = None, None, None, 0, 0
    last = {"epoch": 0, "loss": None, "val": None, "test": None, "monitor": None}
    say(f"fine-tune: {args.pretraining} mode={args.finetune_mode} classes={n_class} counts={counts} "
        f"trainable={sum(p.numel() for p in model.parameters() if p.requires_grad)} batch={args.batch_size} ({per_rank}/GPU x {world})")
    for epoch in range(1, args.epochs + 1):
        if hasattr(train_loader, "set_epoch"):
            train_loader.set_epoch(epoch)
        elif hasattr(getattr(train_loader, "sampler", None), "set_epoch"):
            train_loader.sampler.set_epoch(epoch)
        t0 = time.time()
        loss, steps = train_epoch(engine, train_loader, device, epoch, args, lr=lr, global_step=global_step, log=say, rank=rank, world=world)
        global_step += steps
        tau = None
        val = evaluate(model, val_loader, device, n_class, pos_weight=pos_w, class_weights=cls_w,
                       limit=_limit(args.limit_val_batches)) if val_loader is not None else None
        if val is not None and use_youden and val["count"] and len(set(val["_targets"].tolist())) == 2:
            tau = M.youden_j_threshold(val["_probabilities"], val["_targets"])          # :2605-2663, policy on the val split
            p, y = val["_probabilities"], val["_targets"]
            val.update(tau=tau, f1=M.mean_f1(p, y, n_class, tau), precision=M.mean_precision(p, y, n_class, tau),
                       recall=M.mean_recall(p, y, n_class, tau), balanced_accuracy=M.balanced_accuracy(p, y, n_class, tau), mcc=M.mcc(p, y, tau))
        test = evaluate(model, test_loader, device, n_class, pos_weight=pos_w, class_weights=cls_w, tau=tau,
                        limit=_limit(args.limit_test_batches)) if test_loader is not None else None
        source = val if val is not None else {"loss": loss}
        monitor_value = float(source.get(monitor_key, float("nan")))
        say(f"epoch {epoch}: train_loss={loss:.5f} lr={lr:.3e} steps={steps} "
            + (f"val_loss={val['loss']:.5f} val_auroc={val['auroc']:.4f} " if val else "")
            + (f"test_auroc={test['auroc']:.4f} " if test else "") + f"monitor[{monitor_name}]={monitor_value:.5f} ({time.time() - t0:.1f}s)")
        last.update(epoch=epoch, loss=loss, val=val, test=test, monitor=monitor_value)
        if improved(monitor_value, best, mode=monitor_mode, min_delta=float(args.early_stop_min_delta)):
            best, best_epoch, no_improve = monitor_value, epoch, 0
            if rank == 0:
                best_path = stem.parent / f"{stem.name}_e{epoch:02d}_{tag}.pth"
                payload = _payload(engine, model, epoch, loss, monitor_name, monitor_value, val, test)
                if tau is not None:
                    payload["thresholds"] = {f"{args.dataset or 'custom'}_val_youden": float(tau)}
                torch.save(payload, best_path)
                update_pointer(stem.parent / (stem.name + ".pth"), best_path)
        else:
            no_improve += 1
        # schedulers step once per epoch that took optimiser steps (:6816-6833): LambdaLR cosine, or
        # ReduceLROnPlateau(mode="max") fed the MONITORED value (whatever --early-stop-monitor names)
        if steps > 0 and args.scheduler == "cosine":
            lr = args.lr * finetune_cosine_factor(epoch, args.epochs, args.warmup_epochs)
        elif steps > 0 and plateau is not None:
            lr = plateau.step(monitor_value)
        max_steps = _limit(args.max_train_steps)
        if should_stop(no_improve, args.early_stop_patience, epoch, args.early_stop_min_epochs):
            say(f"early stop at epoch {epoch}: no improvement in {monitor_name} for {no_improve} epochs")
            break
        if max_steps is not None and global_step >= max_steps:
            say(f"reached --max-train-steps {max_steps}")
            break

    summary = {"seed": seed, "epochs_completed": last["epoch"], "global_step": global_step, "best_epoch": best_epoch,
               "monitor_metric": monitor_name, "best_monitor_value": best, "train_loss": last["loss"],
               "val": _public(last["val"]), "test": _public(last["test"]), "checkpoint": None if best_path is None else str(best_path)}
    if rank == 0 and last["epoch"] > 0:
        last_path = stem.parent / f"{stem.name}_last_e{last['epoch']:02d}_{tag}.pth"
        torch.save(_payload(engine, model, last["epoch"], last["loss"], monitor_name, last["monitor"], last["val"], last["test"]), last_path)
        update_pointer(stem.parent / (stem.name + "_last.pth"), last_path)
        (stem.parent / (stem.name + ".metrics.json")).write_text(json.dumps(summary, indent=2, default=str))
    if world > 1:
        dist.barrier()
    return summary


def main(argv=None):
    args = get_args(argv)
    return run(args)


if __name__ == "__main__":
    main()
