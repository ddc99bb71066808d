"""`main_pretrain.py` with the reference's command line on the B200 engine.

    python -m ssl4polyp_b200.main_pretrain --model mae_vit_base_patch16 --norm_pix_loss --batch_size 256 \
        --data_path /data/hyperkvasir --no_train_dir --epochs 400 --warmup_epochs 40 --blr 1.5e-4 --weight_decay 0.05
    torchrun --nproc-per-node 8 -m ssl4polyp_b200.main_pretrain ...            (one process per GPU)

Flags, defaults and their meaning follow /root/reference/src/ssl4polyp/models/mae/main_pretrain.py:47-135; the loop
follows `engine_pretrain.train_one_epoch` (:22-100) and `main` (:139-330):
  * per-ITERATION half-cosine lr with linear warm-up (util/lr_sched.py:9-21), `lr = blr * eff_batch / 256`;
  * `seed + rank` RNG streams (:146-148), `DistributedSampler(seed=args.seed)` + `set_epoch(seed + epoch)`;
  * AdamW(betas 0.9/0.95) with timm's no-decay rule (:216-218);
  * non-finite loss -> `sys.exit(1)` (engine_pretrain.py:59-62);
  * checkpoints `<output_dir>/ckpts/checkpoint-<epoch>.pth` = {"model","optimizer","epoch","scaler","args"} plus the
    `last.pth` symlink, keep-last / milestone rotation, SIGTERM/SIGINT save, `log.txt` json lines (:231-312,
    util/misc.py:306-335); `--resume` restores model, optimiser moments and epoch (:337-352).
What differs, on purpose: the step is `PretrainEngine.step` (our kernels, fused AdamW, bucketed all-reduce, no
GradScaler: bf16 operands need no loss scaling, so "scaler" is an empty dict); `--precision fp32` is refused (the
tensor-core path is bf16 operands / fp32 accumulate by construction); `--data_path synthetic[:N]` yields N random
batches per epoch so the entry point can be exercised without a dataset.  TensorBoard logging is not part of the hot
path (SURVEY.md §8 out of scope): `--log_dir` is accepted and ignored.
"""
from __future__ import annotations

import argparse
import datetime
import glob
import json
import math
import os
import signal
import sys
import time
from pathlib import Path

import numpy as np
import torch
import torch.distributed as dist


def get_args_parser():
    parser = argparse.ArgumentParser('MAE pre-training', add_help=False)
    parser.add_argument('--batch_size', default=64, type=int,
                        help='Batch size per GPU (effective batch size is batch_size * accum_iter * # gpus')
    parser.add_argument('--epochs', default=400, type=int)
    parser.add_argument('--accum_iter', '--acum_iter', dest='accum_iter', default=1, type=int)
    parser.add_argument('--model', default='mae_vit_large_patch16', type=str, metavar='MODEL')
    parser.add_argument('--input_size', default=224, type=int)
    parser.add_argument('--mask_ratio', default=0.75, type=float)
    parser.add_argument('--norm_pix_loss', action='store_true')
    parser.set_defaults(norm_pix_loss=False)
    parser.add_argument('--weight_decay', type=float, default=0.05)
    parser.add_argument('--lr', type=float, default=None, metavar='LR')
    parser.add_argument('--blr', type=float, default=1e-3, metavar='LR')
    parser.add_argument('--min_lr', type=float, default=0., metavar='LR')
    parser.add_argument('--warmup_epochs', type=int, default=40, metavar='N')
    parser.add_argument('--data_path', default='/datasets01/imagenet_full_size/061417/', type=str)
    parser.add_argument('--no_train_dir', action='store_true')
    parser.add_argument('--output_dir', default=str(Path('checkpoints') / 'mae' / 'pretrain'))
    parser.add_argument('--log_dir', default=str(Path('outputs') / 'mae' / 'pretrain'))
    parser.add_argument('--device', default='cuda')
    parser.add_argument('--seed', default=0, type=int)
    parser.add_argument('--resume', default='')
    parser.add_argument('--save-freq-epochs', type=int, default=20)
    parser.add_argument('--save-freq-mins', type=int, default=0)
    parser.add_argument('--keep-last', type=int, default=5)
    parser.add_argument('--keep-every-n-epochs', type=int, default=50)
    parser.add_argument('--start_epoch', default=0, type=int, metavar='N')
    parser.add_argument('--num_workers', default=10, type=int)
    parser.add_argument('--pin_mem', action='store_true')
    parser.add_argument('--no_pin_mem', action='store_false', dest='pin_mem')
    parser.set_defaults(pin_mem=True)
    parser.add_argument('--world_size', default=1, type=int)
    parser.add_argument('--local_rank', default=-1, type=int)
    parser.add_argument('--dist_on_itp', action='store_true')
    parser.add_argument('--dist_url', default='env://')
    parser.add_argument('--precision', choices=['amp', 'fp32'], default='amp')
    # not in the reference: run the Compose's pixel work (crop, bicubic resize, flip, normalise) on the GPU
    parser.add_argument('--device_transform', action='store_true',
                        help='DataLoader workers only decode; RandomResizedCrop/flip/ToTensor/Normalize run on the device (bit-identical)')
    parser.add_argument('--no_graph', action='store_true',
                        help='enqueue every kernel of the step from Python instead of replaying the captured CUDA graph (debugging)')
    return parser


class _SyntheticBatches:
    """`--data_path synthetic[:N]`: N batches per epoch of ImageNet-normalised-looking noise (no dataset on the box).
    A pool of up to 4 distinct pinned batches is drawn once and cycled (two 154 MB batches already exceed the L2), so the
    host never becomes the bottleneck the way a per-step `torch.randn` of 38 M values would."""

    def __init__(self, n, batch, size, seed):
        self.n, self.batch, self.size, self.seed = n, batch, size, seed
        self.sampler = None
        g = torch.Generator().manual_seed(seed)
        self.pool = [torch.randn(batch, 3, size, size, generator=g).pin_memory() for _ in range(min(n, 4))]

    def __len__(self):
        return self.n

    def __iter__(self):
        for i in range(self.n):
            yield self.pool[i % len(self.pool)], None


def _init_distributed(args):
    """util/misc.py init_distributed_mode for the torchrun environment (RANK / WORLD_SIZE / LOCAL_RANK)."""
    if 'RANK' in os.environ and 'WORLD_SIZE' in os.environ:
        args.rank, args.world_size = int(os.environ['RANK']), int(os.environ['WORLD_SIZE'])
        args.gpu = int(os.environ.get('LOCAL_RANK', 0))
        args.distributed = args.world_size > 1
    else:
        args.rank, args.world_size, args.gpu, args.distributed = 0, 1, 0, False
    torch.cuda.set_device(args.gpu)
    if args.distributed:
        from .dp import configure_overlap
        configure_overlap(int(os.environ.get("SSL4P_DP_SMS", "0")))
        dist.init_process_group('nccl', init_method=args.dist_url, device_id=torch.device('cuda', args.gpu))
        dist.barrier()


def _build_loader(args):
    if args.data_path.startswith('synthetic'):
        n = int(args.data_path.split(':')[1]) if ':' in args.data_path else 8
        return _SyntheticBatches(n, args.batch_size, args.input_size, args.seed + args.rank)
    from torchvision import datasets, transforms
    transform_train = transforms.Compose([
        transforms.RandomResizedCrop(args.input_size, scale=(0.2, 1.0), interpolation=3),  # 3 is bicubic
        transforms.RandomHorizontalFlip(),
        transforms.ToTensor(),
        transforms.Normalize(mean=[0.485, 0.456, 0.406], std=[0.229, 0.224, 0.225])])
    data_dir = args.data_path if args.no_train_dir else os.path.join(args.data_path, 'train')
    on_device = getattr(args, 'device_transform', False)
    # --device_transform: the workers only decode (uint8 HWC); crop / bicubic resize / flip / normalise run on the GPU,
    # bit-identical to the Compose above (input_pipeline.TrainTransform), in train_one_epoch
    dataset_train = datasets.ImageFolder(data_dir, transform=_decode_rgb_u8 if on_device else transform_train)
    print(dataset_train)
    sampler = torch.utils.data.DistributedSampler(dataset_train, num_replicas=args.world_size, rank=args.rank,
                                                  shuffle=True, seed=args.seed)
    return torch.utils.data.DataLoader(dataset_train, sampler=sampler, batch_size=args.batch_size,
                                       num_workers=args.num_workers, pin_memory=args.pin_mem and not on_device, drop_last=True,
                                       collate_fn=_collate_frames if on_device else None)


def _decode_rgb_u8(img):
    import numpy as np
    return torch.from_numpy(np.asarray(img.convert('RGB')).copy())


def _collate_frames(batch):
    return [b[0] for b in batch], torch.tensor([b[1] for b in batch])


def train_one_epoch(engine, data_loader, device, epoch, args, save_ckpt_fn=None, last_time_save=None, time_save_sec=None):
    """One epoch of engine_pretrain.train_one_epoch (:22-100) on `PretrainEngine`, fed through `engine.HostPipeline`: the
    step is the captured CUDA graph (when --accum_iter is 1), the next batch is copied to the device while the current
    step runs, and each loss is read from a pinned slot one step late instead of `loss.item()` + `cuda.synchronize()` +
    a scalar all-reduce per iteration (:55,74,81).  Kept observable behaviour: per-iteration lr schedule (:47-48),
    `sys.exit(1)` on a non-finite loss (:59-62, seen one step late), the `Epoch: [e] [it/n]` log every 20 iterations, the
    cross-rank mean loss in the returned stats (MetricLogger.synchronize_between_processes, :95) and the
    `EpochSummary.finalize` block (util/misc.py:439-509) with its `throughput: ... samples/s` line: samples SUM-reduced
    over ranks / epoch wall clock MAX-reduced over ranks."""
    from .engine import HostPipeline, cosine_lr
    accum_iter = args.accum_iter
    n_iter = len(data_loader)
    pipe = getattr(engine, '_host_pipeline', None)
    if pipe is None:                                   # kept on the engine: the captured graph and buffers live across epochs
        pipe = engine._host_pipeline = HostPipeline(engine, graph=(accum_iter == 1) and not getattr(args, 'no_graph', False))
    loss_sum, loss_sq_sum, loss_cnt, samples_seen, lr = 0.0, 0.0, 0, 0, args.lr
    lrs = []
    torch.cuda.synchronize()
    t0 = time.time()

    def account(values, it_now):
        nonlocal loss_sum, loss_sq_sum, loss_cnt
        for v in values:
            if not math.isfinite(v):
                print("Loss is {}, stopping training".format(v))
                sys.exit(1)
            loss_sum += v
            loss_sq_sum += v * v
            loss_cnt += 1
            it_done = loss_cnt - 1
            if it_done % 20 == 0 or it_done == n_iter - 1:
                print(f"Epoch: [{epoch}]  [{it_done:4d}/{n_iter}]  lr: {lrs[it_done]:.6f}  loss: {v:.4f} ({loss_sum / loss_cnt:.4f})  "
                      f"time: {(time.time() - t0) / max(it_now, 1):.4f}")

    for it, (samples, _) in enumerate(data_loader):
        if it % accum_iter == 0:                       # per-iteration (not per-epoch) lr schedule
            lr = cosine_lr(it / n_iter + epoch, args.lr, args.min_lr, args.warmup_epochs, args.epochs)
        lrs.append(lr)
        if isinstance(samples, list):                  # --device_transform: uint8 frames of any size
            if getattr(engine, '_device_tf', None) is None:          # kept off `args`: args is pickled into checkpoints
                from .input_pipeline import TrainTransform
                engine._device_tf = TrainTransform(args.input_size, device=device)
            samples = engine._device_tf(samples)
        samples_seen += int(samples.shape[0])
        kw = {} if accum_iter == 1 else {"accum_iter": accum_iter, "micro_index": it % accum_iter}
        account(pipe.submit(samples, lr=lr, **kw), it + 1)
        if save_ckpt_fn is not None and last_time_save is not None and time_save_sec:
            if (time.time() - last_time_save[0]) >= time_save_sec:
                account(pipe.drain(), it + 1)
                save_ckpt_fn(epoch)
                last_time_save[0] = time.time()
    account(pipe.drain(), n_iter)
    torch.cuda.synchronize()
    epoch_time = time.time() - t0
    bad = engine.nonfinite_count()
    stats = torch.tensor([loss_sum, loss_sq_sum, loss_cnt, samples_seen, n_iter, bad], dtype=torch.float64, device=device)
    tmax = torch.tensor([epoch_time], dtype=torch.float64, device=device)
    if args.distributed:                               # util/misc.py:455-470: one SUM and one MAX per EPOCH
        dist.all_reduce(stats, op=dist.ReduceOp.SUM)
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    g_sum, g_sq, g_cnt, g_samples, g_batches, g_bad = stats.tolist()
    epoch_time = float(tmax.item())
    mean = g_sum / g_cnt if g_cnt else float('nan')
    std = math.sqrt(max(g_sq / g_cnt - mean * mean, 0.0)) if g_cnt > 1 else 0.0
    throughput = g_samples / epoch_time if epoch_time > 0 else 0.0
    if args.rank == 0:
        print(f"Train epoch {epoch} summary:")
        print(f"  loss: mean={mean:.6f}" + (f", std={std:.6f}" if g_cnt > 1 else "") if g_cnt else "  loss: mean=N/A")
        if lrs:
            lo, hi = min(lrs), max(lrs)
            print("  lr: " + (f"{lo:.6f}" if math.isclose(lo, hi, rel_tol=1e-12, abs_tol=1e-12) else f"min={lo:.6f}, max={hi:.6f}"))
        print(f"  throughput: {throughput:.2f} samples/s")
        print(f"  wall clock: {epoch_time:.2f} s")
        print(f"  data: batches={int(round(g_batches))}, samples={int(round(g_samples))}")
        if g_bad:
            print(f"  anomalies:\n    optimiser blocks with a non-finite gradient so far: {int(round(g_bad))}")
    return {"loss": mean, "lr": lr, "throughput": throughput, "epoch_time": epoch_time}


def main(args):
    from . import models_mae
    from .engine import PretrainEngine
    if args.precision == 'fp32':
        raise SystemExit("--precision fp32: the B200 path computes with bf16 tensor-core operands and fp32 accumulation only")
    _init_distributed(args)
    print("{}".format(args).replace(', ', ',\n'))
    device = torch.device('cuda', args.gpu)
    seed = args.seed + args.rank
    torch.manual_seed(seed)
    np.random.seed(seed)

    data_loader_train = _build_loader(args)
    model = models_mae.__dict__[args.model](norm_pix_loss=args.norm_pix_loss)
    model.to(device)
    eff_batch_size = args.batch_size * args.accum_iter * args.world_size
    if args.lr is None:
        args.lr = args.blr * eff_batch_size / 256
    print("base lr: %.2e" % (args.lr * 256 / eff_batch_size))
    print("actual lr: %.2e" % args.lr)
    print("accumulate grad iterations: %d" % args.accum_iter)
    print("effective batch size: %d" % eff_batch_size)
    # Ranks are seeded `seed + rank` before the model is built, as in the reference (:146-148), so their initial weights
    # differ; like DistributedDataParallel's constructor (:212-214) the engine broadcasts rank 0's parameter arena.
    engine = PretrainEngine(model, lr=args.lr, weight_decay=args.weight_decay, betas=(0.9, 0.95), mask_ratio=args.mask_ratio)

    if args.resume:                                    # util/misc.py load_model
        checkpoint = torch.load(args.resume, map_location='cpu', weights_only=False)
        model.load_state_dict(checkpoint['model'])
        model._store.mark_dirty()
        print("Resume checkpoint %s" % args.resume)
        if 'optimizer' in checkpoint and 'epoch' in checkpoint:
            engine.load_optimizer_state_dict(checkpoint['optimizer'])
            args.start_epoch = checkpoint['epoch'] + 1
            print("With optim & sched!")

    print(f"Start training for {args.epochs} epochs")
    last_time_save = [time.time()]
    time_save_sec = args.save_freq_mins * 60
    ckpt_dir = os.path.join(args.output_dir, "ckpts") if args.output_dir else None
    if ckpt_dir:
        os.makedirs(ckpt_dir, exist_ok=True)
    main_process = args.rank == 0

    def _is_milestone(ep):
        n = args.keep_every_n_epochs
        return n > 0 and ((ep + 1) % n == 0 or (ep + 1) == args.epochs)

    def _cleanup_checkpoints():
        paths = sorted(glob.glob(os.path.join(ckpt_dir, "checkpoint-*.pth")))
        survivors = set(paths[-args.keep_last:]) if args.keep_last > 0 else set()
        for p in paths:
            try:
                ep = int(os.path.basename(p).split('-')[1].split('.')[0])
            except Exception:
                ep = -1
            if p in survivors or _is_milestone(ep):
                continue
            try:
                os.remove(p)
            except OSError:
                pass

    def _save_checkpoint(ep):
        if not (ckpt_dir and main_process):
            return
        path = Path(ckpt_dir) / f"checkpoint-{ep}.pth"
        to_save = {"model": model.state_dict(), "optimizer": engine.optimizer_state_dict(), "epoch": ep, "scaler": {}, "args": args}
        torch.save(to_save, path)
        last = Path(ckpt_dir) / "last.pth"
        try:
            if last.is_symlink() or last.exists():
                last.unlink()
            last.symlink_to(path.name)
        except OSError:
            pass
        _cleanup_checkpoints()

    epoch = args.start_epoch

    def _term_handler(signum, frame):
        try:
            _save_checkpoint(epoch)
        finally:
            os._exit(0)

    for sig in (signal.SIGTERM, signal.SIGINT):
        signal.signal(sig, _term_handler)

    start_time = time.time()
    for epoch in range(args.start_epoch, args.epochs):
        if args.distributed and getattr(data_loader_train, "sampler", None) is not None:
            data_loader_train.sampler.set_epoch(args.seed + epoch)
        train_stats = train_one_epoch(engine, data_loader_train, device, epoch, args, save_ckpt_fn=_save_checkpoint,
                                      last_time_save=last_time_save, time_save_sec=time_save_sec)
        do_epoch_save = (args.save_freq_epochs > 0 and ((epoch + 1) % args.save_freq_epochs == 0)) or ((epoch + 1) == args.epochs)
        if args.output_dir and do_epoch_save:
            _save_checkpoint(epoch)
            last_time_save[0] = time.time()
        log_stats = {**{f'train_{k}': v for k, v in train_stats.items()}, 'epoch': epoch}
        if args.output_dir and main_process:
            with open(os.path.join(args.output_dir, "log.txt"), mode="a", encoding="utf-8") as f:
                f.write(json.dumps(log_stats) + "\n")
    total_time_str = str(datetime.timedelta(seconds=int(time.time() - start_time)))
    print('Training time {}'.format(total_time_str))
    if args.distributed:
        torch.cuda.synchronize()
        dist.barrier()


def cli(argv=None):
    """Console entry point (`ssl4polyp-pretrain`); the reference runs the file as a script (main_pretrain.py:325-330)."""
    args = get_args_parser().parse_args(argv)
    if args.output_dir:
        Path(args.output_dir).mkdir(parents=True, exist_ok=True)
    main(args)


if __name__ == '__main__':
    cli()
