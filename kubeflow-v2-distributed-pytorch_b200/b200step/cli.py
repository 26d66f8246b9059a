"""The reference's launcher and training driver (task.py:48-315), with the loop body replaced by TrainStep.

Everything a pipeline operator can observe is kept: the 20 flags and their defaults (task.py:55-94), the env
contract (WORLD_SIZE = nodes, RANK = node rank, MASTER_ADDR/PORT via env://, AIP_MODEL_DIR), `mp.spawn` with one
process per GPU and global rank = node_rank * ngpus + gpu (task.py:117-127,139-149), seeding (task.py:22-28,162),
the stdout lines, evaluation + `state_dict` save on rank 0 every 10th epoch *before* training that epoch
(task.py:279-298), fail-fast error behaviour (any exception -> non-zero exit, the only signal the pipeline reads).

Conscious differences: `google.cloud.storage` is imported lazily inside the upload branch
(B-11); the dataset is the synthetic tabular set instead of a CIFAR-10 download (B-2); `--arch` additionally
accepts the tabular models, and torchvision names are accepted by the parser but not built.
"""
from __future__ import annotations

import argparse
import dataclasses
import json
import os
import random
import sys
from datetime import datetime
from typing import List, Optional

import numpy as np
import torch

from .model import ARCHS, ModelSpec

TABULAR_ARCHS = sorted(ARCHS)


def set_random_seeds(random_seed: int = 0) -> None:
    """task.py:22-28."""
    torch.manual_seed(random_seed)
    torch.backends.cudnn.deterministic = True
    torch.backends.cudnn.benchmark = False
    np.random.seed(random_seed)
    random.seed(random_seed)


def _torchvision_model_names() -> List[str]:
    """task.py:50-52 builds the --arch choices from torchvision; keep them valid when torchvision is there."""
    try:
        import torchvision.models as models

        return sorted(name for name in models.__dict__
                      if name.islower() and not name.startswith("__") and callable(models.__dict__[name]))
    except Exception:
        return ["resnet18"]


def build_parser() -> argparse.ArgumentParser:
    model_names = _torchvision_model_names()
    parser = argparse.ArgumentParser(formatter_class=argparse.ArgumentDefaultsHelpFormatter)
    # ---- the reference's 20 flags, same order / dest / type / default (task.py:56-94) ----
    parser.add_argument("--local_rank", type=int, help="Local rank. Necessary for using the torch.distributed.launch utility.")
    parser.add_argument("--num_epochs", type=int, help="Number of training epochs.", default=100)
    parser.add_argument("--batch_size", type=int, help="Training batch size for one process.", default=1024)
    parser.add_argument("--learning_rate", dest="learning_rate", type=float, help="Learning rate.", default=0.1)
    parser.add_argument("--random_seed", type=int, help="Random seed.", default=0)
    parser.add_argument("--model_dir", type=str, help="Directory for saving models.",
                        default=os.environ["AIP_MODEL_DIR"] if "AIP_MODEL_DIR" in os.environ else "")
    parser.add_argument("--model_filename", type=str, help="Model filename.", default="resnet_distributed.pth")
    parser.add_argument("-a", "--arch", metavar="ARCH", default="resnet18", choices=model_names + TABULAR_ARCHS,
                        help="model architecture: " + " | ".join(TABULAR_ARCHS) +
                             " | resnet18 | resnet34 (B200 paths) | other torchvision names (accepted, not built) "
                             "(default: resnet18)")
    parser.add_argument("--momentum", default=0.9, type=float, metavar="M", help="momentum")
    parser.add_argument("--wd", "--weight-decay", default=1e-4, type=float, metavar="W",
                        help="weight decay (default: 1e-4)", dest="weight_decay")
    parser.add_argument("--pretrained", dest="pretrained", action="store_true", help="use pre-trained model")
    parser.add_argument("--local_training", dest="local_training", action="store_true",
                        help="use local machine for training")
    parser.add_argument("--rank", default=-1, type=int, help="node rank for distributed training")
    parser.add_argument("--resume", action="store_true", help="Resume training from saved checkpoint.")
    parser.add_argument("--world-size", default=int(os.getenv("WORLD_SIZE", -1)), type=int,
                        help="number of nodes for distributed training")
    parser.add_argument("--dist-url", default="http://localhost:8082", type=str,
                        help="url used to set up distributed training")
    parser.add_argument("--dist-backend", default="nccl", type=str, help="distributed backend")
    parser.add_argument("--multiprocessing-distributed", action="store_true",
                        help="Use multi-processing distributed training to launch N processes per node, which has "
                             "N GPUs. This is the fastest way to use PyTorch for either single node or multi node "
                             "data parallel training")
    parser.add_argument("--gpu", default=None, type=int, help="GPU id to use.")
    parser.add_argument("--workers", default=4, type=int, metavar="N", help="number of data loading workers (default: 4)")
    # ---- additive flags (defaults = reference behaviour) ----
    add = parser.add_argument_group("B200 step (additive)")
    add.add_argument("--optimizer", choices=["sgd", "adam"], default="sgd")
    add.add_argument("--precision", choices=["fp32", "tf32", "bf16"], default="fp32",
                     help="fp32 (default: the reference trains in fp32, task.py:309): fp32 storage; tf32: fp32 storage, "
                          "single-pass TF32 tensor-core products; bf16: bf16 storage + fp32 accumulate / master weights")
    add.add_argument("--dropout", type=float, default=0.0)
    add.add_argument("--synthetic_rows", type=int, default=10000,
                     help="rows of the synthetic tabular dataset (the reference's CIFAR-10 test split has 10000)")
    add.add_argument("--num_dense", type=int, default=None)
    add.add_argument("--num_cat", type=int, default=None)
    add.add_argument("--cardinality", type=int, default=None)
    add.add_argument("--embed_dim", type=int, default=None)
    add.add_argument("--hidden", type=str, default=None, help="comma-separated hidden widths, e.g. 512,512")
    add.add_argument("--num_classes", type=int, default=None)
    add.add_argument("--bench_steps", type=int, default=0,
                     help="after training, time this many steps with CUDA events and print one JSON line (rank 0)")
    add.add_argument("--no_graph", action="store_true", help="launch kernels eagerly instead of replaying a CUDA graph")
    add.add_argument("--plan_only", action="store_true",
                     help="rendezvous, build the sample partition and print the launch plan, then exit (needs no GPU)")
    add.add_argument("--save_final", action="store_true",
                     help="also save the trained state_dict to <model_dir>/<model_filename>.final at the end "
                          "(the reference only ever saves before training an epoch, SURVEY.md App. B-8)")
    add.add_argument("--quiet_steps", action="store_true", help="suppress the per-step 'training with gpu' line")
    add.add_argument("--set_epoch", action="store_true",
                     help="reshuffle each epoch like DistributedSampler.set_epoch(epoch) would; the reference never "
                          "calls it, so by default every epoch sees the same order (SURVEY.md App. B-4)")
    return parser


def spec_from_args(args) -> ModelSpec:
    if args.arch not in ARCHS:
        raise NotImplementedError(
            f"--arch {args.arch}: of the torchvision models only the BasicBlock ResNets (resnet18, resnet34) have a "
            f"B200 path (DESIGN.md §7); the tabular models are {TABULAR_ARCHS}")
    spec = ARCHS[args.arch]
    over = {}
    for name in ("num_dense", "num_cat", "cardinality", "embed_dim", "num_classes"):
        v = getattr(args, name)
        if v is not None:
            over[name] = v
    if args.hidden:
        over["hidden"] = tuple(int(t) for t in args.hidden.split(",") if t)
    if args.dropout:
        over["dropout"] = float(args.dropout)
    return dataclasses.replace(spec, **over) if over else spec


def main(argv: Optional[List[str]] = None) -> None:
    """task.py:48-130."""
    parser = build_parser()
    args = parser.parse_args(argv)

    if args.dist_url == "env://" and args.world_size == -1:
        args.world_size = int(os.environ["WORLD_SIZE"])

    args.distributed = args.world_size > 1 or args.multiprocessing_distributed

    ngpus_per_node = torch.cuda.device_count()
    if args.plan_only and ngpus_per_node == 0:
        ngpus_per_node = int(os.environ.get("B200STEP_PLAN_GPUS", "1"))

    # debugging (same lines as task.py:105-113)
    print(f"os WORLD_SIZE={os.getenv('WORLD_SIZE', -1)}")
    print(f"os RANK={os.getenv('RANK', 0)}")
    print(f"os MASTER_ADDR={os.getenv('MASTER_ADDR', 'localhost')}")
    print(f"os MASTER_PORT={os.getenv('MASTER_PORT', '8082')}")
    print(f"Arg - distributed={args.distributed}")
    print(f"Arg - multiprocessing_distributed={args.multiprocessing_distributed}")
    print(f"Arg - dist_backend={args.dist_backend}")
    print(f"Arg - dist_url={args.dist_url}")
    print(f"ngpus_per_node={ngpus_per_node}")

    start = datetime.now().strftime("%Y_%m_%d_%H_%M_%S")
    print(f"Starting training: {start}")
    if args.multiprocessing_distributed:
        args.world_size = ngpus_per_node * args.world_size
        print("GPU x WORLD SIZE = {}".format(args.world_size))
        import torch.multiprocessing as mp

        mp.spawn(main_worker, nprocs=ngpus_per_node, args=(ngpus_per_node, args))
    else:
        main_worker(args.gpu, ngpus_per_node, args)

    end = datetime.now().strftime("%Y_%m_%d_%H_%M_%S")
    print(f"Training complete: {end}")


def main_worker(gpu, ngpus_per_node, args) -> None:
    """task.py:132-315."""
    import torch.distributed as dist

    args.gpu = gpu
    if args.gpu is not None:
        print("Use GPU: {args.gpu} for training")  # sic: the reference prints the literal (task.py:137)

    if args.distributed:
        if args.dist_url == "env://" and args.rank == -1:
            args.rank = int(os.environ["RANK"])
            print(f"Distributed and getting rank from os.environ: rank={args.rank}")
        if args.multiprocessing_distributed:
            args.rank = args.rank * ngpus_per_node + gpu
            print(f"Distributed and Multiprocesing. Setting rank for each worker. rank={args.rank}")
        dist.init_process_group(backend=args.dist_backend, init_method=args.dist_url,
                                world_size=args.world_size, rank=args.rank)
        print("Process group initialized")
    world = args.world_size if args.distributed else 1
    rank = args.rank if args.distributed else 0

    num_epochs = args.num_epochs
    batch_size = args.batch_size          # per process: task.py:153 / SURVEY.md App. B-3
    model_filepath = os.path.join(args.model_dir, args.model_filename)

    set_random_seeds(random_seed=args.random_seed)
    print("random seeds")

    if args.pretrained:
        raise NotImplementedError("--pretrained needs the network and a torchvision model; not available here")
    print(f"=> creating model '{args.arch}'")
    from . import convnet
    from .data import make_synthetic
    from .model import init_parameters
    from .sampler import batch_bounds, rank_indices

    conv = args.arch in convnet.ARCH_BLOCKS                    # the reference's own model family (task.py:171)
    if conv:
        lay = convnet.resnet_layout(args.arch, num_classes=args.num_classes or 1000)
        params = convnet.init_state(args.arch, args.random_seed, lay.num_classes)
        dataset = convnet.load_cifar10_test_split("data") or convnet.make_synthetic_images(
            args.synthetic_rows, seed=args.random_seed)        # task.py:256-257 (train=False for both loaders)
        arch_name, n_params = args.arch, sum(v.numel() for k, v in params.items() if "running" not in k and
                                             "num_batches" not in k)
    else:
        spec = spec_from_args(args)
        params = init_parameters(spec, args.random_seed)
        dataset = make_synthetic(args.synthetic_rows, spec)
        arch_name, n_params = spec.name, spec.num_params
    my_rows = rank_indices(len(dataset), world, rank)           # DistributedSampler(train_set), task.py:260
    bounds = batch_bounds(my_rows.numel(), batch_size)          # DataLoader batching, task.py:263

    if args.plan_only:
        plan = {"rank": rank, "world": world, "gpu": gpu, "arch": arch_name, "params": n_params,
                "samples_this_rank": int(my_rows.numel()), "batches_per_epoch": len(bounds),
                "first_indices": my_rows[:4].tolist(), "last_batch_rows": len(bounds[-1]),
                "precision": args.precision, "optimizer": args.optimizer}
        print("PLAN " + json.dumps(plan))
        if args.distributed:
            dist.barrier()
            dist.destroy_process_group()
        return

    if not torch.cuda.is_available():
        raise RuntimeError("the B200 training step needs a CUDA device (like the reference, there is no CPU path)")
    if args.gpu is None:
        args.gpu = 0
    torch.cuda.set_device(args.gpu)
    if args.distributed:
        print(f"Distributed GPU device=cuda:{args.gpu}")
    else:
        print(f"Non-distributed GPU device id ={args.gpu}")

    from .engine import Communicator, StepConfig, TrainStep

    comm = None
    if world > 1:
        uid = [Communicator.make_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0, device=torch.device("cuda", args.gpu)
                                   if args.dist_backend == "nccl" else None)
        comm = Communicator(rank, world, uid[0])

    cfg = StepConfig(batch_size=batch_size, world=world, rank=rank, precision=args.precision,
                     optimizer=args.optimizer, lr=args.learning_rate, momentum=args.momentum,
                     weight_decay=args.weight_decay, dropout_seed=args.random_seed, use_graph=not args.no_graph)
    if conv:
        step = convnet.ConvTrainStep(lay, params, dataset, my_rows, cfg, comm, device=torch.device("cuda", args.gpu))
    else:
        step = TrainStep(spec, params, dataset, my_rows, cfg, comm, device=torch.device("cuda", args.gpu))

    for epoch in range(num_epochs):
        epoch_start = datetime.now().strftime("%Y_%m_%d_%H_%M_%S")
        print("Rank: {}, Epoch: {}, Training start: {}".format(rank, epoch, epoch_start))

        if epoch % 10 == 0:                                         # task.py:279
            if rank == 0:
                # task.py:281 (test set == train set, task.py:256-257); the image path's test loader applies the random
                # training transform too (task.py:257 passes `transform`)
                accuracy = step.evaluate(draws=convnet.draw_augmentation(len(dataset))) if conv else step.evaluate()
                state = step.state_dict()
                if args.local_training:
                    torch.save(state, model_filepath)               # task.py:283
                    print("saving model to local folders")
                else:
                    torch.save(state, args.model_filename)          # task.py:287
                    model_directory = os.environ["AIP_MODEL_DIR"]
                    print(f"AIP_MODEL_DIR={model_directory}")
                    storage_path = os.path.join(model_directory, "model", args.model_filename)
                    print(f"storage_path={storage_path}")
                    from google.cloud import storage               # lazy: SURVEY.md App. B-11

                    blob = storage.blob.Blob.from_string(storage_path, client=storage.Client())
                    blob.upload_from_filename(args.model_filename)
                print("-" * 75)
                epoch_middle = datetime.now().strftime("%Y_%m_%d_%H_%M_%S")
                print(f"Epoch: {epoch}, Accuracy: {accuracy}, Time: {epoch_middle}")
                print("-" * 75)

        if args.set_epoch and epoch > 0:
            step.set_row_order(rank_indices(len(dataset), world, rank, epoch=epoch))
        if conv:     # this epoch's RandomCrop / RandomHorizontalFlip draws, from torch's generator like the transforms
            step.set_draws(convnet.draw_augmentation(int(my_rows.numel())))
        for b in bounds:                                            # task.py:303
            if not args.quiet_steps:
                print(f"training with gpu {args.gpu}")              # task.py:306 (host-side only, no device sync)
            if len(b) == batch_size:
                step.step()
            else:
                step.step_tail()                                    # drop_last=False: the partial last batch
        # once per epoch, on every rank: surface device-side reports (bad ids / labels) instead
        # of training on for epochs with neutralised input
        step.synchronize()

    step.synchronize()
    epoch_end = datetime.now().strftime("%Y_%m_%d_%H_%M_%S")
    print(f"Epoch complete: {epoch_end}")

    if args.bench_steps > 0 and step.steps_per_epoch_full > 0:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        if args.distributed:
            dist.barrier()
        torch.cuda.synchronize()
        e0.record(step.main)
        for _ in range(args.bench_steps):
            step.step()
        e1.record(step.main)
        step.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1)], device="cuda")
        if args.distributed:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        if rank == 0:
            print(json.dumps({"metric": "train samples/sec", "unit": "samples/s", "n_gpus": world,
                              "value": world * batch_size * args.bench_steps / (ms.item() / 1e3),
                              "ms_per_step": ms.item() / args.bench_steps, "steps": args.bench_steps}))
    final_loss = step.loss_value()
    if rank == 0:
        print(f"final local loss (rank 0): {final_loss}")
        if args.save_final:
            torch.save(step.state_dict(), model_filepath + ".final")
    step.close()
    if comm is not None:
        comm.close()
    if args.distributed:
        dist.destroy_process_group()
