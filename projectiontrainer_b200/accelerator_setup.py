"""Drop-in for ref:Stage1/accelerator_setup.py.  HF `accelerate` is not a dependency of this build: the
fourteen Accelerator members `ProjectionTrainerStage1` touches are provided by
`Stage1Accelerator` directly over `torch.distributed` (NCCL on GPUs, gloo in the CPU tests), one process
per GPU, launched by torchrun / `python -m torch.distributed.run`."""
from __future__ import annotations

import logging
from collections import deque
import os

import torch
import torch.distributed as dist

logger = logging.getLogger(__name__)


class Stage1Accelerator:
    """Single-node data-parallel runtime: rank/device discovery from the torchrun environment, the sum
    all-reduce of the flat projector gradient (the one exchange step of the path), loss
    gather and main-process logging."""

    def __init__(self, gradient_accumulation_steps: int = 1, log_with=None, device=None, backend=None):
        self.gradient_accumulation_steps = int(gradient_accumulation_steps)
        self.sync_gradients = True        # ref never enters accumulate(): every micro-batch steps (F7)
        self.log_with = log_with
        self.trackers: list = []
        # the last records handed to log(): bounded (both trainers log every micro-batch; an unbounded list grows by
        # millions of dicts over a long run and is inherited by every forked DataLoader worker)
        self.logged = deque(maxlen=int(os.environ.get("PT_STAGE1_LOG_KEEP", "4096")))
        world = int(os.environ.get("WORLD_SIZE", "1"))
        self.process_index = int(os.environ.get("RANK", "0"))
        self.local_process_index = int(os.environ.get("LOCAL_RANK", "0"))
        if device is not None:
            self.device = torch.device(device)
        elif torch.cuda.is_available():
            torch.cuda.set_device(self.local_process_index)
            self.device = torch.device("cuda", self.local_process_index)
        else:
            self.device = torch.device("cpu")
        if world > 1 and not dist.is_initialized():
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            backend = backend or ("nccl" if self.device.type == "cuda" else "gloo")
            kw = {"device_id": self.device} if backend == "nccl" else {}
            dist.init_process_group(backend, **kw)
        self.num_processes = dist.get_world_size() if dist.is_initialized() else 1
        if dist.is_initialized():
            self.process_index = dist.get_rank()
        self.is_main_process = self.process_index == 0

    # ---- objects pass through: no DDP wrapper (the engine all-reduces the flat gradient itself)
    def prepare(self, *objs):
        return objs if len(objs) != 1 else objs[0]

    def prepare_model(self, model, device_placement=True):
        return model

    def unwrap_model(self, model):
        return getattr(model, "module", model)

    # ---- collectives
    def all_reduce_sum_(self, t: torch.Tensor) -> torch.Tensor:
        if self.num_processes > 1:
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return t

    def gather(self, t: torch.Tensor) -> torch.Tensor:
        t = t.detach().reshape(-1)
        if self.num_processes == 1:
            return t
        out = [torch.empty_like(t) for _ in range(self.num_processes)]
        dist.all_gather(out, t)
        return torch.cat(out)

    def wait_for_everyone(self):
        if self.num_processes > 1:
            dist.barrier()

    def gather_object(self, obj: list) -> list:
        """`accelerate.utils.gather_object` (ref:Stage2/trainer.py:13, 629): the per-process lists
        concatenated in rank order."""
        if self.num_processes == 1:
            return list(obj)
        out = [None] * self.num_processes
        dist.all_gather_object(out, list(obj))
        return [x for part in out for x in part]

    @property
    def backward_divisor(self) -> int:
        """`accelerate.Accelerator.backward(loss)` divides the loss by `gradient_accumulation_steps`
        before calling `.backward()` (every backend but DeepSpeed).  Both reference trainers have
        already divided it (Stage1/projector_trainer.py:236, Stage2/trainer.py:419), so the gradients
        the reference's optimizer sees are those of loss / GA**2; the trainers here ask for the same.
        (Behaviour of the pinned accelerate==1.6.0 stated from its source as known; the package is
        not in the build container — DESIGN.md lists it under "verify".)"""
        return self.gradient_accumulation_steps

    # ---- autograd-facing members (used when MLPProjector is driven through torch autograd)
    def backward(self, loss):
        loss.backward()

    def clip_grad_norm_(self, parameters, max_norm):
        return torch.nn.utils.clip_grad_norm_(list(parameters), max_norm)

    # ---- tracking
    def init_trackers(self, project_name=None, config=None, init_kwargs=None):
        if self.log_with == "wandb" and self.is_main_process:
            import wandb  # raises if absent: the caller logs and disables tracking, like the reference
            wandb.init(project=project_name, config=config, **(init_kwargs or {}).get("wandb", {}))
            self.trackers = ["wandb"]

    def log(self, values: dict, step=None):
        if not self.is_main_process:
            return
        self.logged.append((step, dict(values)))
        if "wandb" in self.trackers:
            import wandb
            wandb.log(values, step=step)

    def end_training(self):
        if dist.is_initialized():
            dist.destroy_process_group()


def setup_accelerator_and_logging(args):
    """Same contract as ref:Stage1/accelerator_setup.py:7-54."""
    accelerator = Stage1Accelerator(
        gradient_accumulation_steps=args.gradient_accumulation_steps,
        log_with="wandb" if not getattr(args, "disable_wandb", True) else None)
    log_level = logging.INFO if accelerator.is_main_process else logging.WARNING
    logging.basicConfig(level=log_level, format="%(asctime)s - %(name)s - %(levelname)s - %(message)s", force=True)
    if accelerator.is_main_process:
        logger.info(f"Using {accelerator.num_processes} processes.")
        logger.info("Effective batch size: "
                    f"{args.batch_size * accelerator.num_processes * args.gradient_accumulation_steps}")
    if accelerator.is_main_process and not getattr(args, "disable_wandb", True):
        kw = {"name": args.wandb_run_name} if getattr(args, "wandb_run_name", None) else {}
        try:
            accelerator.init_trackers(project_name=args.wandb_project, config=vars(args), init_kwargs={"wandb": kw})
        except Exception as e:  # same behaviour as the reference: log and carry on without tracking
            logger.error(f"Failed to initialize WandB tracker: {e}. Disabling WandB.")
            accelerator.log_with = None
    logger.info(f"Process {accelerator.process_index} using device: {accelerator.device}")
    return accelerator
